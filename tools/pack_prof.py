"""NVFP4 pack on a [16384, 14336] bf16 activation: a few launches for ncu, or `--time` to time it on several
independently allocated buffer sets (the pack has three streams - x, codes, scales - and its time depends on where the
allocator places them)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402

torch.manual_seed(0)
x = torch.randn(16384, 14336, device="cuda", dtype=torch.bfloat16)
AMAX = [float(a.split("=")[1]) for a in sys.argv if a.startswith("--amax=")]
if AMAX:  # force the tensor's amax (its bf16 mantissa decides how many block scales are exact E4M3 ties)
    x.clamp_(-AMAX[0], AMAX[0])
    x[0, 0] = AMAX[0]
sf2 = (x.float().abs().max() / (6.0 * 448.0)).reshape(1)
if "--time" in sys.argv:
    keep = []
    for rep in range(2):
        xs = torch.empty_like(x).copy_(x)
        keep.append(torch.empty(rep * 7919 * 1024 + 1024, device="cuda", dtype=torch.uint8))  # shift later allocations
        for _ in range(3):
            ops.nvfp4_pack(xs, 16, sf2)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(15)]
        for a, b in ev:
            a.record(); ops.nvfp4_pack(xs, 16, sf2); b.record()
        torch.cuda.synchronize()
        ts = sorted(a.elapsed_time(b) for a, b in ev)
        nbytes = x.numel() * 2.5625
        print(f"buffer set {rep}: median {ts[7]:.3f} ms ({nbytes / ts[7] / 1e6:.0f} GB/s)  best {ts[0]:.3f} ms")
        keep.append(xs)
else:
    for _ in range(5):
        ops.nvfp4_pack(x, 16, sf2)
    torch.cuda.synchronize()
    print("ok")
