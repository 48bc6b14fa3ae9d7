"""cta_group::2 FP8 linear (mo_linear_fp8_pair) against the single-CTA kernel: equality and TFLOP/s."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402


def timeit(fn, n=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


shapes = [(256, 256, 128), (256, 256, 512), (300, 264, 144), (1, 8, 16), (512, 512, 1024), (2048, 4096, 4096),
          (16384, 4096, 4096), (16384, 14336, 4096), (16384, 4096, 14336), (8192, 8192, 8192)]
for M, N, K in shapes:
    torch.manual_seed(0)
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
    sa = (x.float().abs().amax() / 448).reshape(1)
    sb = (w.float().abs().amax() / 448).reshape(1)
    a, b = ops.q8_pack(x, sa, "fp8_sat"), ops.q8_pack(w, sb, "fp8_sat")
    bias = (torch.randn(N, device="cuda") * 0.1).to(torch.bfloat16)
    one = ops.linear_fp8(a, b, sa, sb, bias, pair=False)
    two = ops.linear_fp8(a, b, sa, sb, bias, pair=True)
    torch.cuda.synchronize()
    same = (one == two).float().mean().item()
    line = f"{M}x{N}x{K}: identical {100 * same:.3f} %  max diff {(one.float() - two.float()).abs().max().item():.3e}"
    if M * N * K > 1e10:
        flops = 2.0 * M * N * K
        o1, o2 = torch.empty_like(one), torch.empty_like(one)
        t1 = timeit(lambda: ops.linear_fp8(a, b, sa, sb, out=o1, pair=False))
        t2 = timeit(lambda: ops.linear_fp8(a, b, sa, sb, out=o2, pair=True))
        tr = timeit(lambda: torch._scaled_mm(a, b.t(), scale_a=sa, scale_b=sb, out_dtype=torch.bfloat16, use_fast_accum=True))
        line += (f" | 1-CTA {flops / t1 / 1e9:7.1f}  pair {flops / t2 / 1e9:7.1f}  torch._scaled_mm {flops / tr / 1e9:7.1f} TFLOP/s")
    print(line, flush=True)
