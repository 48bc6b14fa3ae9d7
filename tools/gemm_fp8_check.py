"""FP8 per-tensor linear: TFLOP/s of mo_linear_fp8 (tcgen05 kind::f8f6f4) beside torch._scaled_mm on the same operands.
Developer tool: `gpurun -- python tools/gemm_fp8_check.py > gpurun_out/gemm_fp8_check.txt`."""
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


for M, N, K in ((16384, 4096, 4096), (16384, 14336, 4096), (16384, 4096, 14336), (8192, 8192, 8192)):
    torch.manual_seed(0)
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
    sa = (x.float().abs().amax() / 448).reshape(1)
    sb = (w.float().abs().amax() / 448).reshape(1)
    a, b = ops.q8_pack(x, sa, "fp8_sat"), ops.q8_pack(w, sb, "fp8_sat")
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    flops = 2.0 * M * N * K
    t = timeit(lambda: ops.linear_fp8(a, b, sa, sb, out=out))
    tr = timeit(lambda: torch._scaled_mm(a, b.t(), scale_a=sa, scale_b=sb, out_dtype=torch.bfloat16, use_fast_accum=True))
    tp = timeit(lambda: ops.q8_pack(x, sa, "fp8_sat"))
    ref = torch._scaled_mm(a, b.t(), scale_a=sa, scale_b=sb, out_dtype=torch.bfloat16, use_fast_accum=True)
    same = (out == ref).float().mean().item()
    print(f"{M}x{N}x{K}: mo_linear_fp8 {t:.3f} ms = {flops / t / 1e9:7.1f} TFLOP/s | torch._scaled_mm {tr:.3f} ms = "
          f"{flops / tr / 1e9:7.1f} TFLOP/s | activation pack {tp:.3f} ms | identical outputs {100 * same:.2f} %")
