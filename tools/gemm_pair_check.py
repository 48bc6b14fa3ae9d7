"""cta_group::2 bf16 / f16 linear (mo_linear_pair) against the single-CTA kernel and cuBLAS: equality and TFLOP/s."""
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


shapes = [(256, 256, 64), (256, 256, 512), (300, 264, 144), (1, 8, 8), (512, 512, 1024), (2048, 4096, 4096),
          (16384, 4096, 4096), (16384, 14336, 4096), (16384, 4096, 14336), (8192, 8192, 8192)]
for dt in (torch.bfloat16, torch.float16):
    for M, N, K in shapes:
        if dt == torch.float16 and M * N * K > 1e10:
            continue
        torch.manual_seed(0)
        x = torch.randn(M, K, device="cuda", dtype=dt)
        w = (torch.randn(N, K, device="cuda") * 0.05).to(dt)
        bias = (torch.randn(N, device="cuda") * 0.1).to(dt)
        one = ops.linear(x, w, bias, pair=False)
        two = ops.linear(x, w, bias, pair=True)
        ref = torch.randn(M, N, device="cuda", dtype=dt)
        l1 = torch.zeros(1, dtype=torch.float64, device="cuda")
        l2 = torch.zeros(1, dtype=torch.float64, device="cuda")
        ops.linear_loss_into(x, w, bias, ref, l1, pair=False)
        ops.linear_loss_into(x, w, bias, ref, l2, pair=True)
        torch.cuda.synchronize()
        same = (one == two).float().mean().item()
        line = (f"{str(dt)[6:]} {M}x{N}x{K}: identical {100 * same:.3f} %  loss {l1.item():.9g} vs {l2.item():.9g}")
        if M * N * K > 1e10:
            flops = 2.0 * M * N * K
            o1, o2 = torch.empty_like(one), torch.empty_like(one)
            t1 = timeit(lambda: ops.linear(x, w, out=o1, pair=False))
            t2 = timeit(lambda: ops.linear(x, w, out=o2, pair=True))
            tr = timeit(lambda: torch.nn.functional.linear(x, w))
            line += f" | 1-CTA {flops / t1 / 1e9:7.1f}  pair {flops / t2 / 1e9:7.1f}  cuBLAS {flops / tr / 1e9:7.1f} TFLOP/s"
        print(line, flush=True)
