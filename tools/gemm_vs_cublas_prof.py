"""One launch each of mo_linear_pair and cuBLAS (F.linear) on the same bf16 operands, for an ncu side-by-side."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402

M, N, K = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "16384x4096x4096").split("x"))
x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) * 0.05
y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    ops.linear(x, w, None, out=y, pair=True)
    torch.nn.functional.linear(x, w)
torch.cuda.synchronize()
print("ok")
