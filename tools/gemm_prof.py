"""Three launches of the tcgen05 linear on the Llama-3-8B q_proj shape (16384 x 4096 x 4096 bf16) for ncu."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402

x = torch.randn(16384, 4096, device="cuda", dtype=torch.bfloat16)
w = torch.randn(4096, 4096, device="cuda", dtype=torch.bfloat16) * 0.05
y = torch.empty(16384, 4096, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    ops.linear(x, w, None, out=y)
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
