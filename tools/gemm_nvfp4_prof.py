"""Three launches of the block-scaled NVFP4 linear on 16384 x 4096 x 4096 for ncu."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402

x = torch.randn(16384, 4096, device="cuda", dtype=torch.bfloat16)
w = torch.randn(4096, 4096, device="cuda", dtype=torch.bfloat16) * 0.05
a = ops.nvfp4_quantize_operand(x, x.float().abs().max(), 128)
b = ops.nvfp4_quantize_operand(w, w.float().abs().max(), 256)
y = torch.empty(16384, 4096, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    ops.linear_nvfp4(*a, *b, None, out=y)
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
