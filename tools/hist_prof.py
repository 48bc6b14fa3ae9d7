"""mo_hist on a [16384, 14336] bf16 activation (2048 bins, fused amax) for ncu: 4 launches."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402

torch.manual_seed(0)
x = torch.randn(16384, 14336, device="cuda", dtype=torch.bfloat16)
rm = x.float().abs().max().reshape(1)
cnt = torch.zeros(2048, dtype=torch.int64, device="cuda")
bits = torch.zeros(1, dtype=torch.int32, device="cuda")
for _ in range(4):
    ops.hist_into(x, rm, cnt, bits)
torch.cuda.synchronize()
print("ok", int(cnt.sum()))
