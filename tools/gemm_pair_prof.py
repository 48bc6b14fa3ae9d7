"""One launch each of the cta_group::2 linears on the Llama-3-8B q_proj shape (16384 x 4096 x 4096) for ncu."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402

x = torch.randn(16384, 4096, device="cuda", dtype=torch.bfloat16)
w = torch.randn(4096, 4096, device="cuda", dtype=torch.bfloat16) * 0.05
y = torch.empty(16384, 4096, device="cuda", dtype=torch.bfloat16)
sa = (x.float().abs().amax() / 448).reshape(1)
sb = (w.float().abs().amax() / 448).reshape(1)
a, b = ops.q8_pack(x, sa, "fp8_sat"), ops.q8_pack(w, sb, "fp8_sat")
for _ in range(2):
    ops.linear(x, w, None, out=y, pair=True)
    ops.linear_fp8(a, b, sa, sb, out=y, pair=True)
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
