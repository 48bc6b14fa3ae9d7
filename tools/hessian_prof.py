"""One launch each of the local-Hessian kernels and the 4-bit packers on Llama-3-8B shapes, for ncu
(block_hessian_kernel, hessian_reduce_kernel, hessian_sweep_kernel, pack4_vec_kernel)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402

torch.manual_seed(0)
x = torch.randn(16384, 4096, device="cuda", dtype=torch.bfloat16)
w = torch.randn(8192, 4096, device="cuda", dtype=torch.bfloat16) * 0.05
H = torch.zeros(256, 16, 16, device="cuda")
ga = w.float().abs().amax().reshape(1)
for _ in range(2):
    ops.block_hessian_into(x, H, 16)
    am = ops.nvfp4_fp8_scale_sweep_hessian(w.view(-1, 16), ga, H / 32768, 16)
    q, s = ops.int4_pack(w.reshape(-1), 128)
    q2, s2 = ops.nf4_pack(w.reshape(-1), 64)
torch.cuda.synchronize()
print("ok", float(am.mean()), int(q.sum()), int(q2.sum()))
