"""A few launches of the round's new / changed streaming kernels on a [16384, 14336] bf16 tensor, for
`ncu --set full --clock-control none --import-source on` (one GPU, never a bench value)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from modelopt_b200 import ops  # noqa: E402

torch.manual_seed(0)
x = (torch.randn(16384, 14336, device="cuda") * 2).to(torch.bfloat16)
y = torch.empty_like(x)
am = ops.reduce_amax(x).float()
amb = ops.reduce_amax(x.view(-1, 128), axis=1).float()
bam = ops.reduce_amax(x.view(-1, 16), axis=1).float().reshape(-1)
for _ in range(2):
    ops.qdq_mx(x, 32, (4, 3), out=y)                      # mx_e8m0_kernel<bf16, 32, E4M3>
    ops.qdq_mx(x, 32, (2, 1), out=y)                      # mx_e8m0_kernel<bf16, 32, E2M1>
    ops.qdq_nvfp4_dynamic(x, am, 16, out=y)               # nvfp4_dynamic_kernel (div.full hoisted)
    ops.qdq_nvfp4_static(x, bam, am)                      # nvfp4_static_kernel<.., kFromAmax>
    ops.qdq_int(x.view(-1, 128), amb, 4, False, False)    # qdq_stream_kernel<IntOp> per-chunk, amax prefetched
    ops.mxfp8_pack(x)                                     # mxfp8_pack_kernel
    ops.int4_pack(x.reshape(-1), 128)                     # pack4_vec_kernel<INT4>
torch.cuda.synchronize()
print("ok")
