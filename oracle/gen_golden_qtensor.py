"""Golden vectors for the real-quant packers, from the REAL reference's qtensor classes on CPU tensors.
TEST INFRASTRUCTURE ONLY.  Build container only (needs /root/reference):

    PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden_qtensor.py      # -> tests/golden/qtensor_cpu.npz
"""

from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle.gen_golden import DTYPES, OUT, to_np  # noqa: E402  (also puts the reference + shims on sys.path)


def base_tensor():
    torch.manual_seed(77)
    x = torch.randn(24, 256) * torch.rand(24, 1) * 4
    x[1] = 0.0  # all-zero blocks
    x[2, :32] = 0.0
    x[3, 5] = 300.0  # outlier
    x[4] *= 1e-5
    x[5, :8] = torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5.0, 6.0])  # E2M1 ties (scale 1 when amax = 6)
    x[5, 8:] = x[5, 8:].clamp(-5.9, 5.9)
    x[6, :16] = torch.arange(-8, 8).float() * 0.5 + 0.25  # .5 ties for the integer formats
    return x


def main():
    from modelopt.torch.quantization.qtensor import (FP8QTensor, INT4QTensor, INT8QTensor, MXFP4QTensor, MXFP8QTensor,
                                                     NF4QTensor)

    g = {}
    base = base_tensor()
    for dn, dt in DTYPES.items():
        x = base.to(dt)
        g[f"x_{dn}"] = to_np(x)
        nz = x.clone()
        nz[1] = x[0]
        nz[2, :32] = x[0, :32]  # INT4 / NF4 divide by the block amax: keep the zero blocks out (NaN -> uint8 is UB)
        g[f"xnz_{dn}"] = to_np(nz)
        for block in (128, 32):
            qt, s = INT4QTensor.quantize(nz, block)
            g[f"int4_q_{dn}_{block}"], g[f"int4_s_{dn}_{block}"] = to_np(qt._quantized_data), to_np(s)
            g[f"int4_dq_{dn}_{block}"] = to_np(qt.dequantize(scale=s, block_sizes={-1: block}))
            sbs = nz.numel() // block  # ONE scale block: the only layout the reference's dequantize broadcast supports
            qt, s = NF4QTensor.quantize(nz, block, sbs)
            g[f"nf4_q_{dn}_{block}"], g[f"nf4_s_{dn}_{block}"] = to_np(qt._quantized_data), to_np(s)
            if dt != torch.float16:  # reference: bf16 table * f16 -> f32 + f16 mix raises nothing but is untested; skip
                sq, ds, zp = NF4QTensor.double_quantization(s.float(), sbs, 8)
                g[f"nf4_sq_{dn}_{block}"], g[f"nf4_ds_{dn}_{block}"], g[f"nf4_zp_{dn}_{block}"] = to_np(sq), to_np(ds), to_np(zp)
                g[f"nf4_dq_{dn}_{block}"] = to_np(qt.dequantize(scale=sq, double_scale=ds, scale_zeros=zp,
                                                                block_sizes={-1: block}))
        for name, cls in (("int8", INT8QTensor), ("fp8", FP8QTensor)):
            for tag, kw in (("pt", {}), ("ax0", {"axis": 0}), ("ax1", {"axis": -1}), ("b128", {"block_sizes": {-1: 128}}),
                            ("b2d", {"block_sizes": {-1: 64, -2: 8}})):
                qt, s = cls.quantize(nz, **kw)
                g[f"{name}_q_{dn}_{tag}"], g[f"{name}_s_{dn}_{tag}"] = to_np(qt._quantized_data.view(torch.uint8)), to_np(s)
                dkw = {"block_sizes": kw["block_sizes"]} if "block_sizes" in kw else {}
                g[f"{name}_dq_{dn}_{tag}"] = to_np(qt.dequantize(scale=s, **dkw))
        # ragged FP8 block (padding path, fp8_tensor.py:61-63)
        rag = nz[:, :200].contiguous()
        qt, s = FP8QTensor.quantize(rag, block_sizes={-1: 128})
        g[f"fp8_q_{dn}_rag"], g[f"fp8_s_{dn}_rag"] = to_np(qt._quantized_data.view(torch.uint8)), to_np(s)
        g[f"fp8_dq_{dn}_rag"] = to_np(qt.dequantize(scale=s, block_sizes={-1: 128}))
        # MX packers (zero blocks are well defined there)
        qt, s = MXFP4QTensor.quantize(x, 32)
        g[f"mxfp4_q_{dn}"], g[f"mxfp4_s_{dn}"] = to_np(qt._quantized_data), to_np(s)
        g[f"mxfp4_dq_{dn}"] = to_np(qt.dequantize(scale=s, block_sizes={-1: 32}))
        qt, s = MXFP8QTensor.quantize(x)
        g[f"mxfp8_q_{dn}"], g[f"mxfp8_s_{dn}"] = to_np(qt._quantized_data.view(torch.uint8)), to_np(s)
        g[f"mxfp8_dq_{dn}"] = to_np(qt.dequantize(scale=s))
        rag = x[:, :200].contiguous()
        qt, s = MXFP8QTensor.quantize(rag)
        g[f"mxfp8_q_{dn}_rag"], g[f"mxfp8_s_{dn}_rag"] = to_np(qt._quantized_data.view(torch.uint8)), to_np(s)
        g[f"mxfp8_dq_{dn}_rag"] = to_np(qt.dequantize(scale=s))
        g[f"mxfp8_sf_{dn}"] = to_np(MXFP8QTensor.get_weights_scaling_factor(x))
    path = os.path.join(OUT, "qtensor_cpu.npz")
    np.savez_compressed(path, **g)
    print("wrote", path, len(g), "arrays", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
