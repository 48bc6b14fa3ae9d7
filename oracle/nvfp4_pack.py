"""Oracle: NVFP4 real-quant packer / unpacker (CPU torch).  TEST INFRASTRUCTURE ONLY.

Restates ``NVFP4QTensor`` — modelopt/torch/quantization/qtensor/nvfp4_tensor.py:
``_cast_per_block_scale_to_fp8`` :32-46, ``get_weights_scaling_factor`` :169-201,
``get_weights_scaling_factor_2`` :204-207, ``_cast_fp4`` :230-250, ``quantize`` :253-342,
``dequantize`` :344-407 and the Triton ``fp4_dequantize`` (gemm/fp4_kernel.py:54-190).

Layout produced (bit-exact requirement): ``codes`` uint8 ``[..., N/2]`` with element ``2j`` in the
LOW nibble and ``2j+1`` in the HIGH nibble, nibble = ``sign<<3 | e2m1_index``; ``sf`` float8_e4m3fn
``[..., N/16]``; ``sf2`` fp32 scalar ``amax/(6*448)``.
"""

from __future__ import annotations

import torch
import torch.nn.functional as F

E2M1_BOUNDS = torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5.0])
E2M1_VALUES = torch.tensor([0, 0.5, 1, 1.5, 2, 3, 4, 6, 0, -0.5, -1, -1.5, -2, -3, -4, -6])


def weights_scaling_factor_2(x: torch.Tensor) -> torch.Tensor:
    """nvfp4_tensor.py:204-207."""
    amax = torch.maximum(x.max().abs(), x.min().abs())
    return amax.float() / (6.0 * 448.0)


def cast_per_block_scale_to_fp8(s: torch.Tensor) -> torch.Tensor:
    """nvfp4_tensor.py:32-46 (no ``per_block_scale_max`` rescale on the dynamic path)."""
    return s.clamp(min=2**-9, max=448.0).to(torch.float8_e4m3fn)


def weights_scaling_factor(x: torch.Tensor, block: int, sf2: torch.Tensor) -> torch.Tensor:
    """nvfp4_tensor.py:169-201: block amax / (6*sf2); 0 -> 1.0; clamp [2^-9, 448]; cast e4m3."""
    assert x.shape[-1] % block == 0
    xb = x.reshape(*x.shape[:-1], -1, block)
    bam = torch.maximum(xb.amax(-1).abs(), xb.amin(-1).abs()).float()
    s = bam / (6.0 * sf2)
    s[s == 0] = 1.0
    return cast_per_block_scale_to_fp8(s)


def cast_fp4(w: torch.Tensor) -> torch.Tensor:
    """nvfp4_tensor.py:230-250: sign<<3 | searchsorted(bounds, |w|) (+1 on the odd bounds)."""
    sign = (w < 0).to(torch.uint8)
    wa = w.abs()
    o = torch.searchsorted(E2M1_BOUNDS, wa, out_int32=True).to(torch.uint8)
    odd = torch.any(wa.unsqueeze(-1) == E2M1_BOUNDS[[1, 3, 5]], dim=-1).to(torch.uint8)
    return (sign << 3) + o + odd


def quantize(x: torch.Tensor, block: int = 16, sf: torch.Tensor | None = None,
             sf2: torch.Tensor | None = None):
    """nvfp4_tensor.py:253-342 (torch path).  Returns (codes uint8, sf e4m3, sf2 fp32)."""
    pad = (-x.shape[-1]) % block
    if pad:
        x = F.pad(x, (0, pad))
    if sf2 is None:
        sf2 = weights_scaling_factor_2(x)
    if sf is None:
        sf = weights_scaling_factor(x, block, sf2)
    xb = x.reshape(*x.shape[:-1], -1, block)
    scaled = xb / (sf.to(torch.float32) * sf2).unsqueeze(-1)  # promotes to fp32
    scaled = scaled.reshape(x.shape)
    q = cast_fp4(scaled)
    packed = (q[..., 1::2] << 4) | q[..., 0::2]
    return packed, sf, sf2


def dequantize(codes: torch.Tensor, sf: torch.Tensor, sf2: torch.Tensor, block: int, dtype,
               shape=None, fast: bool = False) -> torch.Tensor:
    """nvfp4_tensor.py:344-407.  ``fast=False``: value*(sf*sf2); ``fast=True``: (value*sf)*sf2
    (Triton fp4_dequantize, gemm/fp4_kernel.py:131-133)."""
    lo = (codes & 0x0F).long()
    hi = (codes >> 4).long()
    idx = torch.stack([lo, hi], dim=-1).reshape(*codes.shape[:-1], -1)
    v = E2M1_VALUES[idx]
    vb = v.reshape(*v.shape[:-1], -1, block)
    if fast:
        out = vb * sf.to(torch.float32).unsqueeze(-1) * sf2
    else:
        out = vb * (sf.to(torch.float32) * sf2).unsqueeze(-1)
    out = out.reshape(v.shape)
    if shape is not None:
        out = out.reshape(*out.shape[:-1], -1)[..., : shape[-1]].reshape(shape)
    return out.to(dtype)
