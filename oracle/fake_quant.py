"""Oracle: fake quantize-dequantize functionals (CPU, fp32 torch).  TEST INFRASTRUCTURE ONLY.

Restates, op for op and in the same operation order (``x/s`` is not ``x*(1/s)``):

* INT-N   : ``fake_tensor_quant_device``            modelopt/torch/kernels/quantization/gemm/tensor_quant_gpu.cu:43-56
            (CPU twin ``_tensor_quant``              modelopt/torch/quantization/tensor_quant.py:607-645)
* FP8     : ``fake_e4m3fy`` / ``_fp8_eager``         .../gemm/tensor_quant_gpu_fp8.cu:35-107, tensor_quant.py:46-59
* NVFP4 dynamic : Triton ``fp4_fake_quant_kernel``   .../gemm/fp4_kernel_hopper.py:32-99,
            ``fp8_quantize_scale`` / ``fp4_round_magnitude``  .../common/nvfp4_quant.py:33-63, :104-125
* NVFP4 static  : ``static_blockwise_fp4_fake_quant`` + ``compute_fp4_scales`` + ``nvfp4_scalar_quant``
            .../gemm/fp4_kernel.py:217-316, .../common/nvfp4_quant.py:66-101
* FP4 cast: ``static_blockwise_fp4_cast_kernel``     .../gemm/fp4_kernel.py:319-363

All divisions here are IEEE (``div.rn.f32``).  The reference's Triton NVFP4 kernels divide with
``div.full.f32`` (approximate) on the GPU; the product kernel has an ``ieee`` numerics mode that
must match this file bit for bit, and a ``triton`` mode that is pinned against captured outputs of
the compiled Triton kernel (tests/golden/nvfp4_triton_b200.npz).
"""

from __future__ import annotations

import torch
import torch.nn.functional as F

EPSILON = 1.0 / (1 << 24)  # tensor_quant_gpu.cu:25, tensor_quant.py:631


def _f32(v, like: torch.Tensor) -> torch.Tensor:
    if isinstance(v, torch.Tensor):
        return v.to(torch.float32)
    return torch.tensor(v, dtype=torch.float32, device=like.device)


# ----------------------------------------------------------------------------------------------
# INT-N
# ----------------------------------------------------------------------------------------------
def int_bounds(num_bits: int, unsigned: bool, narrow_range: bool, cuda_semantics: bool = True):
    """tensor_quant_gpu.cu:38-41, :67-69 (CUDA) / tensor_quant.py:622-628 (CPU)."""
    bound = float((1 << (num_bits - 1 + int(unsigned))) - 1)
    if cuda_semantics:
        min_bound = -(bound + (0.0 if narrow_range else 1.0))
    elif unsigned:
        min_bound = 0.0
    else:
        min_bound = -bound if narrow_range else -bound - 1.0
    return min_bound, bound


def int_qdq(
    x: torch.Tensor,
    amax,
    num_bits: int = 8,
    unsigned: bool = False,
    narrow_range: bool = True,
    cuda_semantics: bool = True,
) -> torch.Tensor:
    """INT-N quantize-dequantize; ``amax`` broadcastable to ``x``.

    ``cuda_semantics=True`` follows the CUDA kernel (``amax < 2^-24`` -> 0, strict,
    tensor_quant_gpu.cu:46); ``False`` follows the CPU ``_tensor_quant`` (``<=``, tensor_quant.py:632).
    They differ only at ``amax == 2^-24``.
    """
    dt = x.dtype
    xf = x.to(torch.float32)
    am = _f32(amax, x)
    min_bound, max_bound = int_bounds(num_bits, unsigned, narrow_range, cuda_semantics)
    zero = (am < EPSILON) if cuda_semantics else (am <= EPSILON)
    safe = torch.where(zero, torch.ones_like(am), am)
    scale = torch.tensor(max_bound, dtype=torch.float32) / safe  # div.rn.f32
    q = torch.clamp(torch.round(xf * scale), min_bound, max_bound)  # rint == half-to-even
    out = q / scale  # divide, not multiply (tensor_quant_gpu.cu:55)
    out = torch.where(zero.expand_as(out), torch.zeros_like(out), out)
    return out.to(dt)


# ----------------------------------------------------------------------------------------------
# FP8 E4M3
# ----------------------------------------------------------------------------------------------
def fp8_scales(amax: torch.Tensor):
    """tensor_quant_gpu_fp8.cu:93-98 / tensor_quant.py:49-54: scale = 448/safe(amax), inv = 1/scale."""
    am = amax.to(torch.float32)
    safe = torch.where(am <= EPSILON, torch.ones_like(am), am)
    scale = 448.0 / safe
    inv = 1.0 / scale
    return scale, inv


def fp8_qdq(x: torch.Tensor, amax=None) -> torch.Tensor:
    """FP8 E4M3 quantize-dequantize; ``amax`` None (plain cast), scalar, or broadcastable."""
    dt = x.dtype
    xf = x.to(torch.float32)
    if amax is None:
        return xf.to(torch.float8_e4m3fn).to(dt)  # tensor_quant.py:56 without scaling
    scale, inv = fp8_scales(_f32(amax, x))
    y = (xf * scale).clamp(min=-448.0, max=448.0)  # == cvt.rn.satfinite
    y = y.to(torch.float8_e4m3fn).to(torch.float32) * inv
    return y.to(dt)


# ----------------------------------------------------------------------------------------------
# E2M1 rounding
# ----------------------------------------------------------------------------------------------
def e2m1_round_magnitude(a: torch.Tensor) -> torch.Tensor:
    """common/nvfp4_quant.py:33-63: round |x|/scale to {0,.5,1,1.5,2,3,4,6}; NaN -> 6."""
    f = lambda v: torch.full_like(a, v)
    return torch.where(
        a <= 0.25, f(0.0),
        torch.where(a < 0.75, f(0.5),
        torch.where(a <= 1.25, f(1.0),
        torch.where(a < 1.75, f(1.5),
        torch.where(a <= 2.5, f(2.0),
        torch.where(a < 3.5, f(3.0),
        torch.where(a <= 5.0, f(4.0), f(6.0))))))))  # fmt: skip


def fp4_cast(x: torch.Tensor, out_dtype=None) -> torch.Tensor:
    """gemm/fp4_kernel.py:319-363: round pre-scaled values to the signed E2M1 grid."""
    out_dtype = out_dtype or x.dtype
    xf = x.to(torch.float32)
    q = e2m1_round_magnitude(xf.abs())
    return torch.where(xf >= 0, q, -q).to(out_dtype)


# ----------------------------------------------------------------------------------------------
# NVFP4 dynamic (two-level scale computed from the data)
# ----------------------------------------------------------------------------------------------
def _max_drop_nan(xa: torch.Tensor) -> torch.Tensor:
    """``max.f32`` reduction: NaN operands are dropped unless every operand is NaN."""
    nan = torch.isnan(xa)
    m = torch.where(nan, torch.full_like(xa, -float("inf")), xa).amax(-1, keepdim=True)
    return torch.where(nan.all(-1, keepdim=True), torch.full_like(m, float("nan")), m)


def nvfp4_global_scale(global_amax: torch.Tensor) -> torch.Tensor:
    """fp4_kernel_hopper.py:140 (host) + :73-74 (kernel): amax/(6*448), <=0 -> 1e-12."""
    gs = global_amax.to(torch.float32) / (6.0 * 448.0)
    return torch.where(gs > 0.0, gs, torch.tensor(1e-12, dtype=torch.float32))


def nvfp4_block_scale(block_max: torch.Tensor, gs: torch.Tensor) -> torch.Tensor:
    """nvfp4_quant.py:104-125 + fp4_kernel_hopper.py:83-84."""
    t = block_max / (6.0 * gs)
    t = torch.where(torch.isnan(t), torch.full_like(t, 448.0), torch.minimum(t, torch.tensor(448.0)))  # min.f32
    s = t.to(torch.float8_e4m3fn).to(torch.float32) * gs
    return torch.where(s >= torch.tensor(1e-5, dtype=torch.float32), s, torch.ones_like(s))


def nvfp4_dynamic_qdq(x: torch.Tensor, global_amax: torch.Tensor, block: int = 16,
                      return_scales: bool = False):
    """Dynamic NVFP4 fake-quant over the last dim (ragged tail zero-padded per block)."""
    shape, dt = x.shape, x.dtype
    x2 = x.reshape(-1, shape[-1]).to(torch.float32)
    M, N = x2.shape
    pad = (-N) % block
    if pad:
        x2 = F.pad(x2, (0, pad))
    gs = nvfp4_global_scale(global_amax.reshape(()))
    xb = x2.view(M, -1, block)
    xa = xb.abs()
    s = nvfp4_block_scale(_max_drop_nan(xa), gs)
    q = e2m1_round_magnitude(xa / s)
    o = q * s
    o = torch.where(xb >= 0, o, -o).view(M, -1)[:, :N]
    out = o.to(dt).reshape(shape)
    if return_scales:
        return out, s.view(M, -1)
    return out


# ----------------------------------------------------------------------------------------------
# NVFP4 static (per-block amax supplied)
# ----------------------------------------------------------------------------------------------
def compute_fp4_scales(amax: torch.Tensor, global_amax=None, quantize_block_scales: bool = True,
                       fp8_max_for_normalization: float = 448.0) -> torch.Tensor:
    """gemm/fp4_kernel.py:217-251."""
    amax = amax.to(torch.float32)
    scale = amax / 6.0
    if quantize_block_scales:
        if global_amax is None:
            global_amax = amax.abs().max()  # reduce_amax(amax), core_utils.py:146-183
        ga = global_amax.to(torch.float32)
        sa = ga * (448.0 / fp8_max_for_normalization) / 6.0
        scale = fp8_qdq(scale, sa)
    return scale


def nvfp4_static_qdq(x: torch.Tensor, amax: torch.Tensor, global_amax=None,
                     quantize_block_scales: bool = True, fp8_max_for_normalization: float = 448.0,
                     out_dtype=None) -> torch.Tensor:
    """gemm/fp4_kernel.py:254-316 + common/nvfp4_quant.py:66-101."""
    shape = x.shape
    nb = amax.numel()
    assert x.numel() % nb == 0
    bs = x.numel() // nb
    out_dtype = out_dtype or x.dtype
    scale = compute_fp4_scales(amax, global_amax, quantize_block_scales, fp8_max_for_normalization)
    scale = scale.reshape(nb, 1)
    xf = x.reshape(nb, bs).to(torch.float32)
    zero = scale == 0.0
    safe = torch.where(zero | torch.isnan(scale) | torch.isinf(scale), torch.ones_like(scale), scale)
    q = e2m1_round_magnitude(xf.abs() / safe)
    o = q * safe
    o = torch.where(xf >= 0, o, -o)
    o = torch.where(zero, torch.zeros_like(o), o)
    return o.to(out_dtype).reshape(shape)
