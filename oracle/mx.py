"""CPU restatement of the reference's MX dynamic block fake-quant (`fused_amax_convert`,
modelopt/torch/kernels/quantization/gemm/tensor_quant_mx.cu:36-387, tensor_quant_mx.h:39-245).
TEST INFRASTRUCTURE ONLY.

Parity: the element rounding is pinned against the reference's own header compiled with g++ (oracle/_ref/libref_mx.so,
oracle/build_ref.py); whole-tensor results are pinned against the reference's CUDA extension run on a B200
(tests/golden/mx_b200.npz, oracle/gen_golden_gpu_mx.py).  The reference has no CPU path for this op, and its GPU build
uses --use_fast_math (approximate fp32 quotients): this restatement divides exactly (= the product's MO_NUMERICS_CPU),
which differs from the captured GPU outputs only in blocks whose amax / format_max is within 2 ulp of a power of two.
"""

from __future__ import annotations

import numpy as np
import torch

FMT = {"E4M3": 0, "E5M2": 1, "INT8": 2, "E0M3": 3, "E1M2": 4, "E3M0": 5, "E2M1": 6, "E3M2": 7, "E2M3": 8, "E8M0": 9}
FMT_MAX = {"E4M3": 448.0, "E5M2": 57344.0, "INT8": 127.0, "E2M1": 6.0, "E0M3": 7.0, "E3M0": 16.0, "E1M2": 3.5,
           "E3M2": 28.0, "E2M3": 7.5}  # tensor_quant_mx.h:177-201

_TABLES = {  # tensor_quant_mx.h:42-68
    "E2M1": [0, 0.5, 1, 1.5, 2, 3, 4, 6],
    "E1M2": [0, 0.5, 1, 1.5, 2, 2.5, 3, 3.5],
    "E0M3": [0, 1, 2, 3, 4, 5, 6, 7],
    "E3M0": [0, 0.25, 0.5, 1, 2, 4, 8, 16],
    "E3M2": [0, 0.0625, 0.125, 0.1875, 0.25, 0.3125, 0.375, 0.4375, 0.5, 0.625, 0.75, 0.875, 1, 1.25, 1.5, 1.75, 2, 2.5,
             3, 3.5, 4, 5, 6, 7, 8, 10, 12, 14, 16, 20, 24, 28],
    "E2M3": [0, 0.125, 0.25, 0.375, 0.5, 0.625, 0.75, 0.875, 1, 1.125, 1.25, 1.375, 1.5, 1.625, 1.75, 1.875, 2, 2.25,
             2.5, 2.75, 3, 3.25, 3.5, 3.75, 4, 4.5, 5, 5.5, 6, 6.5, 7, 7.5],
}


def convert(a: torch.Tensor, fmt: str) -> torch.Tensor:
    """convert_to_types on NON-NEGATIVE fp32 magnitudes (tensor_quant_mx.h:70-175)."""
    a = a.float()
    if fmt == "E4M3":  # saturating RNE
        return a.clamp(max=448.0).to(torch.float8_e4m3fn).float()
    if fmt == "E5M2":
        return a.clamp(max=57344.0).to(torch.float8_e5m2).float()
    if fmt == "INT8":  # rint then clamp to +-127
        return torch.where(torch.isnan(a), torch.zeros_like(a), torch.round(a).clamp(max=127.0))
    vals = torch.tensor(_TABLES[fmt], dtype=torch.float32)
    bounds = (vals[:-1] + vals[1:]) / 2  # the header's *_table_bounds are the midpoints
    idx = torch.searchsorted(bounds, a.contiguous(), right=False)  # number of bounds < a
    tie = torch.zeros_like(a, dtype=torch.bool)
    hit = idx < len(bounds)
    tie[hit] = bounds[idx[hit]] == a[hit]
    if fmt == "E3M0":  # ties away from zero (:84-93)
        idx = idx + tie.long()
    else:  # ties to the entry with an even index (:96-117)
        idx = idx + (tie & (idx % 2 == 1)).long()
    idx = torch.where(torch.isnan(a) | torch.isinf(a), torch.full_like(idx, len(vals) - 1), idx)
    return vals[idx.clamp(max=len(vals) - 1)]


def _pow2(e: torch.Tensor) -> torch.Tensor:
    return torch.ldexp(torch.ones_like(e, dtype=torch.float32), e.to(torch.int32))


def e8m0_exponent(amax: torch.Tensor, dmax: float) -> torch.Tensor:
    """compute_scale_e8m0_NV (tensor_quant_mx.cu:104-132) on the bits of the (exact) quotient."""
    ratio = (amax.float() / torch.tensor(dmax, dtype=torch.float32)).contiguous()
    u = ratio.view(torch.int32)
    ef = (u >> 23) & 0xFF
    sf = u & 0x7FFFFF
    up = (sf > 0) & (ef != 0xFE) & ~((ef == 0) & (sf <= 0x400000))
    return torch.where(up, ef - 126, ef - 127)


def mx_fake_quant(x: torch.Tensor, block: int, fmt: str, scale_fmt: str = "E8M0", global_amax=None) -> torch.Tensor:
    """fused_amax_convert: last dim in blocks of `block` (zero-padded), per-block amax -> (scale, unscale) ->
    y = sign(x) * convert(|x| * scale) * unscale, cast to x.dtype."""
    shape, dt = x.shape, x.dtype
    n = shape[-1]
    xf = x.float().reshape(-1, n)
    pad = (-n) % block
    if pad:
        xf = torch.nn.functional.pad(xf, (0, pad))
    b = xf.reshape(xf.shape[0], -1, block)
    amax = b.abs().amax(dim=-1, keepdim=True)
    bad = (amax == 0) | torch.isinf(amax) | torch.isnan(amax)
    safe = torch.where(bad, torch.ones_like(amax), amax)
    if scale_fmt == "E8M0":
        e = e8m0_exponent(safe, FMT_MAX[fmt])
        scale, unscale = _pow2(-e), _pow2(e)
    elif global_amax is None:  # compute_scale general branch (:145-151)
        s = (torch.tensor(FMT_MAX[fmt]) / safe).double()
        inv = convert((1.0 / s).float(), scale_fmt).double()
        scale, unscale = (1.0 / inv).float(), inv.float()
    else:  # compute_scale_with_global (:154-181)
        ga = global_amax.float().reshape(())
        if ga == 0 or torch.isinf(ga) or torch.isnan(ga):
            bad = torch.ones_like(bad)
        emax, smax = torch.tensor(FMT_MAX[fmt]), torch.tensor(FMT_MAX[scale_fmt])
        local = safe / emax
        two = smax.double() * (emax / ga).double()
        lq = convert((local.double() * two).float(), scale_fmt).double() / two
        scale, unscale = (1.0 / lq).float(), lq.float()
    one = torch.ones_like(scale)
    scale, unscale = torch.where(bad, one, scale), torch.where(bad, one, unscale)
    q = convert((b.abs() * scale).reshape(-1), fmt).reshape(b.shape) * unscale
    y = torch.where(b < 0, -q, q).reshape(xf.shape)[:, :n]
    return y.reshape(shape).to(dt)


def ref_lib():
    """The reference's own header behind a C ABI (None when oracle/_ref/libref_mx.so has not been built)."""
    import ctypes
    import os

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_mx.so")
    if not os.path.exists(path):
        return None
    lib = ctypes.CDLL(path)
    lib.ref_mx_convert_many.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long, ctypes.c_int]
    lib.ref_mx_format_max.restype = ctypes.c_float
    lib.ref_mx_format_max.argtypes = [ctypes.c_int]
    return lib


def ref_convert(a: np.ndarray, fmt: str) -> np.ndarray:
    lib = ref_lib()
    a = np.ascontiguousarray(a, dtype=np.float32)
    out = np.empty_like(a)
    lib.ref_mx_convert_many(a.ctypes.data, out.ctypes.data, a.size, FMT[fmt])
    return out
