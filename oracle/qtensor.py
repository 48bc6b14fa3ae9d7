"""CPU restatement of the reference's real-quant packers.  TEST INFRASTRUCTURE ONLY (imported by tests/, never by the
product).  Parity: pinned against the reference's own CPU outputs (tests/golden/qtensor_cpu.npz, generator
oracle/gen_golden_qtensor.py) and the known-answer vectors of its tests/gpu/torch/quantization/test_qtensor_cuda.py.

Every function works on torch CPU tensors in the tensor's own dtype, so bf16 / fp16 intermediates round where the
reference's eager ops round."""

from __future__ import annotations

import math

import torch
import torch.nn.functional as F

E2M1_BOUNDS = torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5.0])
E2M1_VALUES = torch.tensor([0, 0.5, 1, 1.5, 2, 3, 4, 6.0])
NF4 = torch.tensor([-1.0, -0.6962, -0.5251, -0.3949, -0.2844, -0.1848, -0.0911, 0.0, 0.0796, 0.1609, 0.2461, 0.3379,
                    0.4407, 0.5626, 0.7230, 1.0], dtype=torch.bfloat16)  # nf4_tensor.py:25-45 (held in bf16)


def _absmax(t, dim):
    return t.abs().amax(dim=dim, keepdim=True)  # == max(|max|, |min|), core_utils.py:146-183


def _pad_flat(x, block):
    flat = x.reshape(-1)
    pad = (-flat.numel()) % block
    return F.pad(flat, (0, pad)) if pad else flat


# ---------------------------------------------------------------------------------------------------------------------
def int4_quantize(x, block):
    """int4_tensor.py:39-88, eager branch: scales = 7 / amax (reciprocal * 7 in x.dtype), q = round(x * s).clamp(-8, 7) + 8,
    byte = even << 4 | odd."""
    blocks = _pad_flat(x, block).view(-1, block)
    scales = 7 / _absmax(blocks, -1)
    u = ((blocks * scales).flatten().round().clamp(-8, 7) + 8).to(torch.uint8)
    return (u[0::2] << 4 | u[1::2]).reshape(*x.shape[:-1], -1), scales


def int4_dequantize(q, scales, block, shape, dtype):
    """int4_tensor.py:108-129."""
    hi = (q >> 4).to(torch.long) - 8
    lo = (q & 0xF).to(torch.long) - 8
    hi = hi.view(-1, block // 2) / scales.view(-1, 1)
    lo = lo.view(-1, block // 2) / scales.view(-1, 1)
    out = torch.stack([hi.flatten(), lo.flatten()], dim=-1).view(-1)
    return out[: math.prod(shape)].reshape(shape).to(dtype)


def nf4_quantize(x, block, scale_block=None):
    """nf4_tensor.py:75-124, eager branch: index of the first nearest table entry of x / amax, all in x.dtype."""
    blocks = _pad_flat(x, block).view(-1, block)
    scales = _absmax(blocks, -1)
    scaled = (blocks / scales).flatten()
    idx = (scaled[:, None] - NF4.to(x.dtype)).abs().argmin(dim=-1).to(torch.uint8)
    scales = _pad_flat(scales, scale_block) if scale_block else scales.view(-1)
    return idx[0::2] << 4 | idx[1::2], scales


def nf4_double_quantization(scales, scale_block, bits):
    """nf4_tensor.py:126-151."""
    bound = 2 ** (bits - 1) - 1
    b = scales.view(-1, scale_block)
    zp = b.mean()
    b = b - zp
    ds = bound / _absmax(b, -1)
    return (b * ds).round().clamp(-bound, bound).to(torch.int8), ds.flatten(), zp


def nf4_dequantize(q, scales_q, double_scale, zeros, block, shape, dtype):
    """nf4_tensor.py:153-199, eager branch."""
    s = scales_q.view(-1)[: q.numel() * 2 // block]
    s = ((s / double_scale.unsqueeze(-1)).to(dtype) + zeros).flatten()
    hi = NF4[(q >> 4).to(torch.long)].view(-1, block // 2) * s.view(-1, 1)
    lo = NF4[(q & 0xF).to(torch.long)].view(-1, block // 2) * s.view(-1, 1)
    out = torch.stack([hi.flatten(), lo.flatten()], dim=-1).view(-1)
    return out[: math.prod(shape)].reshape(shape).to(dtype)


# ---------------------------------------------------------------------------------------------------------------------
def _pad_dims(x, block_sizes):
    pad = [0] * (2 * x.dim())
    for d, b in block_sizes.items():
        d = d % x.dim()
        pad[(x.dim() - 1 - d) * 2 + 1] = (-x.shape[d]) % b
    return F.pad(x, pad) if any(pad) else x


def _block_amax(x, block_sizes):
    a = x
    for d, b in block_sizes.items():
        d = d % a.dim()
        a = a.reshape(*a.shape[:d], a.shape[d] // b, b, *a.shape[d + 1:]).abs().amax(dim=d + 1)
    return a


def _expand(scales, block_sizes):
    for d, b in (block_sizes or {}).items():
        scales = scales.repeat_interleave(b, dim=d)
    return scales


def scaled8_quantize(x, kind, axis=None, block_sizes=None, scales=None):
    """int8_tensor.py:35-93 (kind 'int8': scales = amax / 127, q = round(x / s).clamp(-128, 127)) and
    fp8_tensor.py:38-115 (kind 'fp8': scales = amax / 448, q = e4m3(x / s), input padded to whole blocks)."""
    orig = x
    if block_sizes and (kind == "fp8" or scales is None):
        x = _pad_dims(x, block_sizes)
    if scales is None:
        if block_sizes:
            amax = _block_amax(x, block_sizes)
        elif axis is None:
            amax = x.abs().amax()
        else:
            ax = [axis] if isinstance(axis, int) else list(axis)
            red = [i for i in range(x.dim()) if i not in ax and i - x.dim() not in ax]
            amax = x.abs().amax(dim=red, keepdim=True)
        scales = amax / (127.0 if kind == "int8" else 448.0)
    y = x / _expand(scales, block_sizes)
    q = y.round().clamp(-128, 127).to(torch.int8) if kind == "int8" else y.to(torch.float8_e4m3fn)
    return q[tuple(slice(0, d) for d in orig.shape)], scales


def scaled8_dequantize(q, scales, dtype, shape, block_sizes=None, kind="int8"):
    """int8_tensor.py:95-124, fp8_tensor.py:117-151: q.to(dtype) * scales.to(dtype)."""
    if block_sizes and kind == "fp8":
        q = _pad_dims(q.view(torch.uint8), block_sizes).view(q.dtype)
    s = _expand(scales, block_sizes)
    return (q.to(dtype) * s.to(dtype))[tuple(slice(0, d) for d in shape)]


# ---------------------------------------------------------------------------------------------------------------------
def mxfp4_quantize(x, block=32):
    """mxfp4_tensor.py:36-83: e = ceil(max(log2(amax / 6), -127)); code = sign_bit << 3 | #{bounds < |x / 2^e|};
    sign_bit = (2 - sign(x)) // 2 (so a zero carries it); even element in the low nibble."""
    b = x.reshape(-1, block)
    e = torch.ceil(torch.maximum(torch.log2(b.float().abs().amax(dim=-1, keepdim=True) / 6.0), torch.tensor(-127.0)))
    v = (b / torch.exp2(e)).view(x.shape)
    code = (((2 - torch.sign(v)) // 2) * 8 + ((v.abs().unsqueeze(-1) - E2M1_BOUNDS) > 0).sum(-1)).to(torch.uint8)
    return code[..., 1::2] << 4 | code[..., 0::2], (e + 127).to(torch.uint8)


def mxfp4_dequantize(q, scale, block, dtype):
    """mxfp4_tensor.py:85-144."""
    nib = torch.stack([q & 0xF, q >> 4], dim=-1).reshape(*q.shape[:-1], -1)
    val = E2M1_VALUES[(nib & 7).long()] * (1 - 2 * ((nib >> 3) & 1).float())
    out = val.reshape(-1, block) * torch.exp2(scale.float() - 127).reshape(-1, 1)
    return out.reshape(nib.shape).to(dtype)


def mxfp8_scale(amax):
    """mxfp8_tensor.py:41-66."""
    d = amax.float() / 448.0
    e = torch.ceil(torch.where(d > 0, torch.log2(d), torch.tensor(-127.0)))
    return (e.clamp(-127, 127) + 127).to(torch.uint8)


def mxfp8_quantize(x, scale=None):
    """mxfp8_tensor.py:141-229: q = e4m3(clamp(x * 2^(127 - sf), +-448)) per 32-block of the last dim (zero-padded)."""
    n = x.shape[-1]
    xp = F.pad(x, (0, (-n) % 32))
    blocks = xp.reshape(*xp.shape[:-1], -1, 32)
    if scale is None:
        scale = mxfp8_scale(blocks.abs().amax(dim=-1))
    y = (blocks * torch.exp2(127 - scale.float()).unsqueeze(-1)).clamp(-448.0, 448.0).to(torch.float8_e4m3fn)
    return y.reshape(xp.shape)[..., :n], scale


def mxfp8_dequantize(q, scale, dtype):
    """mxfp8_tensor.py:231-262."""
    n = q.shape[-1]
    qp = F.pad(q.float(), (0, (-n) % 32))
    out = qp.reshape(*qp.shape[:-1], -1, 32) * torch.exp2(scale.float() - 127).unsqueeze(-1)
    return out.reshape(qp.shape)[..., :n].to(dtype)
