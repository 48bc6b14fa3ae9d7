"""CPU restatement of the SmoothQuant / AWQ-lite arithmetic.  TEST INFRASTRUCTURE ONLY.

Follows modelopt/torch/quantization/model_calib.py:
  * smoothquant.postprocess (1304-1330) and apply_pre_quant_scale_and_smooth (1226-1270)
  * awq_lite: get_weight_scale (1453-1469), get_act_scale (1471-1472), get_scale (1474-1483), the search forward
    (1537-1551: weight_quantizer(W * s) with per-call block amax, input * (1/s)), update_loss (1485-1491)
  * awq_clip: clip ratios (1761-1763), the blocked `_clip_search` (1813-1858), update_best_params (1787-1798)
Parity pinning: tests/test_oracle_smooth.py reproduces the pre_quant_scale / folded weight / amax the REAL reference
produced with mtq.quantize on CPU (tests/golden/quantize_cpu.npz, awq_clip_cpu.npz) from these functions.
"""

from __future__ import annotations

import torch
import torch.nn.functional as F

from .fake_quant import int_qdq


def smooth_scale(act_amax: torch.Tensor, weight: torch.Tensor, alpha: float = 1.0) -> torch.Tensor:
    act_amax = act_amax.float()
    weight_scale = weight.abs().amax(dim=0, keepdim=True)
    scale_a = (weight_scale.pow(1 - alpha) / act_amax.pow(alpha)).squeeze()
    eps = 1.0 / (1 << 31)
    if scale_a.min() <= eps:
        scale_a[(act_amax <= eps).reshape(-1)] = 1
    return scale_a.clamp(min=1e-4, max=1e4)


def fold_weight(weight: torch.Tensor, col_scale: torch.Tensor) -> torch.Tensor:
    """_apply_weight_pre_quant_scale (1208-1223)."""
    return (weight * col_scale.squeeze()[None, :]).to(weight.dtype)


def awq_weight_scale(weight: torch.Tensor, block_size=None) -> torch.Tensor:
    org_shape, sl, w = weight.shape, None, weight
    if block_size:
        if org_shape[-1] % block_size != 0:
            sl = slice(org_shape[-1])
            w = F.pad(w, (0, block_size - org_shape[-1] % block_size), "constant", 0)
            org_shape = w.shape
        w = w.contiguous().view(-1, block_size)
    wa = w.abs()
    scale = (wa / (wa.amax(dim=1, keepdim=True) + torch.finfo(w.dtype).tiny)).view(org_shape)
    if sl is not None:
        scale = scale[..., sl]
    return scale.mean(0).to(torch.float32)


def awq_act_scale(x: torch.Tensor) -> torch.Tensor:
    return x.abs().contiguous().view(-1, x.shape[-1]).mean(0).to(torch.float32)


def awq_get_scale(x_max, w_max, alpha: float) -> torch.Tensor:
    s = (x_max.pow(alpha) / (w_max.pow(1 - alpha) + torch.finfo(torch.float32).tiny)).clamp(min=1e-4, max=1e4).view(-1)
    return (s / (s.max() * s.min()).sqrt()).view(-1)


def int_block_dynamic_qdq(x: torch.Tensor, block: int, num_bits: int = 4, prescale=None, unsigned=False,
                          narrow_range=False) -> torch.Tensor:
    """TensorQuantizer.forward with pre_quant_scale, static block_sizes {-1: block} and no stored amax
    (tensor_quantizer.py:1143, :736-751, :975-1061): amax of every block is taken from the scaled data."""
    p = x if prescale is None else x * prescale.to(x.dtype)
    v = p.contiguous().view(-1, block)
    amax = torch.maximum(v.amax(dim=1, keepdim=True).abs(), v.amin(dim=1, keepdim=True).abs())
    return int_qdq(v, amax, num_bits, unsigned, narrow_range).view(x.shape)


def awq_lite_linear(weight, bias, batches, block: int, num_bits: int = 4, alpha_step: float = 0.1):
    """The whole awq_lite search for ONE weight-only linear.  Returns (alphas, losses, best_alpha, best_scale)."""
    w_scale = awq_weight_scale(weight, block)
    act = sum(awq_act_scale(b) for b in batches) / len(batches)
    alphas = [k.item() for k in torch.arange(0, 1.0 + alpha_step, alpha_step)]
    losses = [torch.zeros(()) for _ in alphas]
    for b in batches:
        ref = F.linear(b, weight, bias)
        for i, a in enumerate(alphas):
            s = awq_get_scale(act, w_scale, a)
            wq = int_block_dynamic_qdq(weight, block, num_bits, prescale=s.to(weight.dtype))
            out = F.linear(b * (1 / s).to(weight.dtype), wq, bias)
            losses[i] += (out - ref).float().pow(2).mean()
    vals = [float(l) for l in losses]
    best = alphas[min(range(len(alphas)), key=vals.__getitem__)]
    return alphas, vals, best, awq_get_scale(act, w_scale, best)


def awq_clip_ratios(min_clip_ratio: float = 0.5, shrink_step: float = 0.05) -> list[float]:
    """model_calib.py:1761-1763."""
    return [round(float(k), 2) for k in torch.arange(min_clip_ratio, 1.0, shrink_step)] + [1.0]


def awq_clip_block_losses(weight, tokens, w_amax, ratios, block: int, num_bits: int = 4) -> torch.Tensor:
    """Blocked `_clip_search` for ONE batch of already sampled tokens (model_calib.py:1813-1858), without the
    out-channel batching (it only bounds memory).  Returns [len(ratios), co, n_blocks] in fp32."""
    co, ci = weight.shape
    pad = (-ci) % block
    w = F.pad(weight, (0, pad)) if pad else weight
    x = F.pad(tokens, (0, pad)) if pad else tokens
    wb = w.reshape(co, 1, -1, block)
    xb = x.reshape(1, x.shape[0], -1, block)
    org = (xb * wb).sum(dim=-1)                                           # [co, T, nb]
    out = []
    for r in ratios:
        amax = (w_amax * r).reshape(co, -1, 1)
        wq = int_qdq(w.reshape(co, -1, block), amax, num_bits, False, False).reshape(co, 1, -1, block)
        cur = (xb * wq).sum(dim=-1)
        out.append((cur - org).float().pow(2).mean(dim=1))                # [co, nb]
    return torch.stack(out)


def awq_clip_select(losses: torch.Tensor, w_amax: torch.Tensor, ratios) -> torch.Tensor:
    """update_best_params (model_calib.py:1787-1798): first strict minimum over the ratios, in order."""
    best_loss = torch.full_like(w_amax, float("inf"))
    best = torch.zeros_like(w_amax)
    for k, r in enumerate(ratios):
        l = losses[k].view_as(w_amax)
        take = l < best_loss
        best_loss = torch.where(take, l, best_loss)
        best = torch.where(take, w_amax * r, best)
    return best
