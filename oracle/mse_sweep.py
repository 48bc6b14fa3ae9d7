"""CPU restatement of the reference's amax searches by squared error.  TEST INFRASTRUCTURE ONLY.

Follows modelopt/torch/quantization/calib/mse.py:
  * MseCalibrator (31-172): candidates linspace(0.25, 4.0, 39); per candidate fake-quantise the fp32 copy of the
    tensor with amax * candidate, accumulate sum((x - xq)^2) over the reduce axes, take the FIRST argmin.
  * NVFP4MSECalibrator (175-311) / Triton nvfp4_fp8_scale_sweep (gemm/nvfp4_fp8_sweep.py:57-110): candidates are the
    126 positive finite E4M3 values / 448 (gemm/_fp8_scale_candidates.py:26-31); scale = c * global_amax / 6;
    loss = sum((|w| - e2m1(|w| / scale) * scale)^2) per block, strict '<' keeps the first minimum;
    best_amax = global_amax * c_best.

Parity pinning: `mse_sweep` is checked against the reference's own MseCalibrator run on CPU (tests/golden/mse_cpu.npz,
made by oracle/gen_golden.py); `nvfp4_fp8_sweep` against outputs of the reference's Triton kernel captured on a B200
(tests/golden/nvfp4_sweep_b200.npz, made by oracle/gen_golden_gpu.py) - the reference has no CPU path for it.
"""

from __future__ import annotations

import math

import torch

from .fake_quant import e2m1_round_magnitude


def mse_candidates(start: float = 0.25, stop: float = 4.0, step: float = 0.1) -> torch.Tensor:
    """calib/mse.py:62,73-74."""
    return torch.linspace(start, stop, steps=math.ceil((stop - start) / step) + 1)


def mse_sweep(x: torch.Tensor, amax: torch.Tensor, qdq, reduce_axis=None, candidates=None):
    """Returns (best_amax, losses[K, ...]).  `qdq(x_f32, amax)` is the quantizer under search."""
    cand = mse_candidates() if candidates is None else candidates
    xf = x.detach().to(torch.float32)
    rows = []
    for c in cand:
        err = (xf - qdq(xf, amax * c)) ** 2
        rows.append(err.sum() if reduce_axis is None else (err.sum(dim=reduce_axis) if reduce_axis else err))
    losses = torch.stack(rows)
    best = torch.argmin(losses, dim=0)
    pick = cand[best]
    if pick.ndim != 0:
        pick = pick.view_as(amax)
    return amax * pick, losses


def fp8_scale_candidates() -> torch.Tensor:
    v = torch.arange(0, 128, dtype=torch.uint8).view(torch.float8_e4m3fn).float()
    return v[torch.isfinite(v) & (v > 0)] / 448.0


def nvfp4_fp8_sweep(x: torch.Tensor, global_amax, block: int = 16, return_losses: bool = False):
    """IEEE-division restatement of the fused sweep; x viewed as [n_blocks, block]."""
    w = x.detach().to(torch.float32).reshape(-1, block).abs()
    ga = torch.as_tensor(global_amax, dtype=torch.float32).reshape(())
    cand = fp8_scale_candidates()
    best_loss = torch.full((w.shape[0],), float("inf"))
    best_idx = torch.zeros(w.shape[0], dtype=torch.long)
    all_losses = []
    for k, c in enumerate(cand):
        scale = c * ga / 6.0
        scale = torch.where(scale == 0, torch.ones_like(scale), scale)
        d = w - e2m1_round_magnitude(w / scale) * scale
        loss = (d * d).sum(dim=1)
        better = loss < best_loss
        best_loss = torch.where(better, loss, best_loss)
        best_idx = torch.where(better, torch.full_like(best_idx, k), best_idx)
        if return_losses:
            all_losses.append(loss)
    best = ga * cand[best_idx]
    return (best, torch.stack(all_losses)) if return_losses else best


def nvfp4_block_loss(x: torch.Tensor, block_amax: torch.Tensor, block: int = 16) -> torch.Tensor:
    """fp64 loss of a chosen per-block amax (scale = amax / 6): the size-independent check the reference's
    tests/gpu/torch/quantization/test_nvfp4_fp8_sweep_kernel.py uses (per-block MSE gap)."""
    w = x.detach().to(torch.float64).reshape(-1, block).abs()
    scale = (block_amax.detach().to(torch.float32) / 6.0).to(torch.float64).reshape(-1, 1)
    scale = torch.where(scale == 0, torch.ones_like(scale), scale)
    q = e2m1_round_magnitude((w / scale).to(torch.float32)).to(torch.float64)
    return ((w - q * scale) ** 2).sum(dim=1)
