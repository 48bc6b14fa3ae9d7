"""MseCalibrator / NVFP4MSECalibrator on the B200 path (interface of modelopt/torch/quantization/calib/mse.py).

``MseCalibrator`` searches the amax multiplier in ``linspace(0.25, 4.0, 39)`` that minimises the squared error.
When it is told what the quantizer does (``fused=dict(num_bits=..., unsigned=..., narrow_range=...)``, set by
``model_calib._make_weight_mse_calibrator`` for INT-N / FP8 quantizers) ``collect`` is ONE kernel that reads the
tensor once and evaluates every candidate (``mo_mse_sweep``); otherwise it falls back to the reference's loop over
``quant_func`` (one fake-quant kernel per candidate, each still a CUDA kernel of this package).

``NVFP4MSECalibrator.collect`` is ONE kernel (``mo_nvfp4_fp8_scale_sweep``) replacing the Triton sweep.
"""

from __future__ import annotations

import math
from collections.abc import Callable

import torch
import torch.nn.functional as F

from .. import ops
from ..utils import convert_quantization_axis_to_reduce_axis
from .calibrator import _Calibrator

__all__ = ["MseCalibrator", "NVFP4MSECalibrator", "fp8_scale_candidates"]


def fp8_scale_candidates(device="cpu") -> torch.Tensor:
    """The 126 positive finite E4M3 values / 448 (gemm/_fp8_scale_candidates.py:26-31)."""
    v = torch.arange(0, 128, dtype=torch.uint8, device=device).view(torch.float8_e4m3fn).float()
    return v[torch.isfinite(v) & (v > 0)] / 448.0


class MseCalibrator(_Calibrator):
    def __init__(self, amax: torch.Tensor, axis=None, step_size: float = 0.1, start_multiplier: float = 0.25,
                 stop_multiplier: float = 4.0, quant_func: Callable | None = None,
                 error_func: Callable | None = None, fused: dict | None = None):
        super().__init__(num_bits=None, axis=axis, unsigned=None)
        self._initial_amax = amax
        self._step_size = step_size
        self._start_multiplier = start_multiplier
        self._stop_multiplier = stop_multiplier
        self._num_steps = math.ceil((stop_multiplier - start_multiplier) / step_size) + 1
        self._quant_func = quant_func
        self._error_func = error_func
        self._fused = fused
        self._losses_sum = None      # list of tensors (loop path) or one float64 [K, C] tensor (fused path)
        self._candidates = None
        self._amax = None

    def _generate_candidates(self, device) -> torch.Tensor:
        return torch.linspace(self._start_multiplier, self._stop_multiplier, steps=self._num_steps, device=device)

    def _compute_candidate_amax(self, candidates: torch.Tensor) -> torch.Tensor:
        if candidates.ndim != 0:
            candidates = candidates.view_as(self._initial_amax)
        return self._initial_amax * candidates

    def _can_fuse(self, x) -> bool:
        if self._fused is None or self._error_func is not None or not x.is_cuda:
            return False
        a = self._initial_amax
        if a.numel() == 1:
            return True
        return a.dim() == x.dim() and list(a.shape).count(1) == x.dim() - 1  # single-axis amax

    @torch.no_grad()
    def collect(self, x: torch.Tensor):
        if self._quant_func is None and self._fused is None:
            raise RuntimeError("Quantization function not set.")
        device = x.device
        candidates = self._generate_candidates(device)
        if self._candidates is None:
            self._candidates = candidates
            self._num_steps = len(candidates)
        if self._can_fuse(x):
            xs = x.detach()
            if xs.dtype == torch.float64:
                xs = xs.float()
            if self._losses_sum is None:
                self._losses_sum = torch.zeros((self._num_steps, self._initial_amax.numel()), dtype=torch.float64,
                                               device=device)
            # the [K, *amax.shape] table uses the reference's own arithmetic (amax dtype x fp32 0-dim candidate)
            table = torch.stack([self._compute_candidate_amax(c) for c in candidates.unbind(0)])
            ops.mse_sweep_into(xs, table, self._losses_sum, **self._fused)
            return
        x = x.detach().to(dtype=torch.float32)
        if self._losses_sum is None:
            self._losses_sum = [None] * self._num_steps
        reduce_axis = convert_quantization_axis_to_reduce_axis(x, self._axis)
        for step, candidate in enumerate(candidates):
            xq = self._quant_func(x, self._compute_candidate_amax(candidate))
            error = self._error_func(x, xq) if self._error_func is not None else F.mse_loss(x, xq, reduction="none")
            loss = error.sum() if reduce_axis is None else (error.sum(dim=reduce_axis) if reduce_axis else error)
            if self._losses_sum[step] is None:
                self._losses_sum[step] = loss.clone()
            else:
                self._losses_sum[step] += loss

    def reset(self):
        self._losses_sum = None
        self._candidates = None
        self._amax = None
        self._initial_amax = None

    @torch.no_grad()
    def compute_amax(self, verbose: bool = False):
        if self._losses_sum is None:
            return None
        if isinstance(self._losses_sum, torch.Tensor):
            losses = self._losses_sum
            best = torch.argmin(losses, dim=0)  # first minimum, like the reference's torch.argmin over stacked losses
            best = best.reshape(()) if self._initial_amax.numel() == 1 else best
        else:
            if not any(l is not None for l in self._losses_sum):
                return None
            first = next(l for l in self._losses_sum if l is not None)
            rows = [l if l is not None else torch.full_like(first, float("inf")) for l in self._losses_sum]
            best = torch.argmin(torch.stack(rows), dim=0)
        self._amax = self._compute_candidate_amax(self._candidates[best])
        if verbose:
            ratio = self._amax / self._initial_amax
            print(f"MSE Calibrator: best_amax/initial_amax ratio - mean: {ratio.float().mean().item():.4f}")
        return self._amax


class NVFP4MSECalibrator(MseCalibrator):
    """Per-block FP8-scale sweep for static NVFP4 weights (calib/mse.py:175-311); one-shot between resets."""

    def __init__(self, amax: torch.Tensor, global_amax: torch.Tensor, axis=None, quant_func: Callable | None = None,
                 error_func: Callable | None = None, hessian: torch.Tensor | None = None):
        super().__init__(amax=amax, axis=axis, quant_func=quant_func, error_func=error_func)
        self._hessian = hessian  # [cin // block, block, block]: the Hessian-weighted sweep (local_hessian), mse.py:205-207
        self._global_amax = global_amax.to(dtype=torch.float32)
        self._best_amax = None

    def _generate_candidates(self, device):
        return fp8_scale_candidates(device)

    def _compute_candidate_amax(self, candidates):
        if candidates.ndim != 0:
            candidates = candidates.view_as(self._initial_amax)
        return torch.ones_like(self._initial_amax, dtype=torch.float32) * (self._global_amax * candidates)

    @torch.no_grad()
    def collect(self, x: torch.Tensor):
        if self._best_amax is not None:
            raise RuntimeError("NVFP4MSECalibrator: a previous collect() call produced a final amax; "
                               "multi-collect is not supported. Call reset() to start a fresh cycle.")
        blocked = x.is_cuda and x.ndim == 2 and x.shape[0] == int(self._initial_amax.numel())
        if self._hessian is not None and blocked and x.shape[-1] == 16:  # wins over the plain path (mse.py:258-266)
            best = ops.nvfp4_fp8_scale_sweep_hessian(x.detach(), self._global_amax, self._hessian, block_size=x.shape[-1])
            self._best_amax = best.reshape(self._initial_amax.shape).to(torch.float32)
            return
        if self._error_func is None and blocked:
            best = ops.nvfp4_fp8_scale_sweep(x.detach(), self._global_amax, block_size=x.shape[-1])
            self._best_amax = best.reshape(self._initial_amax.shape).to(torch.float32)
            return
        super().collect(x)
        best = super().compute_amax(verbose=False)
        self._best_amax = best.to(torch.float32) if best is not None else None
        self._losses_sum = None

    def compute_amax(self, verbose: bool = False):
        return self._best_amax

    def reset(self):
        self._best_amax = None
        self._losses_sum = None
        self._candidates = None
        self._amax = None
