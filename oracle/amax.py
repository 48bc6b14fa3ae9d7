"""Oracle: amax reductions and the max calibrator (CPU torch).  TEST INFRASTRUCTURE ONLY.

Restates:
* ``reduce_amax``                                 modelopt/torch/quantization/utils/core_utils.py:146-183
* ``reduce_block_amax``                           .../utils/core_utils.py:43-89
* ``convert_quantization_axis_to_reduce_axis``    .../utils/core_utils.py:128-143
* ``MaxCalibrator``                               modelopt/torch/quantization/calib/max.py:26-110
"""

from __future__ import annotations

import torch


def quant_axis_to_reduce_axis(x: torch.Tensor, axis):
    """core_utils.py:128-143."""
    if axis is None:
        return None
    axis = axis if isinstance(axis, (list, tuple)) else [axis]
    return [i for i in range(x.dim()) if i not in axis and (i - x.dim()) not in axis]


def reduce_amax(x: torch.Tensor, axis=None, keepdims=True, squeeze_scalar=True) -> torch.Tensor:
    """core_utils.py:146-183: max(|max(x)|, |min(x)|); NaN propagates (torch.max semantics)."""
    if x.dtype in (torch.float8_e4m3fn, torch.float8_e5m2):
        x = x.to(torch.get_default_dtype())
    if axis is None:
        return torch.maximum(torch.abs(torch.max(x)), torch.abs(torch.min(x)))
    if isinstance(axis, int):
        axis = (axis,)
    axis = tuple(axis)
    if len(axis) == 0:
        out = x.abs()
    else:
        mx = torch.amax(x, dim=axis, keepdim=keepdims)
        mn = torch.amin(x, dim=axis, keepdim=keepdims)
        out = torch.maximum(torch.abs(mx), torch.abs(mn))
    if squeeze_scalar and out.numel() == 1:
        out = out.squeeze()
    return out


def reduce_block_amax(x: torch.Tensor, block_sizes: dict) -> torch.Tensor:
    """core_utils.py:43-89."""
    amax = x.clone()
    for dim, bs in block_sizes.items():
        dim = dim if dim >= 0 else amax.dim() + dim
        assert amax.shape[dim] % bs == 0
        new_shape = [*amax.shape[:dim], amax.shape[dim] // bs, bs, *amax.shape[dim + 1:]]
        amax = reduce_amax(amax.reshape(new_shape), dim + 1, keepdims=False, squeeze_scalar=False)
    return amax


class MaxCalibrator:
    """calib/max.py:26-110."""

    def __init__(self, num_bits=8, axis=None, unsigned=False, track_amax=False):
        self._num_bits, self._axis, self._unsigned = num_bits, axis, unsigned
        self._track_amax = track_amax
        self._amaxs = []
        self._calib_amax = None

    def collect(self, x: torch.Tensor):
        local = reduce_amax(x, axis=quant_axis_to_reduce_axis(x, self._axis)).detach()
        assert not torch.any(torch.isnan(local)), "detected nan values in amax"
        assert torch.all(local >= 0), "detected negative values after abs"
        assert not torch.any(torch.isinf(local)), "detected inf values in amax"
        if self._calib_amax is None:
            self._calib_amax = local
        else:
            if local.shape != self._calib_amax.shape:
                raise RuntimeError("amax shape changed!")
            self._calib_amax = torch.max(self._calib_amax, local)
        if self._track_amax:
            self._amaxs.append(local.cpu().numpy())

    def reset(self):
        self._calib_amax = None

    def compute_amax(self):
        return self._calib_amax
