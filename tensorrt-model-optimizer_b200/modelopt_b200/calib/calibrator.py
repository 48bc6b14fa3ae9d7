"""Calibrator base class (interface of modelopt/torch/quantization/calib/calibrator.py:25-68)."""

from __future__ import annotations

from typing import Any

import torch

__all__ = ["_Calibrator"]


class _Calibrator:
    """collect(x) gathers statistics, compute_amax() turns them into an amax tensor, reset() drops them.

    ``axis`` is the quantization axis (the dims that KEEP their own amax), as in
    ``QuantizerAttributeConfig.axis``; ``None`` is per-tensor."""

    def __init__(self, num_bits=8, axis=None, unsigned=False):
        self._num_bits = num_bits
        self._axis = axis
        self._unsigned = unsigned

    def __init_subclass__(cls, **kwargs: Any) -> None:
        super().__init_subclass__(**kwargs)
        torch.serialization.add_safe_globals([cls])  # loadable with torch.load(weights_only=True)

    def collect(self, x):
        raise NotImplementedError

    def reset(self):
        raise NotImplementedError

    def compute_amax(self, *args, **kwargs):
        raise NotImplementedError

    def __repr__(self):
        return f"num_bits={self._num_bits} axis={self._axis} unsigned={self._unsigned}"
