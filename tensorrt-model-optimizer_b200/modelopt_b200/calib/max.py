"""MaxCalibrator on the B200 path (interface of modelopt/torch/quantization/calib/max.py:26-110).

The reference's ``collect`` runs ``reduce_amax`` (two full reduction passes: max and min), three
``assert torch.any(...)`` device->host syncs and a ``torch.max`` merge per call.  Here the running amax
is a word (or vector) of fp32 BIT PATTERNS — an ``AmaxArena`` slot when the calibration driver provides
an arena, else a private tensor — and ``collect`` is ONE kernel (``mo_amax``: one read of ``x``, one
``atomicMax`` per CTA and channel straight into the running state).  No host sync happens per batch;
the NaN / inf / negative assertions of max.py:69-77 run once, in ``compute_amax`` (same AssertionError
text), because unsigned-max on the bit pattern keeps NaN and inf like ``torch.max`` does.
"""

from __future__ import annotations

import math

import torch

from .. import ops
from .calibrator import _Calibrator

__all__ = ["MaxCalibrator"]


def _kept_axes(nd: int, axis):
    if axis is None:
        return []
    axis = axis if isinstance(axis, (list, tuple)) else [axis]
    return sorted({a % nd for a in axis}) if nd else []


class MaxCalibrator(_Calibrator):
    def __init__(self, num_bits=8, axis=None, unsigned=False, track_amax=False):
        super().__init__(num_bits, axis, unsigned)
        self._track_amax = track_amax
        if self._track_amax:
            self._amaxs = []
        self._arena = None      # AmaxArena shared by every quantizer of a model (set by the driver)
        self._slot = None
        self._slot_arena = None
        self._bits = None       # private state when there is no arena
        self._shape = None      # keepdims shape of the running amax
        self._dtype = None
        self._checked = False

    # ---- state ---------------------------------------------------------------------------------
    def attach_arena(self, arena):
        """Place the running amax in ``arena`` (an AmaxArena or AmaxArenaPool; only before the first collect)."""
        if self._shape is None:
            self._arena = arena

    @property
    def amaxs(self):
        return self._amaxs

    def _state(self, n, device):
        if self._arena is not None:
            arena = self._arena.get(device) if hasattr(self._arena, "get") else self._arena
            if arena.device == device:
                if self._slot is None or self._slot_arena is not arena or arena.bits(self._slot).numel() != n:
                    self._slot, self._slot_arena = arena.alloc(n), arena
                return arena.bits(self._slot)
        if self._bits is None:
            self._bits = torch.zeros(n, dtype=torch.int32, device=device)
        return self._bits

    @property
    def _calib_amax(self):
        """The running amax as the reference exposes it (tensor in the input dtype, keepdims shape)."""
        if self._shape is None:
            return None
        bits = self._slot_arena.bits(self._slot) if self._slot is not None else self._bits
        out = bits.view(torch.float32).to(self._dtype).reshape(self._shape)
        return out.squeeze() if out.numel() == 1 else out

    # ---- interface ------------------------------------------------------------------------------
    @torch.no_grad()
    def collect(self, x):
        if not x.is_cuda:
            raise RuntimeError("modelopt_b200 MaxCalibrator: expected a CUDA tensor; there is no CPU fallback")
        if x.dtype in (torch.float8_e4m3fn, torch.float8_e5m2):
            x = x.to(torch.get_default_dtype())
        x = x.detach()
        nd = x.dim()
        keep = _kept_axes(nd, self._axis)
        shape = tuple(x.shape[d] if d in keep else 1 for d in range(nd)) if self._axis is not None else ()
        if self._shape is None:
            self._shape, self._dtype = shape, x.dtype
        elif self._norm(shape) != self._norm(self._shape):
            raise RuntimeError("amax shape changed!")
        if x.numel() == 0:
            return
        view = ops._collapse_keep_axes(list(x.shape), keep)
        n = math.prod(self._shape) if self._shape else 1
        bits = self._state(n, x.device)
        self._checked = False
        if view is None or x.dtype == torch.float64:
            # kept dims not adjacent / fp64: reduce with the generic wrapper, merge on the bit patterns
            red = [d for d in range(nd) if d not in keep]
            local = ops.reduce_amax(x, axis=red if self._axis is not None else None, keepdims=True,
                                    squeeze_scalar=False).float().reshape(-1)
            torch.maximum(bits, local.view(torch.int32), out=bits)
        else:
            outer, axis_size, inner = view
            ops.amax_into(x, outer, axis_size, inner, bits)
        if self._track_amax:
            red = [d for d in range(nd) if d not in keep] if self._axis is not None else None
            self._amaxs.append(ops.reduce_amax(x, axis=red).cpu().numpy())

    @staticmethod
    def _norm(shape):
        return tuple(shape) if math.prod(shape) != 1 else ()

    def reset(self):
        if self._slot is not None:
            self._slot_arena.reset(self._slot)
        self._bits = None
        self._shape = None
        self._checked = False

    def compute_amax(self):
        amax = self._calib_amax
        if amax is not None and not self._checked:
            bad = torch.stack([torch.isnan(amax).any(), torch.isinf(amax).any()]).tolist()  # one sync
            assert not bad[0], "detected nan values in amax."
            assert not bad[1], "detected inf values in amax."
            self._checked = True
        return amax

    def __str__(self):
        return f"MaxCalibrator(track_amax={self._track_amax})"

    def __repr__(self):
        s = f"MaxCalibrator({super().__repr__()} calib_amax={self._calib_amax} track_amax={self._track_amax}"
        if self._track_amax:
            s += f" amaxs={self._amaxs}"
        return s + ")"
