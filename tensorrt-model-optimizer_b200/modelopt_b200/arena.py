"""Calibration-statistics arena: every quantizer's running amax lives in ONE contiguous device buffer.

The reference keeps one ``_calib_amax`` tensor per ``MaxCalibrator`` and, after ``forward_loop``,
issues one ``all_reduce(MAX)`` per quantizer (``TensorQuantizer.sync_amax_across_distributed_group``,
modelopt/torch/quantization/nn/modules/tensor_quantizer.py:1373-1385 — about 450 latency-bound NCCL
calls for Llama-3-8B).  Here a quantizer's statistic is a *slot* (a view) of a shared int32 buffer that
holds fp32 bit patterns; ``mo_amax`` merges into the slot with one ``atomicMax`` per CTA, and the
data-parallel exchange is ONE ``all_reduce(MAX)`` over the whole arena.

The reduction runs on the int32 view: for non-negative floats the integer order is the numeric order,
+inf sorts above every finite value and NaN patterns above +inf, so the result equals the per-quantizer
float MAX of the reference, NaN/inf survive to the finish-time check, and it is exact (max is
order-independent).
"""

from __future__ import annotations

import torch


class AmaxArena:
    """Growable pool of int32 words (fp32 bit patterns of running |x| maxima) on one device."""

    def __init__(self, device, capacity: int = 4096):
        self.device = torch.device(device)
        self._buf = torch.zeros(capacity, dtype=torch.int32, device=self.device)
        self._used = 0
        self._slots: list[tuple[int, int]] = []

    # -- allocation ------------------------------------------------------------------------------
    def alloc(self, n: int) -> int:
        """Reserve ``n`` words; returns the slot id.  Slots are 8-word aligned so that two quantizers
        never share a 32-byte sector (their atomics do not collide in L2)."""
        start = (self._used + 7) // 8 * 8
        if start + n > self._buf.numel():
            new = torch.zeros(max(2 * self._buf.numel(), start + n), dtype=torch.int32, device=self.device)
            new[: self._used] = self._buf[: self._used]
            self._buf = new  # views handed out earlier are re-derived on access (see bits())
        self._used = start + n
        self._slots.append((start, n))
        return len(self._slots) - 1

    def bits(self, slot: int) -> torch.Tensor:
        """int32 view of a slot (what ``mo_amax`` merges into)."""
        s, n = self._slots[slot]
        return self._buf[s : s + n]

    def amax(self, slot: int) -> torch.Tensor:
        """fp32 view of a slot."""
        return self.bits(slot).view(torch.float32)

    def reset(self, slot: int | None = None):
        if slot is None:
            self._buf.zero_()
        else:
            self.bits(slot).zero_()

    @property
    def used(self) -> torch.Tensor:
        return self._buf[: self._used]

    # -- exchange --------------------------------------------------------------------------------
    def sync(self, group=None):
        """One all-reduce(MAX) over every slot (data-parallel ranks saw different batches)."""
        import torch.distributed as dist

        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return
        if self._used:
            dist.all_reduce(self.used, op=dist.ReduceOp.MAX, group=group)

    # -- finish-time validation (replaces the three per-collect host syncs of calib/max.py:69-77) --
    def check(self):
        """One device->host read: raise like MaxCalibrator.collect does on NaN / inf amax."""
        a = self.used.view(torch.float32)
        bad = torch.stack([torch.isnan(a).any(), torch.isinf(a).any()]).tolist()
        assert not bad[0], "detected nan values in amax"
        assert not bad[1], "detected inf values in amax"


class AmaxArenaPool:
    """One arena per device, created on first use (calibrators learn their device at the first collect)."""

    def __init__(self):
        self._arenas: dict[torch.device, AmaxArena] = {}

    def get(self, device) -> AmaxArena:
        device = torch.device(device)
        if device not in self._arenas:
            self._arenas[device] = AmaxArena(device)
        return self._arenas[device]

    def arenas(self):
        return list(self._arenas.values())

    def sync(self, group=None):
        for a in self._arenas.values():
            a.sync(group)

    def check(self):
        for a in self._arenas.values():
            a.check()
