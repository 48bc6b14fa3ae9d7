"""Calibration algorithms (interface of modelopt/torch/quantization/model_calib.py).

``max_calibrate`` keeps the reference's sequence — enable_stats_collection, weight_only_quantize, forward_loop,
finish_stats_collection, data-parallel sync (model_calib.py:310-498) — with two B200-specific changes that do not
alter any result:

* every MaxCalibrator's running amax lives in one ``AmaxArena`` per device, so the data-parallel exchange is ONE
  all-reduce(MAX) over the arena instead of one NCCL call per quantizer, and the NaN/inf validation is ONE
  device->host read instead of three per ``collect``;
* nothing on the per-batch path synchronises with the host.
"""

from __future__ import annotations

import warnings

import torch
import torch.distributed as dist
from torch import nn

from .arena import AmaxArenaPool
from .calib import MaxCalibrator
from .nn.quant_module import QuantModule
from .nn.tensor_quantizer import SequentialQuantizer, TensorQuantizer

__all__ = ["max_calibrate", "enable_stats_collection", "finish_stats_collection", "weight_only_quantize",
           "disable_pre_quant_scale_and_resmooth", "sync_histograms"]


def weight_only_quantize(model: nn.Module):
    """model_calib.py:187-199: run every weight through its weight quantizer once."""
    seen = set()
    for module in model.modules():
        if module in seen:
            continue
        if isinstance(module, QuantModule):
            for weight, q in module.iter_weights_for_calibration():
                q(weight)
        seen.add(module)


def enable_stats_collection(model: nn.Module, arena_pool: AmaxArenaPool | None = None):
    """model_calib.py:1128-1141 (+ attach the shared amax arena to every max calibrator)."""
    for _, module in model.named_modules():
        if isinstance(module, TensorQuantizer) and not module._disabled:
            if module._use_constant_amax or module._constant_amax is not None:
                module.disable_quant()
                continue
            if module._calibrator is not None:
                module.disable_quant()
                module.enable_calib()
                if arena_pool is not None and isinstance(module._calibrator, MaxCalibrator):
                    module._calibrator.attach_arena(arena_pool)
            else:
                module.disable()


def finish_stats_collection(model: nn.Module, method: str | None = None, **kwargs):
    """model_calib.py:1144-1167."""
    for _, module in model.named_modules():
        if not isinstance(module, TensorQuantizer) or module._disabled:
            continue
        if module._use_constant_amax or module._constant_amax is not None:
            module.enable_quant()
            continue
        cal = getattr(module, "_calibrator", None)
        if cal and not getattr(module, "_dynamic", False):
            if method == "entropy":
                if cal.compute_amax(method) is not None:
                    module.load_calib_amax("entropy", **kwargs)
            elif cal.compute_amax(**kwargs) is not None:
                module.load_calib_amax(**kwargs)
        module.enable_quant()
        module.disable_calib()


def _dp_group(model):
    for m in model.modules():
        if isinstance(m, QuantModule) and m.parallel_state is not None:
            g = m.parallel_state.data_parallel_group
            return None if g == -1 else g, g != -1
    return None, True


@torch.no_grad()
def max_calibrate(model: nn.Module, forward_loop=None, distributed_sync=True, **_unused):
    """model_calib.py:310-498 (data-parallel sync only; TP / EP groups are outside this path)."""
    pool = AmaxArenaPool()
    enable_stats_collection(model, pool)
    weight_only_quantize(model)
    if forward_loop is not None:
        forward_loop(model)
    # the one exchange: amax of every quantizer, all-reduce MAX over the data-parallel group
    if distributed_sync and dist.is_available() and dist.is_initialized():
        group, ok = _dp_group(model)
        if ok:
            pool.sync(group)
    pool.check()  # one host read: AssertionError on NaN / inf like MaxCalibrator.collect (max.py:69-77)
    for m in model.modules():
        if isinstance(m, TensorQuantizer) and isinstance(getattr(m, "_calibrator", None), MaxCalibrator):
            m._calibrator._checked = True
    finish_stats_collection(model)
    if distributed_sync and dist.is_available() and dist.is_initialized():
        # quantizers calibrated outside the arena (custom calibrators): the reference's per-quantizer sync
        group, ok = _dp_group(model)
        for m in model.modules():
            if ok and isinstance(m, TensorQuantizer) and not m._disabled and getattr(m, "_amax", None) is not None \
                    and not isinstance(getattr(m, "_calibrator", None), MaxCalibrator):
                m.sync_amax_across_distributed_group(group)


def sync_histograms(model: nn.Module, group=None):
    """All-reduce-SUM of the histograms of every HistogramCalibrator over the data-parallel group.

    New behaviour (the reference does not synchronise histograms, calib/histogram.py:158-163).  Ranks first agree
    on a common range (all-reduce MAX of the number of bins: every rank's bins have the width fixed by ITS first
    batch, so the widths are exchanged too and must agree — they do when every rank saw the same first batch
    statistics, otherwise this raises), then integer counts are summed exactly.  World size 1 is a no-op, i.e. the
    reference's behaviour bit for bit."""
    from .calib import HistogramCalibrator

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    cals = [m._calibrator for m in model.modules()
            if isinstance(m, TensorQuantizer) and isinstance(getattr(m, "_calibrator", None), HistogramCalibrator)
            and m._calibrator._counts is not None]
    if not cals:
        return
    dev = cals[0]._counts.device
    meta = torch.tensor([[c._counts.numel(), float(c._calib_bin_edges[1] - c._calib_bin_edges[0])] for c in cals],
                        dtype=torch.float64, device=dev)
    hi, lo = meta.clone(), meta.clone()
    dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
    if not torch.equal(hi[:, 1], lo[:, 1]):
        raise RuntimeError("sync_histograms: ranks disagree on the bin width (their first batches had different "
                           "ranges); use HistogramCalibrator with an agreed initial range")
    sizes = hi[:, 0].long().tolist()
    flat = torch.zeros(sum(sizes), dtype=torch.int64, device=dev)
    off = 0
    for c, n in zip(cals, sizes):
        flat[off:off + c._counts.numel()] = c._counts
        off += n
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    off = 0
    for c, n in zip(cals, sizes):
        if n != c._counts.numel():  # grow this rank's edges to the common range, same width
            width = c._calib_bin_edges[1] - c._calib_bin_edges[0]
            c._calib_bin_edges = torch.arange(0, n + 1, device=c._calib_bin_edges.device) * width
            c._num_bins = n
        c._counts = flat[off:off + n].clone()
        off += n


def disable_pre_quant_scale_and_resmooth(*a, **k):  # pragma: no cover
    raise NotImplementedError("modelopt_b200: resmoothing is not implemented")


# =================================================================================================
# MSE weight calibration (model_calib.py:639-826)
# =================================================================================================
def _mse_quant_func(x, amax, quantizer):
    """model_calib.py:639-661: fake-quantize ``x`` with a temporary amax."""
    original = quantizer._amax.clone() if hasattr(quantizer, "_amax") else None
    quantizer._amax = amax
    state = (quantizer._if_quant, quantizer._if_calib, quantizer._fake_quant)
    quantizer._if_quant, quantizer._if_calib, quantizer._fake_quant = True, False, True
    try:
        if hasattr(quantizer, "_original_shape"):
            x = quantizer._reset_to_original_shape(x)
        xq = quantizer(x)
        if hasattr(quantizer, "_block_reshape_size"):
            xq = quantizer._process_for_blockquant(xq)
    finally:
        quantizer._if_quant, quantizer._if_calib, quantizer._fake_quant = state
        if original is not None:
            quantizer._amax = original
        else:
            delattr(quantizer, "_amax")
    return xq


def _make_weight_mse_calibrator(q, step_size, start_multiplier, stop_multiplier, fp8_scale_sweep):
    """model_calib.py:664-730."""
    from functools import partial

    from .calib import MseCalibrator, NVFP4MSECalibrator

    if (not isinstance(q, TensorQuantizer) or not q.is_enabled or q._dynamic or q._calibrator is None
            or getattr(q, "_amax", None) is None):
        return None
    initial = q._amax.clone().detach()
    axis = q._calibrator._axis
    quant_func = partial(_mse_quant_func, quantizer=q)
    if fp8_scale_sweep:
        if q.is_nvfp4_static and q.backend is None and getattr(q, "global_amax", None) is not None:
            return NVFP4MSECalibrator(amax=initial, axis=axis, global_amax=q.global_amax, quant_func=quant_func)
        return None
    fused = None
    if q.backend is None and q.pre_quant_scale is None and not q.is_nvfp4_static and not q.is_nvfp4_dynamic:
        if isinstance(q._num_bits, int):
            fused = {"num_bits": q._num_bits, "unsigned": q._unsigned, "narrow_range": q._narrow_range}
        elif q._num_bits == (4, 3):
            fused = {"num_bits": (4, 3)}
    return MseCalibrator(amax=initial, axis=axis, step_size=step_size, start_multiplier=start_multiplier,
                         stop_multiplier=stop_multiplier, quant_func=quant_func, fused=fused)


@torch.no_grad()
def mse_calibrate(model: nn.Module, forward_loop=None, distributed_sync=True, step_size: float = 0.1,
                  start_multiplier: float = 0.25, stop_multiplier: float = 4.0, fp8_scale_sweep: bool = False, **_):
    """model_calib.py:733-826: max-calibrate everything, then refine every weight amax by MSE search."""
    max_calibrate(model, forward_loop, distributed_sync)
    seen = set()
    for module in model.modules():
        if id(module) in seen or not isinstance(module, QuantModule):
            continue
        seen.add(id(module))
        for weight, q in module.iter_weights_for_calibration():
            cal = _make_weight_mse_calibrator(q, step_size, start_multiplier, stop_multiplier, fp8_scale_sweep)
            if cal is None:
                continue
            old = q._calibrator
            q._calibrator = cal
            q.disable_quant()
            q.enable_calib()
            q(weight)
            if cal.compute_amax() is not None:
                q.load_calib_amax()
            q.enable_quant()
            q.disable_calib()
            cal.reset()
            q._calibrator = old


__all__ += ["mse_calibrate"]
