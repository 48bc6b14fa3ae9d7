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

import math
import warnings

import torch
import torch.distributed as dist
from torch import nn

from .arena import AmaxArenaPool
from .calib import MaxCalibrator
from .nn.quant_module import QuantModule
from .nn.tensor_quantizer import SequentialQuantizer, TensorQuantizer
from .utils import is_quantized_linear

__all__ = ["max_calibrate", "enable_stats_collection", "finish_stats_collection", "weight_only_quantize",
           "disable_pre_quant_scale_and_resmooth", "sync_histograms", "sync_max_calibrators"]


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


def enable_stats_collection(model: nn.Module, arena_pool: AmaxArenaPool | None = None, distributed_sync: bool = True):
    """model_calib.py:1128-1141 (+ attach the shared amax arena to every max calibrator, and tell every histogram
    calibrator which data-parallel group it calibrates with: see `sync_histograms`)."""
    from .calib import HistogramCalibrator

    hist_group = False
    if distributed_sync and dist.is_available() and dist.is_initialized():
        group, ok = _dp_group(model)
        if ok and dist.get_world_size(group) > 1:
            hist_group = group
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
                if isinstance(module._calibrator, HistogramCalibrator) and module._calibrator._state is None:
                    module._calibrator.dp_group = hist_group
            else:
                module.disable()


def finish_stats_collection(model: nn.Module, method: str | None = None, **kwargs):
    """model_calib.py:1144-1167 (+ the data-parallel SUM of the histograms, when the calibrators collected as a group)."""
    from .calib import HistogramCalibrator

    groups = {id(c.dp_group): c.dp_group for c in
              (getattr(m, "_calibrator", None) for m in model.modules() if isinstance(m, TensorQuantizer))
              if isinstance(c, HistogramCalibrator) and c.dp_group is not False and c._state is not None}
    for g in groups.values():
        sync_histograms(model, g)
    _batched_histogram_searches(model, method, kwargs)
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


_SEARCH_STREAMS = {}


def _batched_histogram_searches(model, method, kwargs, n_streams: int = 8):
    """The entropy searches of every HistogramCalibrator, enqueued on a few streams side by side and read back
    with ONE synchronisation; each calibrator remembers its result, so the per-quantizer loop of finish_stats_collection
    (which, like the reference's, asks every calibrator twice) finds it ready.  One search is three small latency-bound
    kernels (~0.2 ms): 224 of them one after the other, each followed by a host read, were 140 ms of a Llama-8B
    calibration's fixed cost; this is ~10 ms."""
    from .calib import HistogramCalibrator

    if method != "entropy":  # the only histogram search this entry point can request (like the reference's, :1158-1163)
        return
    m, kw = "entropy", {}
    cals = [q._calibrator for q in model.modules()
            if isinstance(q, TensorQuantizer) and not q._disabled and not getattr(q, "_dynamic", False)
            and isinstance(getattr(q, "_calibrator", None), HistogramCalibrator)
            and q._calibrator._state is not None and q._calibrator._override is None]
    if len(cals) < 2:
        return
    dev = cals[0]._state.device
    if any(c._state.device != dev for c in cals):
        return
    # one read of every calibrator's state (bin counts, frozen flags) instead of one per calibrator
    snaps = torch.stack([c._state for c in cals]).tolist()
    for c, snap in zip(cals, snaps):
        if c._pending is not None:
            c._pending.synchronize()
            c._pending = None
        c._snap, c._snap_valid = snap, True
        c._raise_if_frozen(snap)
    cur = torch.cuda.current_stream(dev)
    pool = _SEARCH_STREAMS.setdefault((dev.type, dev.index), [])
    while len(pool) < n_streams:
        pool.append(torch.cuda.Stream(device=dev))  # created once per device: cudaStreamCreate costs ~2 ms
    streams = pool[:min(n_streams, len(cals))]
    idx = []
    for i, c in enumerate(cals):
        st = streams[i % len(streams)]
        if i < len(streams):
            st.wait_stream(cur)
        with torch.cuda.stream(st):
            idx.append(c.launch_search(m, **kw))
    for st in streams:
        cur.wait_stream(st)
    for t in idx:
        t.record_stream(cur)
    ks = torch.cat([t.reshape(1).to(torch.int64) for t in idx]).tolist()
    for c, k in zip(cals, ks):
        c.finish_search(m, int(k), **kw)


def _dp_group(model):
    for m in model.modules():
        if isinstance(m, QuantModule) and m.parallel_state is not None:
            g = m.parallel_state.data_parallel_group
            return None if g == -1 else g, g != -1
    return None, True


def _max_calibrators(model):
    """(quantizer, MaxCalibrator) of every enabled quantizer that collects with a MaxCalibrator, in `named_modules` order
    (identical on every rank of a data-parallel replica)."""
    out = []
    for m in model.modules():
        if isinstance(m, TensorQuantizer) and not m._disabled and isinstance(getattr(m, "_calibrator", None), MaxCalibrator):
            out.append((m, m._calibrator))
    return out


def _state_bits(cal):
    """The int32 words holding a MaxCalibrator's running amax (arena slot or private tensor), or None before any collect."""
    if cal._shape is None:
        return None
    return cal._slot_arena.bits(cal._slot) if cal._slot is not None else cal._bits


def sync_max_calibrators(model, pool: AmaxArenaPool, group=None):
    """The data-parallel exchange of the running amax of EVERY MaxCalibrator of `model` (reference: one
    all_reduce(MAX) per quantizer, tensor_quantizer.py:1373-1385).  Fast path: ONE all-reduce over each arena of `pool` plus
    ONE over the concatenated state of the calibrators that live outside it (they collected before this pool existed:
    `enable_stats_collection` + `forward_loop` of the AWQ passes, a second `calibrate()` without `reset_amax`).  The fast path
    needs the same arena layout on every rank; slots are created at a quantizer's first non-empty collect, so a rank that
    never fed one (an empty MoE expert, an uneven forward_loop) has another layout: ranks first agree on a layout
    signature and otherwise take the ordered per-quantizer exchange, which also hands the amax to the rank that saw no
    data (the reference would hang in NCCL there).  Returns the set of calibrators whose state went through `pool`."""
    cals = _max_calibrators(model)
    arenas = pool.arenas()
    in_pool = {id(c) for _, c in cals if c._slot is not None and any(c._slot_arena is a for a in arenas)}
    strays = [c for _, c in cals if id(c) not in in_pool and _state_bits(c) is not None]
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return in_pool
    dev = next((b.device for b in (_state_bits(c) for _, c in cals) if b is not None), None)
    if dev is None:
        dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")

    def h(seq):
        v = 1469598103934665603
        for x in seq:
            v = ((v ^ (int(x) & 0xFFFFFFFF)) * 1099511628211) & 0x7FFFFFFFFFFFFFF
        return v

    sig = torch.tensor([len(arenas), sum(a._used for a in arenas), h(x for a in arenas for s in a._slots for x in s),
                        len(strays), h(_state_bits(c).numel() for c in strays),
                        h(i for i, (_, c) in enumerate(cals) if id(c) in in_pool)], dtype=torch.int64, device=dev)
    lo, hi = sig.clone(), sig.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    if torch.equal(lo, hi):
        pool.sync(group)
        if strays:
            flat = torch.cat([_state_bits(c).reshape(-1) for c in strays])
            dist.all_reduce(flat, op=dist.ReduceOp.MAX, group=group)
            off = 0
            for c in strays:
                b = _state_bits(c)
                b.copy_(flat[off:off + b.numel()].view_as(b))
                off += b.numel()
                c._checked = False
        return in_pool
    # ---- layouts differ: ordered exchange, quantizer by quantizer in one flat buffer ----
    sizes = torch.tensor([0 if _state_bits(c) is None else _state_bits(c).numel() for _, c in cals],
                         dtype=torch.int64, device=dev)
    agreed = sizes.clone()
    dist.all_reduce(agreed, op=dist.ReduceOp.MAX, group=group)
    agreed_l, sizes_l = agreed.tolist(), sizes.tolist()
    flat = torch.zeros(sum(agreed_l), dtype=torch.int32, device=dev)
    off = 0
    for (_, c), n, mine in zip(cals, agreed_l, sizes_l):
        if mine:
            if mine != n:
                raise RuntimeError("amax shape differs between data-parallel ranks")
            flat[off:off + n] = _state_bits(c).reshape(-1)
        off += n
    dist.all_reduce(flat, op=dist.ReduceOp.MAX, group=group)
    off = 0
    for (q, c), n, mine in zip(cals, agreed_l, sizes_l):
        if n and mine:
            _state_bits(c).copy_(flat[off:off + n].view_as(_state_bits(c)))
        elif n == 1:  # this rank never fed the quantizer: adopt the group's per-tensor amax
            c._bits, c._slot, c._shape = flat[off:off + 1].clone(), None, ()
            c._dtype = c._dtype or torch.float32
        elif n:
            warnings.warn("a per-channel quantizer saw no data on this rank; its amax stays unset here")
        c._checked = False
        off += n
    return set()


@torch.no_grad()
def max_calibrate(model: nn.Module, forward_loop=None, distributed_sync=True, **_unused):
    """model_calib.py:310-498 (data-parallel sync only; TP / EP groups are outside this path)."""
    pool = AmaxArenaPool()
    enable_stats_collection(model, pool)
    weight_only_quantize(model)
    if forward_loop is not None:
        forward_loop(model)
    # the one exchange: amax of every quantizer, all-reduce MAX over the data-parallel group
    validated = {id(c) for _, c in _max_calibrators(model)
                 if c._slot is not None and any(c._slot_arena is a for a in pool.arenas())}
    if distributed_sync and dist.is_available() and dist.is_initialized():
        group, ok = _dp_group(model)
        if ok:
            validated = sync_max_calibrators(model, pool, group)
    pool.check()  # one host read: AssertionError on NaN / inf like MaxCalibrator.collect (max.py:69-77)
    for _, c in _max_calibrators(model):
        if id(c) in validated:
            c._checked = True  # covered by pool.check(); every other calibrator validates itself in compute_amax()
    finish_stats_collection(model)
    if distributed_sync and dist.is_available() and dist.is_initialized():
        # quantizers whose amax did not come from a MaxCalibrator that was enabled here — custom calibrators, and DISABLED
        # quantizers that already hold an amax (AWQ's patched forwards calibrate the input quantizer per batch and leave it
        # disabled): the reference's per-quantizer sync covers every quantizer with an ``_amax``, enabled or not
        # (model_calib.py:117-130)
        group, ok = _dp_group(model)
        for m in model.modules():
            if ok and isinstance(m, TensorQuantizer) and getattr(m, "_amax", None) is not None \
                    and (m._disabled or not isinstance(getattr(m, "_calibrator", None), MaxCalibrator)):
                m.sync_amax_across_distributed_group(group)
    _finalize_with_shared_state(model)


def _finalize_with_shared_state(model: nn.Module, weight_patterns=None):
    """model_calib.py:143-157: once every ``_amax`` is final (after the data-parallel sync), static-block weight
    quantizers are promoted and fusible siblings (q/k/v, gate/up, w1/w3) get one shared NVFP4 ``global_amax``."""
    from .utils import SHARED_PATTERNS, promote_static_block_weight_quantizers

    promote_static_block_weight_quantizers(model, SHARED_PATTERNS if weight_patterns is None else weight_patterns)


def sync_histograms(model: nn.Module, group=None):
    """Merge the HistogramCalibrators of a data-parallel group: all-reduce(MAX) of the ranges, then all-reduce(SUM) of the
    counts of every calibrator in ONE flat int64 buffer.

    New behaviour (the reference does not synchronise histograms, calib/histogram.py:158-163).  Every rank derived the same
    bin WIDTH from the all-reduced maximum of the first batch (`HistogramCalibrator.collect`), and a range only ever grows
    in whole bins of that width (histogram.py:121-126), so bin i is the same interval on every rank and a rank whose range
    stayed smaller simply has zeros above its last bin.  The merge therefore is: (1) verify that the widths agree (MIN and
    MAX of the state words: one exchange each) and raise if not - e.g. calibrators that collected before `dp_group` was set;
    (2) every calibrator takes the widest rank's (bin count, range); (3) the exact integer counts are summed.  The result is
    the bin-by-bin sum of what the reference computes on each rank; with `dp_sync = "step"` the ranks already agree batch
    by batch and the sum is bit-identical to ONE process fed the union of the ranks' batches.  World size 1 is a no-op (the
    reference's behaviour bit for bit); a second call without a `collect` in between does nothing."""
    from .calib import HistogramCalibrator
    from .calib.histogram import _FIRST_AMAX, _FLAGS, _LAST_END, _NBINS, _RANGE_MAX, _WIDTH

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    cals = [m._calibrator for m in model.modules()
            if isinstance(m, TensorQuantizer) and isinstance(getattr(m, "_calibrator", None), HistogramCalibrator)
            and m._calibrator._state is not None and not m._calibrator._synced]
    n_cals = torch.tensor([len(cals)], dtype=torch.int64, device=cals[0]._state.device if cals else None)
    if cals:
        lo, hi = n_cals.clone(), n_cals.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
        if int(lo) != int(hi):
            raise RuntimeError("sync_histograms: ranks hold different numbers of histogram calibrators with data")
    if not cals:
        return
    # per calibrator: bin count, range, width, end of the last growth, first-batch maximum (fp32 bit patterns of non-negative
    # values order like the values), and the two freeze flags as separate columns (NCCL has no bitwise OR)
    st = torch.stack([c._state[:6] for c in cals]).to(torch.int64)
    meta = torch.cat([st[:, [_NBINS, _RANGE_MAX, _WIDTH, _LAST_END, _FIRST_AMAX]], st[:, _FLAGS:_FLAGS + 1] & 1,
                      (st[:, _FLAGS:_FLAGS + 1] >> 1) & 1], dim=1)
    lo, hi = meta.clone(), meta.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    if not torch.equal(lo[:, [2, 4]], hi[:, [2, 4]]):
        raise RuntimeError("sync_histograms: ranks disagree on the bin width of a histogram; calibrate with "
                           "HistogramCalibrator.dp_group set from the first batch (enable_stats_collection does it) so "
                           "that every rank derives the width from the all-reduced maximum of the first batch")
    own, merged = meta.tolist(), hi.tolist()  # two host reads for all calibrators
    sizes = [row[0] for row in merged]
    for c, mine, row, n in zip(cals, own, merged, sizes):
        if n > c._buf.numel():
            raise RuntimeError("sync_histograms: a rank's histogram has more bins than this rank's counts buffer holds "
                               "(different `headroom` / `num_bins` on the ranks)")
        if mine != row:  # another rank's range grew further, or a rank froze: take over the merged state
            st_new = c._state.tolist()
            st_new[_NBINS], st_new[_RANGE_MAX], st_new[_LAST_END], st_new[_FLAGS] = row[0], row[1], row[3], row[5] | (row[6] << 1)
            c._state.copy_(torch.tensor(st_new, dtype=torch.int32))
        c._snap_valid = False
        c._edges_cache = None
    flat = torch.cat([c._buf[:n] for c, n in zip(cals, sizes)])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    off = 0
    for c, n in zip(cals, sizes):
        c._buf[:n] = flat[off:off + n]
        c._synced = True
        off += n
        c._num_bins  # reads the merged state: a calibrator that froze on ANY rank is reported on every rank


@torch.no_grad()
def disable_pre_quant_scale_and_resmooth(linear: nn.Module, delete_pre_quant_scale: bool = False):
    """model_calib.py:1169-1204: fold the input quantizer's pre_quant_scale back into the weight (W <- dtype(W * s)),
    re-calibrate the weight quantizer and switch the scale off; the fold and the new weight amax are ONE kernel
    (mo_scale_cols_amax) for per-row / per-tensor weight quantizers."""
    from .utils import is_quantized_linear

    assert is_quantized_linear(linear), "Only quantized linear modules are supported"
    iq = linear.input_quantizer
    assert iq._enable_pre_quant_scale, "pre_quant_scale should be enabled first!"
    assert hasattr(iq, "_pre_quant_scale"), "pre_quant_scale should be available"
    _apply_weight_pre_quant_scale(linear, iq._pre_quant_scale.to(torch.float32))
    # the scale is kept (it may be useful later) but disabled
    iq._enable_pre_quant_scale = False
    if iq.amax is not None:
        assert hasattr(iq, "_amax_for_smoothing")
        w = linear._parameters["weight"]
        iq.amax = iq._amax_for_smoothing.amax().to(device=w.device, dtype=w.dtype)
    if delete_pre_quant_scale:
        delattr(iq, "_pre_quant_scale")
        iq._enable_pre_quant_scale = False


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


def _make_weight_mse_calibrator(q, step_size, start_multiplier, stop_multiplier, fp8_scale_sweep, error_func=None,
                                hessian=None):
    """model_calib.py:664-730.  ``error_func`` / ``hessian``: the local-Hessian metric (closure for the generic sweep, raw
    per-cin-block tensor for the fused NVFP4 kernel)."""
    from functools import partial

    from .calib import MseCalibrator, NVFP4MSECalibrator

    if (not isinstance(q, TensorQuantizer) or not q.is_enabled or q._dynamic or q.is_mx_format or q._calibrator is None
            or getattr(q, "_amax", None) is None):
        return None
    initial = q._amax.clone().detach()
    axis = q._calibrator._axis
    quant_func = partial(_mse_quant_func, quantizer=q)
    if fp8_scale_sweep:
        if q.is_nvfp4_static and q.backend is None and getattr(q, "global_amax", None) is not None:
            return NVFP4MSECalibrator(amax=initial, axis=axis, global_amax=q.global_amax, quant_func=quant_func,
                                      error_func=error_func, hessian=hessian)
        return None
    fused = None
    if (error_func is None and q.backend is None and q.pre_quant_scale is None and not q.is_nvfp4_static
            and not q.is_nvfp4_dynamic):
        if isinstance(q._num_bits, int):
            fused = {"num_bits": q._num_bits, "unsigned": q._unsigned, "narrow_range": q._narrow_range}
        elif q._num_bits == (4, 3):
            fused = {"num_bits": (4, 3)}
    return MseCalibrator(amax=initial, axis=axis, step_size=step_size, start_multiplier=start_multiplier,
                         stop_multiplier=stop_multiplier, quant_func=quant_func, error_func=error_func, fused=fused)


def _mse_calibrate_weights(model, step_size, start_multiplier, stop_multiplier, fp8_scale_sweep, error_func_for=None,
                           hessian_for=None):
    """model_calib.py:776-826: refine every eligible weight amax with its MSE (or Hessian-weighted) search."""
    seen = set()
    for module in model.modules():
        if id(module) in seen or not isinstance(module, QuantModule):
            continue
        seen.add(id(module))
        for weight, q in module.iter_weights_for_calibration():
            cal = _make_weight_mse_calibrator(q, step_size, start_multiplier, stop_multiplier, fp8_scale_sweep,
                                              error_func_for(q) if error_func_for else None,
                                              hessian_for(q) if hessian_for else None)
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


@torch.no_grad()
def mse_calibrate(model: nn.Module, forward_loop=None, distributed_sync=True, step_size: float = 0.1,
                  start_multiplier: float = 0.25, stop_multiplier: float = 4.0, fp8_scale_sweep: bool = False, **_):
    """model_calib.py:733-826: max-calibrate everything, then refine every weight amax by MSE search."""
    max_calibrate(model, forward_loop, distributed_sync)
    _mse_calibrate_weights(model, step_size, start_multiplier, stop_multiplier, fp8_scale_sweep)


class _LocalHessianAccumulator:
    """model_calib.py:829-900: per-cin-block H = sum X^T X for one weight quantizer.  ``accumulate`` is ONE read of the
    activation (mo_block_hessian) instead of float copy + transpose + reshape + batched matmul."""

    def __init__(self, cout: int, cin: int, block_size: int):
        self.cout, self.cin, self.block_size = cout, cin, block_size
        self.num_blocks_per_cin = cin // block_size
        self.is_enabled = cin % block_size == 0  # not block-divisible -> no Hessian (falls back to plain MSE)
        self.hessian_per_block = None
        self._normalized_hessian = None
        self.num_samples = 0

    @torch.no_grad()
    def accumulate(self, input_tensor: torch.Tensor) -> None:
        if not self.is_enabled:
            return
        from . import ops

        x = input_tensor.reshape(-1, self.cin)
        if self.hessian_per_block is None:
            self.hessian_per_block = torch.zeros(self.num_blocks_per_cin, self.block_size, self.block_size,
                                                 dtype=torch.float32, device=x.device)
        if self.block_size == 16 and x.is_cuda and x.dtype in (torch.float32, torch.float16, torch.bfloat16):
            ops.block_hessian_into(x, self.hessian_per_block, 16)
        else:  # other block sizes: the reference's batched matmul
            xt = x.to(torch.float32).T.reshape(self.num_blocks_per_cin, self.block_size, -1)
            self.hessian_per_block += xt @ xt.transpose(-1, -2)
        self.num_samples += input_tensor.numel() // self.cin

    def normalized_hessian(self):
        if self._normalized_hessian is None and self.hessian_per_block is not None and self.num_samples:
            self._normalized_hessian = self.hessian_per_block / self.num_samples
        return self._normalized_hessian

    def build_error_func(self, keep_buffer: bool = False):
        hessian = self.normalized_hessian()
        if hessian is None:
            return None
        cout, bs = self.cout, self.block_size
        if not keep_buffer:
            self.hessian_per_block = None

        def local_hessian_error(x, xq):
            shape = x.shape
            dw = (x - xq).view(cout, -1, bs)
            loss = torch.einsum("cnb,nbd,cnd->cn", dw, hessian.to(dw.dtype), dw).reshape(-1)
            return loss.unsqueeze(-1).expand(-1, bs).reshape(shape)

        return local_hessian_error


def sync_local_hessians(accumulators, group=None):
    """all-reduce(SUM) of every accumulator's raw Hessian and sample count over the data-parallel group, in ONE flat
    buffer (accumulators are visited in creation order, which is the same on every rank of a data-parallel replica)."""
    accs = [a for a in accumulators if a.is_enabled and a.hessian_per_block is not None]
    if not accs or not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    dev = accs[0].hessian_per_block.device
    flat = torch.cat([a.hessian_per_block.reshape(-1).double() for a in accs]
                     + [torch.tensor([float(a.num_samples) for a in accs], dtype=torch.float64, device=dev)])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    off = 0
    for a in accs:
        n = a.hessian_per_block.numel()
        a.hessian_per_block.copy_(flat[off:off + n].reshape(a.hessian_per_block.shape).float())
        off += n
    for i, a in enumerate(accs):
        a.num_samples = int(round(float(flat[off + i])))
        a._normalized_hessian = None


@torch.no_grad()
def local_hessian_calibrate(model: nn.Module, forward_loop=None, distributed_sync=True, step_size: float = 0.1,
                            start_multiplier: float = 0.25, stop_multiplier: float = 4.0, fp8_scale_sweep: bool = True,
                            block_size: int = 16, debug: bool = False, **_):
    """model_calib.py:1004-1125: max-calibrate, capture H = sum X^T X per weight from a forward with the weight quantizers
    disabled, then run the weight search with dw^T H dw as the metric (static NVFP4 + fp8_scale_sweep: one fused kernel)."""
    if forward_loop is None:
        warnings.warn("forward_loop must be provided for local_hessian; skipping local_hessian")
        return
    max_calibrate(model, forward_loop, distributed_sync)
    accumulators: dict[int, _LocalHessianAccumulator] = {}
    handles = []

    def dense_hook(linear, args):
        if not args:
            return
        q = linear.weight_quantizer
        w = linear._parameters["weight"]
        acc = accumulators.get(id(q))
        if acc is None:
            acc = accumulators[id(q)] = _LocalHessianAccumulator(w.shape[0], w.shape[1], block_size)
        x = args[0]
        acc.accumulate(x.to_local() if hasattr(x, "to_local") else x)

    disabled = []
    for name, module in model.named_modules():
        if is_quantized_linear(module) and isinstance(module.weight_quantizer, TensorQuantizer):
            w = module._parameters.get("weight")
            if w is None or w.dim() != 2 or not module.weight_quantizer.is_enabled:
                continue
            if w.shape[1] % block_size != 0:
                warnings.warn(f"local_hessian: {name} input features ({w.shape[1]}) not divisible by block_size "
                              f"({block_size}); falling back to plain MSE for these weights.")
            handles.append(module.register_forward_pre_hook(dense_hook))
    for module in model.modules():  # forward with weight fake-quant disabled; input quantizers untouched (:1069-1073)
        q = getattr(module, "weight_quantizer", None)
        for sub in (q if isinstance(q, SequentialQuantizer) else [q]):
            if isinstance(sub, TensorQuantizer) and sub.is_enabled:
                sub.disable()
                disabled.append(sub)
    try:
        forward_loop(model)
    finally:
        for h in handles:
            h.remove()
        for sub in disabled:
            sub.enable()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        if distributed_sync:
            # data parallel: every rank saw other batches of the SAME layers, so H = sum X^T X and the sample count add up
            # (the reference leaves this as a TODO and warns, model_calib.py:1081-1087; with one rank nothing changes)
            sync_local_hessians(list(accumulators.values()))
        else:
            warnings.warn("local_hessian: Hessian is not synced across ranks; refined weight amaxes may diverge under "
                          "tensor/data parallelism.")
    error_funcs = {qid: acc.build_error_func(keep_buffer=debug) for qid, acc in accumulators.items()}
    hessians = {qid: acc.normalized_hessian() for qid, acc in accumulators.items()}
    _mse_calibrate_weights(model, step_size, start_multiplier, stop_multiplier, fp8_scale_sweep,
                           error_func_for=lambda q: error_funcs.get(id(q)), hessian_for=lambda q: hessians.get(id(q)))
    if debug:
        model._local_hessian_accumulators = accumulators


__all__ += ["mse_calibrate", "local_hessian_calibrate", "sync_local_hessians"]


# =================================================================================================
# SmoothQuant (model_calib.py:1208-1359)
# =================================================================================================
def _recalibrate_weight(linear):
    """reset_amax + max-calibrate the weight quantizer on the current weight (model_calib.py:1222-1223)."""
    q = linear.weight_quantizer
    q.reset_amax()
    max_calibrate(linear, lambda lin: lin.weight_quantizer(lin.weight), distributed_sync=False)


@torch.no_grad()
def _apply_weight_pre_quant_scale(linear, pre_quant_scale):
    """W <- dtype(W * s[None, :]) and the weight amax of the result.  Per-row / per-tensor weight quantizers get both
    from ONE kernel (mo_scale_cols_amax); other layouts fold with that kernel and re-calibrate with mo_amax."""
    from . import ops

    w = linear.weight
    s = pre_quant_scale.to(w.device).squeeze()
    q = linear.weight_quantizer
    mode = None
    if (isinstance(q, TensorQuantizer) and q.is_enabled and q.block_sizes is None and not q._dynamic
            and isinstance(q._calibrator, MaxCalibrator) and w.is_cuda and w.dim() == 2 and w.is_contiguous()):
        mode = "row" if q.axis in (0, (0,)) else ("tensor" if q.axis is None else None)
    if not (w.is_cuda and w.dim() == 2 and w.is_contiguous()):
        w.data.copy_((w * s[None, :]).to(w.dtype))
        linear.invalidate_weight_cache()  # `.data.copy_` does not move the weight's version counter
        _recalibrate_weight(linear)
        return
    if mode is None:
        ops.scale_cols(w.data, s, out=w.data)
        linear.invalidate_weight_cache()  # raw-pointer in-place write: the weight's version counter did not move
        _recalibrate_weight(linear)
        return
    bits = torch.zeros(w.shape[0] if mode == "row" else 1, dtype=torch.int32, device=w.device)
    ops.scale_cols(w.data, s, out=w.data, amax_bits=bits, amax_mode=mode)
    amax = bits.view(torch.float32)
    amax = amax.view(-1, 1) if mode == "row" else amax.reshape(())
    q.reset_amax()
    q.amax = amax.to(w.dtype)
    linear.invalidate_weight_cache()  # raw-pointer in-place write above: the weight's version counter did not move


@torch.no_grad()
def apply_pre_quant_scale_and_smooth(linear: nn.Module, pre_quant_scale: torch.Tensor | None = None):
    """model_calib.py:1226-1270."""
    from .utils import is_quantized_linear

    assert is_quantized_linear(linear), "Only quantized linear modules are supported"
    assert linear.input_quantizer.pre_quant_scale is None, "pre_quant_scale should be None first!"
    if pre_quant_scale is None:
        pre_quant_scale = getattr(linear.input_quantizer, "_pre_quant_scale", None)
    assert pre_quant_scale is not None, "pre_quant_scale should be provided or already set"
    assert torch.all(pre_quant_scale > 0), "pre_quant_scale should be positive"
    pre_quant_scale = pre_quant_scale.to(torch.float32)
    iq = linear.input_quantizer
    iq._enable_pre_quant_scale = True
    iq.pre_quant_scale = pre_quant_scale.to(linear.weight.dtype)
    _apply_weight_pre_quant_scale(linear, 1.0 / pre_quant_scale)
    if iq.amax is not None:
        assert hasattr(iq, "_amax_for_smoothing")
        device, dtype = linear.weight.device, linear.weight.dtype
        a = iq._amax_for_smoothing.to(device=device, dtype=dtype)
        iq.amax = (a * pre_quant_scale.to(device)).amax().to(dtype)


@torch.no_grad()
def smoothquant(model: nn.Module, forward_loop=None, alpha=1.0, **_):
    """model_calib.py:1273-1359."""
    from .utils import is_quantized_linear

    assert forward_loop is not None, "forward_loop must be provided for smoothquant"
    for module in model.modules():
        if is_quantized_linear(module) and module.input_quantizer.is_enabled and module.input_quantizer.axis is None:
            module.input_quantizer.axis = -1
    max_calibrate(model, forward_loop)
    smoothed = 0
    for name, module in model.named_modules():
        if not is_quantized_linear(module):
            continue
        iq, wq = module.input_quantizer, module.weight_quantizer
        if not hasattr(iq, "_amax"):
            warnings.warn(f"{name} is not calibrated, skip smoothing")
            continue
        if iq.num_bits != 8 or wq.num_bits != 8:
            warnings.warn(f"Only int8 smoothing is supported, skip {name}")
            continue
        if iq.axis != -1:
            warnings.warn(f"Only per-channel smoothing is supported, skip {name}")
            continue
        assert iq._amax.numel() > 1, f"Error: {name} has only one channel to smooth"
        from .utils import reduce_amax

        act_amax = iq.amax.float()
        w = module.weight
        weight_scale = reduce_amax(w.detach(), axis=0, keepdims=True)  # column max of |W| (one mo_amax launch)
        device, dtype = w.device, w.dtype
        scale_a = (weight_scale.pow(1 - alpha) / act_amax.pow(alpha)).squeeze()
        iq._amax_for_smoothing = act_amax.cpu()
        iq.reset_amax()
        iq.axis = None
        iq.amax = act_amax.amax().to(dtype=dtype, device=device)
        epsilon = 1.0 / (1 << 31)
        if scale_a.min() <= epsilon:
            scale_a[(act_amax <= epsilon).reshape(-1)] = 1
        scale_a = scale_a.clamp(min=1e-4, max=1e4)
        apply_pre_quant_scale_and_smooth(module, scale_a)
        smoothed += 1
    print(f"Smoothed {smoothed} modules")


__all__ += ["smoothquant", "apply_pre_quant_scale_and_smooth"]


# =================================================================================================
# AWQ-lite (model_calib.py:1362-1721)
# =================================================================================================
def _get_awq_quantizer_block_size(tensor, quantizer):
    """model_calib.py:1943-1952."""
    if quantizer.block_sizes is None:
        return None
    if -1 in quantizer.block_sizes:
        return quantizer.block_sizes[-1]
    if 1 in quantizer.block_sizes:
        return quantizer.block_sizes[1]
    raise ValueError("AWQ requires block quantization along -1 axis")


def _awq_weight_scale(weight, block_size):
    """get_weight_scale (model_calib.py:1453-1469): ONE kernel when the block tiles the row, else the eager chain."""
    from . import ops

    per_vec = 8 if weight.dtype == torch.float32 else 16
    cols = weight.shape[-1]
    block = block_size or cols
    lanes = block // per_vec
    if (weight.is_cuda and weight.dtype in (torch.float32, torch.float16, torch.bfloat16) and cols % block == 0
            and block % per_vec == 0 and 1 <= lanes <= 32 and lanes & (lanes - 1) == 0
            and weight.data_ptr() % 32 == 0):
        return ops.awq_weight_scale(weight.detach(), block)
    org_shape, sl, w = weight.shape, None, weight
    if block_size:
        if org_shape[-1] % block_size != 0:
            sl = slice(org_shape[-1])
            w = torch.nn.functional.pad(w, (0, block_size - org_shape[-1] % block_size), "constant", 0)
            org_shape = w.shape
        w = w.contiguous().view(-1, block_size)
    wa = w.abs()
    scale = (wa / (wa.amax(dim=1, keepdim=True) + torch.finfo(w.dtype).tiny)).view(org_shape)
    if sl is not None:
        scale = scale[..., sl]
    return scale.mean(0).to(torch.float32)


def _awq_act_scale(x):
    """get_act_scale (model_calib.py:1471-1472)."""
    from . import ops

    per_vec = 8 if x.dtype == torch.float32 else 16
    if x.is_cuda and x.dtype in (torch.float32, torch.float16, torch.bfloat16) and x.shape[-1] % per_vec == 0 \
            and x.is_contiguous() and x.data_ptr() % 32 == 0:
        return ops.awq_act_scale(x.detach())
    return x.abs().contiguous().view(-1, x.shape[-1]).mean(0).to(torch.float32)


def _awq_get_scale(x_max, w_max, alpha):
    """get_scale (model_calib.py:1474-1483); [Cin] vectors, stays in torch."""
    scales = (x_max.pow(alpha) / (w_max.to(x_max.device).pow(1 - alpha) + torch.finfo(torch.float32).tiny)) \
        .clamp(min=1e-4, max=1e4).view(-1)
    return (scales / (scales.max() * scales.min()).sqrt()).view(-1)


class AWQLiteHelper:
    """Per-linear search state (model_calib.py:1414-1451).  Losses are fp64 device accumulators, one per alpha; the
    11 candidate scale vectors are formed once when the search pass starts, not once per batch."""

    cache_mode: bool = False

    def __init__(self, module, name, alpha_step):
        self.name = name
        self.act_scale = 0.0
        self.num_cache_steps = 0
        self.num_search_steps = 0
        self.block_size = _get_awq_quantizer_block_size(module.weight, module.weight_quantizer)
        self.weight_scale = _awq_weight_scale(module.weight, self.block_size)
        self.alphas = [k.item() for k in torch.arange(0, 1.0 + alpha_step, alpha_step)]
        self.loss_acc = torch.zeros(len(self.alphas), dtype=torch.float64, device=module.weight.device)
        self.loss = dict.fromkeys(self.alphas, 0.0)
        self.scales = None
        self.best_scale = None
        self.best_alpha = None
        self.is_input_quantized = module.input_quantizer.is_enabled
        self.num_tokens = 0
        self.module = module
        self.is_enabled = True

    def setup(self):
        module = self.module
        module.__dict__["_forward_no_awq"] = module.forward
        module.forward = _awq_lite_forward.__get__(module)
        if module.input_quantizer.is_enabled:
            module.input_quantizer.disable()
            if module.input_quantizer.axis not in [None, -1]:
                self.is_enabled = False
                return
            module.input_quantizer.axis = -1

    def cleanup(self):
        module = self.module
        module.__dict__.pop("_if_calib", None)
        module.__dict__.pop("forward", None)
        module.__dict__.pop("_forward_no_awq", None)


def _awq_lite_forward(self, input, *args, **kwargs):
    """The search forward (model_calib.py:1507-1560)."""
    from . import ops

    h = self.awq_lite
    self.weight_quantizer.disable()
    for q in (self.input_quantizer, self.weight_quantizer):
        if hasattr(q, "_pre_quant_scale"):
            delattr(q, "_pre_quant_scale")
    out_actual = self._forward_no_awq(input, *args, **kwargs)
    self.weight_quantizer.enable()
    if input.numel() == 0 or not h.is_enabled:
        return out_actual
    if AWQLiteHelper.cache_mode:
        h.act_scale = h.act_scale + _awq_act_scale(self.input_quantizer(input))
        h.num_cache_steps += 1
        h.num_tokens += input.numel() / input.shape[-1]
        if h.is_input_quantized:
            self.input_quantizer._calibrator.collect(input)  # running per-channel amax (one mo_amax launch)
        return out_actual
    if h.scales is None:
        h.scales = []
        for alpha in h.alphas:
            s = _awq_get_scale(h.act_scale, h.weight_scale, alpha)
            h.scales.append(((1 / s).to(self.weight.dtype), s.to(self.weight.dtype)))
    ref = out_actual[0] if isinstance(out_actual, tuple) else out_actual
    self.input_quantizer._enable_pre_quant_scale = True
    self.weight_quantizer._enable_pre_quant_scale = True
    # bf16 / f16 linears: GEMM and loss in ONE tcgen05 kernel per alpha, the candidate output is never written
    fused = (hasattr(self, "_tcgen05_ok") and self._tcgen05_ok(input, args, kwargs)
             and not self.output_quantizer.is_enabled and ref.dtype == input.dtype and not isinstance(out_actual, tuple))
    for k, (inv_s, s) in enumerate(h.scales):
        self.input_quantizer.pre_quant_scale = inv_s
        self.weight_quantizer.pre_quant_scale = s
        if fused:
            ops.linear_loss_into(self.input_quantizer(input), self.weight_quantizer(self._parameters["weight"]),
                                 self.bias, ref, h.loss_acc[k:k + 1])
            continue
        out = self._forward_no_awq(input, *args, **kwargs)
        out = out[0] if isinstance(out, tuple) else out
        if out.is_cuda and out.dtype == ref.dtype and out.dtype != torch.float64:
            ops.sq_diff_mean_into(out, ref, h.loss_acc[k:k + 1])  # one kernel, no host sync
        else:
            h.loss_acc[k] += (out - ref).float().pow(2).mean().double()
    h.num_search_steps += 1
    return out_actual


@torch.no_grad()
def awq_lite(model: nn.Module, forward_loop=None, alpha_step: float = 0.1, debug: bool = False, **_):
    """model_calib.py:1394-1721 (single-GPU / data-parallel; tensor- and expert-parallel syncs are out of scope)."""
    from .conversion import set_quantizer_by_cfg_context
    from .utils import is_quantized_linear

    if forward_loop is None:
        warnings.warn("forward_loop must be provided for awq_lite; skipping awq_lite")
        return
    helpers = []
    for name, module in model.named_modules():
        if is_quantized_linear(module) and module.weight_quantizer.is_enabled:
            module.awq_lite = AWQLiteHelper(module, name, alpha_step)
            module.awq_lite.setup()
            helpers.append((name, module))

    AWQLiteHelper.cache_mode = True
    enable_stats_collection(model)
    forward_loop(model)
    for _, module in helpers:  # input quantizers were collected directly by the patched forward
        h, iq = module.awq_lite, module.input_quantizer
        if h.is_input_quantized and h.num_cache_steps > 0 and iq._calibrator.compute_amax() is not None:
            iq.load_calib_amax()
    with set_quantizer_by_cfg_context(model, [{"quantizer_name": "*weight_quantizer", "enable": False}]):
        max_calibrate(model, lambda m: None, distributed_sync=True)
    finish_stats_collection(model)

    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    for _, module in helpers:
        h = module.awq_lite
        if h.num_cache_steps > 0:
            module.__dict__["_if_calib"] = True
            h.act_scale = h.act_scale / h.num_cache_steps
            bad = bool(torch.any(torch.isnan(h.act_scale)) or torch.any(torch.isnan(h.weight_scale)))
            if world > 1:
                flag = torch.tensor([float(bad)], device=h.act_scale.device)
                dist.all_reduce(flag, op=dist.ReduceOp.MAX)
                bad = bool(flag.item())
            if bad:
                h.is_enabled = False
            elif world > 1:
                dist.all_reduce(h.act_scale, op=dist.ReduceOp.SUM)
                h.act_scale /= world
        else:
            h.is_enabled = False

    AWQLiteHelper.cache_mode = False
    forward_loop(model)

    for name, module in helpers:
        h, iq = module.awq_lite, module.input_quantizer
        if h.num_cache_steps > 0 and h.num_search_steps == 0:
            h.is_enabled = False
            warnings.warn(f"awq_lite: Calling `forward_loop(model)` the second time did not forward data through the "
                          f"{name}. Please provide a valid `forward_loop` function that can be used to forward data "
                          "through the model many times.")
        for q in (module.weight_quantizer, iq):
            if hasattr(q, "_pre_quant_scale"):
                delattr(q, "_pre_quant_scale")
        if h.is_input_quantized:
            if iq.amax is not None:
                act_amax = iq.amax
                iq._amax_for_smoothing = act_amax.cpu()
                iq.reset_amax()
                iq.axis = None
                iq.amax = act_amax.amax()
            iq.enable()
        if h.is_enabled:
            losses = h.loss_acc.tolist()  # the one host read per linear
            h.loss = dict(zip(h.alphas, (float(torch.tensor(v, dtype=torch.float32)) for v in losses)))
            h.best_alpha = min(h.loss, key=h.loss.get)
            h.best_scale = _awq_get_scale(h.act_scale, h.weight_scale, h.best_alpha)
            apply_pre_quant_scale_and_smooth(module, 1.0 / h.best_scale)
        else:
            warnings.warn(f"awq_lite: Forcing pre_quant_scale=1 for {name} because it was not properly exercised "
                          "during calibration; quantizing with max calibration.")
            max_calibrate(module, lambda m: m.weight_quantizer(m.weight), distributed_sync=False)
            iq._enable_pre_quant_scale = True
            iq.pre_quant_scale = torch.ones(module.weight.shape[1], dtype=module.weight.dtype,
                                            device=module.weight.device)
        h.cleanup()
        if not debug:
            delattr(module, "awq_lite")


class AWQClipHelper:
    """Per-linear clip-search state (model_calib.py:1746-1785).  ``loss_acc[k]`` is the accumulated loss of clip ratio
    ``clip_ratios[k]``: one [co, n_blocks] fp32 plane per ratio (block search) or a scalar (per-tensor search)."""

    def __init__(self, module, min_clip_ratio, shrink_step):
        wq = module.weight_quantizer
        self.num_tokens = 0
        self.block_size = _get_awq_quantizer_block_size(module.weight, wq)
        wq.reset_amax()
        max_calibrate(wq, lambda q: q(module.weight), distributed_sync=False)
        self.w_amax = wq.amax.clone()
        co, ci = module.weight.shape
        self.clip_ratios = [round(float(k), 2) for k in torch.arange(min_clip_ratio, 1.0, shrink_step)] + [1.0]
        self.per_tensor = self.is_per_tensor_clip(module)
        dev = module.weight.device
        if self.per_tensor:
            self.loss_acc = torch.zeros(len(self.clip_ratios), dtype=torch.float64, device=dev)
        else:
            self.loss_acc = torch.zeros(len(self.clip_ratios), co, math.ceil(ci / self.block_size), device=dev)
        self.best_clip_val = None
        self.best_loss = None
        self.is_input_quantized = module.input_quantizer.is_enabled
        wq.disable()

    @property
    def loss(self):
        """{ratio: loss} like the reference's dict (views of ``loss_acc``)."""
        return {r: self.loss_acc[k] for k, r in enumerate(self.clip_ratios)}

    @staticmethod
    def is_per_tensor_clip(module):
        q = module.weight_quantizer
        dynamic = q.block_sizes is not None and q.block_sizes.get("type", None) == "dynamic" and q.axis is None
        return dynamic or (q.axis is None and q.block_sizes is None)

    def update_best_params(self):
        """First strict minimum over the ascending ratios (model_calib.py:1787-1798)."""
        loss = self.loss_acc.float() if self.per_tensor else self.loss_acc
        self.best_loss = torch.full_like(self.w_amax, float("inf"), dtype=torch.float32)
        self.best_clip_val = torch.zeros_like(self.w_amax)
        for k, shrink in enumerate(self.clip_ratios):
            cur = loss[k].view_as(self.w_amax)
            take = cur < self.best_loss
            self.best_loss = torch.where(take, cur, self.best_loss)
            self.best_clip_val = torch.where(take, self.w_amax * shrink, self.best_clip_val)  # product in amax's dtype


def _awq_clip_block_losses(module, h, tokens):
    """Blocked branch of `_clip_search` (model_calib.py:1813-1858)."""
    from . import ops

    weight, q = module.weight, module.weight_quantizer
    bs = h.block_size
    fused = (weight.is_cuda and weight.dtype in (torch.float32, torch.float16, torch.bfloat16)
             and tokens.dtype == weight.dtype and isinstance(q.num_bits, int) and bs <= ops.AWQ_CLIP_MAX_BLOCK
             and weight.is_contiguous())
    if fused:
        ops.awq_clip_loss_into(weight.detach(), tokens, h.w_amax.float().contiguous().view(-1), h.clip_ratios,
                               h.loss_acc, bs, num_bits=q.num_bits, unsigned=q.unsigned,
                               narrow_range=q.narrow_range, amax_dtype=h.w_amax.dtype)
        return
    # any other block format: one batched GEMM per ratio on (Wq - W), blocks as the batch dimension
    co, ci = weight.shape
    pad = (-ci) % bs
    x = torch.nn.functional.pad(tokens, (0, pad)) if pad else tokens
    xb = x.reshape(x.shape[0], -1, bs).permute(1, 2, 0).float()                     # [nb, bs, T]
    for k, shrink in enumerate(h.clip_ratios):
        q.amax = h.w_amax * shrink
        d = (q(weight) - weight).float()
        d = torch.nn.functional.pad(d, (0, pad)) if pad else d
        out = torch.bmm(d.reshape(co, -1, bs).permute(1, 0, 2), xb)                 # [nb, co, T]
        h.loss_acc[k] += out.pow(2).mean(dim=-1).t()


def _awq_clip_tensor_losses(module, h, inputs):
    """Per-tensor branch (model_calib.py:1803-1812): only the global amax is clipped (NVFP4)."""
    from . import ops

    weight, q = module.weight, module.weight_quantizer
    x2 = inputs.reshape(-1, inputs.shape[-1])
    h.num_tokens += x2.shape[0]
    fused = ops.linear_supported(x2, weight)
    out_actual = ops.linear(x2, weight) if fused else x2 @ weight.T
    original = q.amax.clone()
    for k, shrink in enumerate(h.clip_ratios):
        q.amax = original * shrink
        wq = q(weight)
        if fused:
            ops.linear_loss_into(x2, wq, None, out_actual, h.loss_acc[k:k + 1])
        else:
            h.loss_acc[k] += (x2 @ wq.T - out_actual).float().pow(2).mean().double()


def _awq_clip_forward(self, input, *args, **kwargs):
    """The patched forward (model_calib.py:1860-1885)."""
    h = self.awq_clip
    if h.is_input_quantized:
        self.input_quantizer.enable()
        max_calibrate(self.input_quantizer, lambda q: q(input), distributed_sync=False)
        self.input_quantizer.disable()
    x = self.input_quantizer(input)
    self.weight_quantizer.enable()
    if h.per_tensor:
        _awq_clip_tensor_losses(self, h, x)
    else:
        x2 = x.reshape(-1, x.shape[-1])
        tokens = x2[0::max(1, x2.shape[0] // h.max_tokens)]
        h.num_tokens += tokens.shape[0]
        world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        if world > 1:  # the reference averages every batch's loss over the data-parallel group (:1849-1852)
            before = h.loss_acc.clone()
            _awq_clip_block_losses(self, h, tokens)
            delta = h.loss_acc - before
            dist.all_reduce(delta, op=dist.ReduceOp.SUM)
            h.loss_acc = before + delta / world
        else:
            _awq_clip_block_losses(self, h, tokens)
    self.weight_quantizer.disable()
    return self._forward_no_awq(input, *args, **kwargs)


@torch.no_grad()
def awq_clip(model: nn.Module, forward_loop=None, max_co_batch_size: int = 1024, max_tokens_per_batch: int = 64,
             min_clip_ratio: float = 0.5, shrink_step: float = 0.05, debug: bool = False, **_):
    """model_calib.py:1724-1940.  ``max_co_batch_size`` is accepted for interface parity; the fused search keeps no
    [co_bsz, tokens, n_block, block] temporaries, so there is nothing to batch."""
    from .conversion import set_quantizer_by_cfg_context
    from .utils import is_quantized_linear

    assert forward_loop is not None, "forward_loop must be provided for awq_clip"
    helpers = []
    for name, module in model.named_modules():
        if is_quantized_linear(module) and module.weight_quantizer.is_enabled \
                and module.weight_quantizer.block_sizes is not None:
            module.__dict__["_forward_no_awq"] = module.forward
            module.forward = _awq_clip_forward.__get__(module)
            module.awq_clip = AWQClipHelper(module, min_clip_ratio, shrink_step)
            module.awq_clip.max_tokens = max_tokens_per_batch
            helpers.append((name, module))

    enable_stats_collection(model)
    forward_loop(model)
    with set_quantizer_by_cfg_context(model, [{"quantizer_name": "*weight_quantizer", "enable": False}]):
        max_calibrate(model, lambda m: None, distributed_sync=True)
    finish_stats_collection(model)

    for name, module in helpers:
        h = module.awq_clip
        if h.num_tokens > 0:
            h.update_best_params()
            module.weight_quantizer.amax = h.best_clip_val
            module.weight_quantizer.enable()
            if h.is_input_quantized:
                module.input_quantizer.enable()
        module.__dict__.pop("forward", None)
        module.__dict__.pop("_forward_no_awq", None)
        if not debug:
            delattr(module, "awq_clip")


@torch.no_grad()
def awq(model: nn.Module, forward_loop=None, algorithm: str = "awq_lite", **kwargs):
    """model_calib.py:1362-1391: ``awq_lite``, ``awq_clip`` or ``awq_full`` (lite, then clip)."""
    import inspect

    if algorithm not in ("awq_lite", "awq_clip", "awq_full"):
        raise ValueError(f"modelopt_b200: unknown AWQ algorithm {algorithm!r}")

    def own(fn):
        names = inspect.signature(fn).parameters
        return {k: v for k, v in kwargs.items() if k in names}

    with SequentialQuantizer.convert_to_single_quantizer(model):
        if algorithm in ("awq_full", "awq_lite"):
            awq_lite(model, forward_loop, **own(awq_lite))
        if algorithm in ("awq_full", "awq_clip"):
            awq_clip(model, forward_loop, **own(awq_clip))
    from .utils import is_quantized_linear

    for module in model.modules():
        if is_quantized_linear(module) and isinstance(module.weight_quantizer, SequentialQuantizer):
            max_calibrate(module, lambda lin: lin.weight_quantizer(lin.weight), distributed_sync=False)


def awq_full(model: nn.Module, forward_loop=None, **kwargs):
    return awq(model, forward_loop, algorithm="awq_full", **kwargs)


__all__ += ["awq_lite", "awq_clip", "awq_full", "awq", "AWQLiteHelper", "AWQClipHelper"]
