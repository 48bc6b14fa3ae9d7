"""HistogramCalibrator on the B200 path (interface of modelopt/torch/quantization/calib/histogram.py:35-343).

``collect`` keeps the reference's contract — 2048 bins over [0, max|x|] on the first batch, the range grows in
whole bin widths afterwards, old counts are added to the prefix of the new histogram — without the reference's
abs / float-copy / max / histc chain and WITHOUT a host round trip per batch (the reference syncs twice per collect,
histogram.py:79, :121-123):

* the calibrator's parameters (bin count, range, width, flags) live in a 16-word device state and the counts in a device
  buffer of fixed capacity (``headroom`` x ``num_bins`` bins); ``collect`` enqueues kernels and returns;
* steady state: ONE read of the tensor bins |x| under the current range and folds max|x| in the same pass; a one-thread
  kernel then takes the reference's ``x_max > edges[-1]`` decision on the device, and only a batch that grows the range
  is read again (taken out of the old bins, put into the new ones);
* data-parallel calibration (``dp_group`` set by ``enable_stats_collection``): the batch maximum is all-reduced(MAX)
  between those two steps — one word, stream-ordered — so every rank takes the same decisions, its bins are the other
  ranks' bins, and the SUM of the ranks' counts (``model_calib.sync_histograms``) is bit-identical to one process
  calibrating on the concatenated batches.  New behaviour: the reference does not synchronise histograms
  (calib/histogram.py:158-163); with one rank nothing changes.

Counts are exact int64 (the reference's fp32 ``torch.histc`` stops counting at 2^24 per bin); ``_calib_hist``
is exposed as fp32 like the reference's.  Bin edges are rebuilt on demand with the same ``torch.linspace`` / ``torch.arange``
calls as the reference so that they are bit-identical.  A range that would need more than ``headroom * num_bins`` bins, or
a non-finite maximum (``torch.histc`` raises there), freezes the device state and is reported by the next state read
(at the latest by ``compute_amax``).

``compute_amax`` runs the percentile / mse / entropy searches on the device (``csrc/hist_search.cu``) and looks the
amax up in the edge array exactly like the reference.  ``method="mse"`` reproduces the reference's mis-slotted
``fake_tensor_quant`` call (histogram.py:308) unless ``intended_mse=True``.
"""

from __future__ import annotations

import os
import warnings

import torch
import torch.distributed as dist

from .. import ops
from .calibrator import _Calibrator

__all__ = ["HistogramCalibrator", "calibrate_weights"]

# include/modelopt_b200.h MO_HIST_*
_NBINS, _RANGE_MAX, _WIDTH, _LAST_END, _FLAGS, _FIRST_AMAX, _BATCH_AMAX = 0, 1, 2, 3, 4, 5, 6
_STATE_WORDS = 16
_FLAG_OVERFLOW, _FLAG_NONFINITE = 1, 2


class HistogramCalibrator(_Calibrator):
    def __init__(self, num_bits=8, axis=None, unsigned=False, num_bins=2048, grow_method=None, skip_zeros=False,
                 torch_hist=True, headroom=32):
        super().__init__(num_bits, axis, unsigned)
        self._init_bins = num_bins
        self._skip_zeros = skip_zeros
        self._headroom = headroom
        self._state = None      # int32[16] device: bin count, range, width, flags (include/modelopt_b200.h MO_HIST_*)
        self._buf = None        # int64[capacity] device counts; bins beyond the current bin count are zero
        self._snap = None       # last host copy of the state: (list of 16 ints)
        self._pending = None    # event of an asynchronous state read-back into self._host
        self._host = None       # pinned int32[16]
        self._snap_valid = False  # self._snap is the CURRENT device state (nothing was enqueued since it was read)
        self._edges_cache = None
        self._override = None   # (counts, edges) set by hand (tests, external restore)
        self._synced = False    # counts already summed over the data-parallel group
        self._n_collect = 0     # collects so far (paces the state read-back, invalidates remembered search results)
        self._memo = None       # (search key, amax) of the last entropy / percentile search
        self._torch_hist = torch_hist
        # data-parallel group whose ranks calibrate together: False = none; None = the default (world) group
        self.dp_group = False
        # "finish": ranks agree on the bin width at the first batch and merge at the end; "step": they exchange the batch
        # maximum on every collect (see collect)
        self.dp_sync = os.environ.get("MODELOPT_B200_HIST_DP_SYNC", "finish")
        # "cpu": build grown bin edges with torch.arange on the CPU (what the reference computes for CPU tensors;
        # CPU arange evaluates start + i*step in fp64, CUDA arange in fp32, so edges can differ in the last bit)
        self.numerics = None
        if axis is not None:
            raise NotImplementedError("Calibrator histogram collection only supports per tensor scaling")

    # ---- device state ------------------------------------------------------------------------------------------------
    def _group_size(self):
        if self.dp_group is False or not (dist.is_available() and dist.is_initialized()):
            return 1
        return dist.get_world_size(self.dp_group)

    @torch.no_grad()
    def collect(self, x):
        if not x.is_cuda:
            raise RuntimeError("modelopt_b200 HistogramCalibrator: expected a CUDA tensor; there is no CPU fallback")
        x = x.detach()
        if x.dtype == torch.float64:
            x = x.float()
        if x.numel() == 0:
            return
        x = x.contiguous()
        self._override = None
        first = self._state is None
        if first:
            self._state = torch.zeros(_STATE_WORDS, dtype=torch.int32, device=x.device)
            self._buf = torch.zeros(self._init_bins * self._headroom, dtype=torch.int64, device=x.device)
            self._snap = None
        else:
            self._poll()  # a finished read-back of an earlier batch: learn the bin count, report a frozen state
        hint = self._hint()
        self._synced = False
        self._edges_cache = None
        self._snap_valid = False
        self._n_collect += 1
        if self._group_size() > 1 and (first or self.dp_sync == "step"):
            # Data-parallel.  The FIRST batch's maximum is all-reduced(MAX): it fixes the bin width, which must be the same on
            # every rank for the counts to be summable at the end.  dp_sync == "finish" (default): later batches are collected
            # locally - each rank grows its own range in units of the common width - and `sync_histograms` merges the ranks
            # at the end (MAX of the ranges, SUM of the counts): no collective on the per-batch path.  dp_sync == "step": the
            # batch maximum is exchanged on EVERY collect, so all ranks take the same grow / no-grow decision and the result is
            # bit-identical to one process fed the union of the ranks' batches step by step - at one blocking one-word
            # all-reduce per collect (224 per Llama-8B forward: 3-4x the histogram overhead on 8 GPUs, moving the exchange
            # to a side stream did not help: profiles/r02_bench_8gpu_hist_step.txt).
            ops.hist_collect_begin(x, self._state, self._buf, hint, self._skip_zeros, first)
            dist.all_reduce(self._state[_BATCH_AMAX:_BATCH_AMAX + 1], op=dist.ReduceOp.MAX, group=self.dp_group)
            ops.hist_collect_end(x, self._state, self._buf, hint, self._skip_zeros, first, self._init_bins, self.numerics)
        else:
            ops.hist_collect(x, self._state, self._buf, hint, self._skip_zeros, first, self._init_bins, self.numerics)
        self._read_back()

    def _read_back(self):
        # asynchronous read-back (64 bytes), looked at by a later collect: after the first batches, then every 8th
        # (not while the stream is being captured into a CUDA graph: the collect itself is capturable, the read-back is not)
        if self._pending is None and (self._n_collect <= 2 or self._n_collect % 8 == 0) \
                and not torch.cuda.is_current_stream_capturing():
            if self._host is None:
                self._host = torch.empty(_STATE_WORDS, dtype=torch.int32).pin_memory()
            self._host.copy_(self._state, non_blocking=True)
            self._pending = torch.cuda.Event()
            self._pending.record()

    def _hint(self):
        """Bins to keep in shared-memory counters: the last bin count the host has seen (any value is correct)."""
        return max(self._init_bins, self._snap[_NBINS] if self._snap else 0)

    def _poll(self):
        if self._pending is None or torch.cuda.is_current_stream_capturing():
            return
        if self._pending.query():
            self._snap = self._host.tolist()
            self._pending = None
            self._raise_if_frozen(self._snap)

    def _raise_if_frozen(self, snap):
        if snap[_FLAGS] & _FLAG_NONFINITE:  # what torch.histc raises in the reference's collect
            raise RuntimeError("torch.histc: range of [0, inf or nan] is not finite (a batch's max|x| was inf or NaN)")
        if snap[_FLAGS] & _FLAG_OVERFLOW:
            raise RuntimeError(
                f"modelopt_b200 HistogramCalibrator: the range grew beyond {self._headroom}x the first batch's "
                f"({self._init_bins * self._headroom} bins of the first batch's width); batches from that point on were "
                "not collected.  Construct the calibrator with a larger `headroom`.")

    def _read_state(self):
        """Blocking read of the device state (compute time, tests): list of 16 ints."""
        if self._state is None:
            return None
        if not self._snap_valid or self._snap is None:
            self._snap = self._state.tolist()
            self._snap_valid = True
            if self._pending is not None:
                self._pending.synchronize()  # the pinned buffer is free again
                self._pending = None
        self._raise_if_frozen(self._snap)
        return self._snap

    @staticmethod
    def _f32(bits: int) -> float:
        return float(torch.tensor([bits], dtype=torch.int32).view(torch.float32)[0])

    # ---- the reference's attributes -------------------------------------------------------------------------------------
    @property
    def _num_bins(self):
        if self._override is not None:
            return self._override[0].numel()
        snap = self._read_state()
        return self._init_bins if snap is None else snap[_NBINS]

    @_num_bins.setter
    def _num_bins(self, value):
        self._init_bins = value

    @property
    def _counts(self):
        if self._override is not None:
            return self._override[0]
        snap = self._read_state()
        return None if snap is None else self._buf[: snap[_NBINS]]

    @property
    def _calib_hist(self):
        c = self._counts
        return None if c is None else c.to(torch.float32)

    @property
    def _calib_bin_edges(self):
        if self._override is not None:
            return self._override[1]
        snap = self._read_state()
        if snap is None:
            return None
        if self._edges_cache is None:
            if snap[_LAST_END] == 0:  # never grown: histogram.py:118 (a CPU tensor in the reference too)
                self._edges_cache = torch.linspace(0, self._f32(snap[_FIRST_AMAX]), snap[_NBINS] + 1)
            else:                     # :124-126
                cpu = ops._mode(self.numerics) == ops.MO_DIV_IEEE
                self._edges_cache = torch.arange(0, self._f32(snap[_LAST_END]), self._f32(snap[_WIDTH]),
                                                 device="cpu" if cpu else self._state.device)
        return self._edges_cache

    def set_histogram(self, counts, edges):
        """Install a histogram by hand (restore, tests): later ``compute_amax`` calls use it; ``collect`` discards it."""
        self._override = (counts, edges)

    def reset(self):
        if self._pending is not None:
            self._pending.synchronize()
        self._state = self._buf = self._snap = self._pending = self._edges_cache = self._override = self._memo = None
        self._synced = self._snap_valid = False
        self._n_collect = 0

    def compute_amax(self, method: str, *, stride: int = 1, start_bin: int = 128, percentile: float = 99.99,
                     intended_mse: bool = False):
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1 and not self._synced:
            warnings.warn(
                "This histogram was not synchronized across distributed processes (the reference never does, "
                "calib/histogram.py:158-163). Use enable_stats_collection() / finish_stats_collection(), or "
                "modelopt_b200.model_calib.sync_histograms(), or the same calibration data on every rank."
            )
        if self._counts is None and self._calib_bin_edges is None:
            return None
        memo = self._memo
        if memo is not None and memo[0] == self._search_key(method, stride, start_bin, percentile):
            return memo[1].clone()  # finish_stats_collection asks twice (compute_amax, then load_calib_amax), as the reference does
        counts, edges = self._counts, self._calib_bin_edges
        if method == "entropy":
            k = int(ops.hist_entropy_index(counts, self._num_bits, self._unsigned, stride, start_bin))
            return self.finish_search(method, k, stride=stride, start_bin=start_bin, percentile=percentile)
        if method == "mse":
            if self._num_bits == (4, 3) and not intended_mse:
                # what the reference does (histogram.py:310 passes 4 arguments to a 5-argument function)
                raise TypeError("scaled_e4m3() missing 1 required positional argument: 'M'")
            idx, _ = ops.hist_mse_index(counts, edges, self._num_bits, self._unsigned, stride, start_bin,
                                        as_written=not intended_mse)
            i = start_bin + int(idx) * stride
            e = edges.to(counts.device).float()
            return ((e[1:] + e[:-1]) / 2)[i]
        if method == "percentile":
            k = int(ops.hist_percentile_index(counts, percentile))
            return self.finish_search(method, k, stride=stride, start_bin=start_bin, percentile=percentile)
        raise TypeError(f"Unknown calibration method {method}")

    # ---- the entropy / percentile search in two halves, so that a caller with many calibrators (finish_stats_collection) can
    # enqueue every search first and read all the indices back with ONE synchronisation ------------------------------------
    def _search_key(self, method, stride, start_bin, percentile):
        return (method, stride, start_bin, percentile, self._n_collect, self._synced, id(self._override))

    def launch_search(self, method: str, *, stride: int = 1, start_bin: int = 128, percentile: float = 99.99):
        """Enqueue the search on the current stream; returns a one-element device tensor (the index) or None (no data)."""
        counts = self._counts
        if counts is None:
            return None
        if method == "entropy":
            return ops.hist_entropy_index(counts, self._num_bits, self._unsigned, stride, start_bin)
        if method == "percentile":
            return ops.hist_percentile_index(counts, percentile)
        raise TypeError(f"launch_search: method {method}")

    def finish_search(self, method: str, k: int, *, stride: int = 1, start_bin: int = 128, percentile: float = 99.99):
        """Index -> amax (what compute_amax returns), remembered until the next collect."""
        edges = self._calib_bin_edges
        amax = torch.tensor(edges[start_bin + k * stride if method == "entropy" else k].item())
        self._memo = (self._search_key(method, stride, start_bin, percentile), amax)
        return amax.clone()

    def __str__(self):
        if self._calib_bin_edges is None:
            return "HistogramCalibrator(calib_bin_edges=None)"
        e = self._calib_bin_edges
        return f"HistogramCalibrator(calib_bin_edges=[{e[0]:.3f}, ..., {e[-1]:.3f}]({len(e)}))"

    def __repr__(self):
        return (f"HistogramCalibrator({super().__repr__()} calib_bin_edges={self._calib_bin_edges} "
                f"calib_hist={self._calib_hist})")


def calibrate_weights(model, method="percentile", perchannel=True, percentile=99.99, num_bins=2048):
    """histogram.py:346-433: calibrate the weight quantizer of every quantized conv / linear from its weight.
    ``max`` is per-channel max|w| (axis 0); ``mse`` / ``entropy`` / ``percentile`` build a histogram per channel."""
    from ..nn.quant_module import QuantLinearConvBase

    for name, module in model.named_modules():
        if not isinstance(module, QuantLinearConvBase) or not hasattr(module, "weight_quantizer"):
            continue
        w = module.weight if not hasattr(module, "_parameters") else module._parameters["weight"]
        w = w.detach()
        q = module.weight_quantizer
        if method == "max":
            red = list(range(1, w.dim())) if perchannel else None
            q.amax = ops.reduce_amax(w, axis=red)
            continue
        chans = range(w.shape[0]) if perchannel else [slice(None)]
        amaxs = []
        for c in chans:
            cal = HistogramCalibrator(q.num_bits, None, q.unsigned, num_bins=num_bins, headroom=1)  # one batch: no growth
            cal.collect(w[c].contiguous())
            amaxs.append(cal.compute_amax(method, percentile=percentile) if method == "percentile"
                         else cal.compute_amax(method))
        if perchannel:
            q.amax = torch.stack([a.to(w.device) for a in amaxs]).reshape(-1, *([1] * (w.dim() - 1)))
        else:
            q.amax = amaxs[0].to(w.device)
