"""HistogramCalibrator on the B200 path (interface of modelopt/torch/quantization/calib/histogram.py:35-343).

``collect`` keeps the reference's contract — 2048 bins over [0, max|x|] on the first batch, the range grows in
whole bin widths afterwards, old counts are added to the prefix of the new histogram — but does it in one or
two reads of the bf16 tensor instead of the reference's abs / float-copy / max / histc chain:

* steady state: ONE kernel (``mo_hist``) adds the bins of |x| under the current range to the running counts and
  folds max|x| in the same pass (plus a 4-byte memset and one 4-byte read); only if that max exceeded the range (the
  reference's ``x_max > edges[-1]``, a host decision there too) is the batch taken out again and re-binned with the
  grown range;
* first batch: an amax pass fixes the range, then one binning pass.

Counts are exact int64 (the reference's fp32 ``torch.histc`` stops counting at 2^24 per bin); ``_calib_hist``
is exposed as fp32 like the reference's.  Bin edges are built with the same ``torch.linspace`` / ``torch.arange``
calls as the reference so that they are bit-identical.

``compute_amax`` runs the percentile / mse / entropy searches on the device (``csrc/hist_search.cu``) and looks the
amax up in the edge array exactly like the reference.  ``method="mse"`` reproduces the reference's mis-slotted
``fake_tensor_quant`` call (histogram.py:308) unless ``intended_mse=True``.
"""

from __future__ import annotations

import warnings

import torch
import torch.distributed as dist

from .. import ops
from .calibrator import _Calibrator

__all__ = ["HistogramCalibrator", "calibrate_weights"]


class HistogramCalibrator(_Calibrator):
    def __init__(self, num_bits=8, axis=None, unsigned=False, num_bins=2048, grow_method=None, skip_zeros=False,
                 torch_hist=True):
        super().__init__(num_bits, axis, unsigned)
        self._num_bins = num_bins
        self._skip_zeros = skip_zeros
        self._calib_bin_edges = None
        self._counts = None  # int64 device tensor
        self._range_cache = None  # (edges object, device range_max, its bit pattern, amax scratch word)
        self._torch_hist = torch_hist
        # "cpu": build grown bin edges with torch.arange on the CPU (what the reference computes for CPU tensors;
        # CPU arange evaluates start + i*step in fp64, CUDA arange in fp32, so edges can differ in the last bit)
        self.numerics = None
        if axis is not None:
            raise NotImplementedError("Calibrator histogram collection only supports per tensor scaling")

    # the reference's attribute: fp32 counts
    @property
    def _calib_hist(self):
        return None if self._counts is None else self._counts.to(torch.float32)

    @torch.no_grad()
    def collect(self, x):
        if not x.is_cuda:
            raise RuntimeError("modelopt_b200 HistogramCalibrator: expected a CUDA tensor; there is no CPU fallback")
        x = x.detach()
        if x.dtype == torch.float64:
            x = x.float()
        dev = x.device
        if self._counts is None:
            amax_bits = torch.zeros(1, dtype=torch.int32, device=dev)
            ops.amax_into(x, 1, 1, x.numel(), amax_bits)
            x_max = amax_bits.view(torch.float32)
            self._check_finite(x_max)
            counts = torch.zeros(self._num_bins, dtype=torch.int64, device=dev)
            ops.hist_into(x, x_max, counts, None, self._skip_zeros)
            self._counts = counts
            self._calib_bin_edges = torch.linspace(0, x_max[0], self._num_bins + 1)  # histogram.py:118 (CPU tensor)
            return
        edges = self._calib_bin_edges
        cache = self._range_cache
        if cache is None or cache[0] is not edges or cache[1].device != dev:
            range_max = edges[-1:].to(device=dev, dtype=torch.float32)
            cache = self._range_cache = (edges, range_max, int(range_max.view(torch.int32).item()),
                                         torch.zeros(1, dtype=torch.int32, device=dev))
        _, range_max, range_bits, amax_bits = cache
        # Steady state: ONE kernel adds this batch straight into the running counts with the current range and folds
        # max|x| in the same read; the host then looks at one word (the reference's `x_max > edges[-1]` branch, :121,
        # is a host decision too).  fp32 bit patterns of non-negative values order like the values, NaN above +inf.
        amax_bits.zero_()
        ops.hist_into(x, range_max, self._counts, amax_bits, self._skip_zeros)
        if int(amax_bits.item()) > range_bits:
            x_max = amax_bits.view(torch.float32)
            # the range grows (rare): take the speculative batch out again, then re-bin it with the new range
            undo = torch.zeros_like(self._counts)
            ops.hist_into(x, range_max, undo, None, self._skip_zeros)
            self._counts -= undo
            self._check_finite(x_max)
            width = edges[1] - edges[0]
            xm = x_max[0].to(width.device)
            self._num_bins = int((xm / width).ceil().item())
            cpu = ops._mode(self.numerics) == ops.MO_DIV_IEEE
            self._calib_bin_edges = torch.arange(0, xm.cpu() + width.cpu() if cpu else xm + width,
                                                 width.cpu() if cpu else width,
                                                 device="cpu" if cpu else dev)  # :124-126
            range_max = self._calib_bin_edges[-1:].to(dev).contiguous()
            new = torch.zeros(self._num_bins, dtype=torch.int64, device=dev)
            ops.hist_into(x, range_max, new, None, self._skip_zeros)
            new[: self._counts.numel()] += self._counts
            self._counts = new

    @staticmethod
    def _check_finite(x_max):
        v = float(x_max[0])
        if v != v or v in (float("inf"), float("-inf")):
            raise RuntimeError(f"torch.histc: range of [0, {v}] is not finite")

    def reset(self):
        self._calib_bin_edges = None
        self._counts = None
        self._range_cache = None

    def compute_amax(self, method: str, *, stride: int = 1, start_bin: int = 128, percentile: float = 99.99,
                     intended_mse: bool = False):
        if dist.is_initialized():
            warnings.warn(
                "This method does not perform any synchronization across distributed processes. Use "
                "modelopt_b200.model_calib.sync_histograms() or the same calibration data on every rank."
            )
        if self._counts is None and self._calib_bin_edges is None:
            return None
        counts, edges = self._counts, self._calib_bin_edges
        if method == "entropy":
            k = int(ops.hist_entropy_index(counts, self._num_bits, self._unsigned, stride, start_bin))
            return torch.tensor(edges[start_bin + k * stride].item())
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
            return torch.tensor(edges[k].item())
        raise TypeError(f"Unknown calibration method {method}")

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
            cal = HistogramCalibrator(q.num_bits, None, q.unsigned, num_bins=num_bins)
            cal.collect(w[c].contiguous())
            amaxs.append(cal.compute_amax(method, percentile=percentile) if method == "percentile"
                         else cal.compute_amax(method))
        if perchannel:
            q.amax = torch.stack([a.to(w.device) for a in amaxs]).reshape(-1, *([1] * (w.dim() - 1)))
        else:
            q.amax = amaxs[0].to(w.device)
