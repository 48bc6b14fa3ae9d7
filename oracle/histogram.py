"""Oracle: HistogramCalibrator and its amax searches (CPU numpy/torch).  TEST INFRASTRUCTURE ONLY.

Restates modelopt/torch/quantization/calib/histogram.py:
* ``HistogramCalibrator.collect``   :77-130  (the ``torch_hist=True`` branch; ``torch.histc`` is torch's own
  arithmetic and is called, not restated)
* ``_compute_amax_entropy``         :210-282
* ``_compute_amax_mse``             :285-325 — INCLUDING the reference's argument-slot bug:
  ``fake_tensor_quant(centers, amax, num_bits, unsigned)`` binds ``bias=num_bits`` and ``num_bits=unsigned``, so
  what is evaluated is ``_tensor_quant(centers - num_bits, amax, num_bits=int(unsigned)) + num_bits``
  (tensor_quant.py:350-393, :607-645).  ``intended=True`` evaluates what the code meant.
* ``_compute_amax_percentile``      :328-343

Parity: pinned against the real reference by tests/golden/hist_cpu.npz (oracle/gen_golden.py).
"""

from __future__ import annotations

import numpy as np
import torch

from .fake_quant import fp8_qdq


class HistogramCalibrator:
    """calib/histogram.py:35-208 (per-tensor only)."""

    def __init__(self, num_bits=8, axis=None, unsigned=False, num_bins=2048, skip_zeros=False):
        if axis is not None:
            raise NotImplementedError("Calibrator histogram collection only supports per tensor scaling")
        self._num_bits, self._unsigned = num_bits, unsigned
        self._num_bins, self._skip_zeros = num_bins, skip_zeros
        self._calib_bin_edges = None
        self._calib_hist = None

    def collect(self, x: torch.Tensor):
        if torch.min(x) < 0.0:
            x = x.abs()
        x = x.float()
        if self._skip_zeros:
            x = x[torch.where(x != 0)]
        x_max = x.max()
        if self._calib_bin_edges is None and self._calib_hist is None:
            self._calib_hist = torch.histc(x, bins=self._num_bins, min=0, max=x_max)
            self._calib_bin_edges = torch.linspace(0, x_max, self._num_bins + 1)
        else:
            if x_max > self._calib_bin_edges[-1]:
                width = self._calib_bin_edges[1] - self._calib_bin_edges[0]
                self._num_bins = int((x_max / width).ceil().item())
                self._calib_bin_edges = torch.arange(0, x_max + width, width, device=x.device)
            hist = torch.histc(x, bins=self._num_bins, min=0, max=self._calib_bin_edges[-1])
            hist[: self._calib_hist.numel()] += self._calib_hist
            self._calib_hist = hist

    def reset(self):
        self._calib_bin_edges = None
        self._calib_hist = None

    def compute_amax(self, method: str, *, stride=1, start_bin=128, percentile=99.99, intended_mse=False):
        hist = self._calib_hist.int().cpu().numpy()
        edges = self._calib_bin_edges.cpu().numpy()
        if method == "entropy":
            return compute_amax_entropy(hist, edges, self._num_bits, self._unsigned, stride, start_bin)
        if method == "mse":
            return compute_amax_mse(hist, edges, self._num_bits, self._unsigned, stride, start_bin, intended_mse)
        if method == "percentile":
            return compute_amax_percentile(hist, edges, percentile)
        raise TypeError(f"Unknown calibration method {method}")


def _entropy(pk: np.ndarray, qk: np.ndarray) -> float:
    """scipy.stats.entropy(pk, qk) (scipy 1.x, base e): both normalised to sum 1, sum(rel_entr(pk, qk))."""
    pk = np.asarray(pk, dtype=np.float64)
    qk = np.asarray(qk, dtype=np.float64)
    pk = pk / np.sum(pk)
    qk = qk / np.sum(qk)
    with np.errstate(divide="ignore", invalid="ignore"):
        t = np.where((pk > 0) & (qk > 0), pk * np.log(pk / qk), np.where((pk == 0) & (qk >= 0), 0.0, np.inf))
    return float(np.sum(t))


def compute_amax_entropy(calib_hist, calib_bin_edges, num_bits, unsigned, stride=1, start_bin=128):
    """histogram.py:210-282, vectorised per candidate.  For a candidate of ``i`` source bins the first ``i`` bins
    are merged into ``nbins = 2^(bits-1+unsigned)`` equal-width groups; every NON-EMPTY source bin of a group gets
    the group's mean over its non-empty bins; KL(reference || that) with the clipped tail folded into bin i-1.
    Quirks kept: ``hist[0] = hist[1]`` is written into the caller's array (``calib_hist[:]`` is a numpy view), the
    two distributions reach ``entropy`` un-normalised (the reference's normaliser discards its result; scipy
    normalises internally) and ties go to the LAST minimum."""
    if calib_bin_edges is None and calib_hist is None:
        return None
    try:
        from scipy.stats import entropy
    except ImportError:  # pragma: no cover
        entropy = _entropy
    hist = calib_hist
    hist[0] = hist[1]
    total = np.sum(hist)
    nbins = 1 << (num_bits - 1 + int(unsigned))
    candidates = list(range(start_bin, len(hist) + 1, stride))
    kl = np.empty(len(candidates), dtype=np.float64)
    for n, i in enumerate(candidates):
        head = hist[:i]
        group = np.digitize(np.arange(i), np.linspace(0, i, num=nbins + 1)) - 1
        live = head != 0
        sums = np.bincount(group[live], weights=head[live].astype(np.float64), minlength=nbins)
        cnts = np.bincount(group[live], minlength=nbins)
        mean = np.divide(sums, cnts, out=np.zeros(nbins, dtype=np.float64), where=cnts > 0)
        q = np.where(live, mean[group], 0.0)
        p = np.array(head)
        p[-1] += np.sum(hist[i:])
        if round(np.sum(q) + np.sum(hist[i:])) != total or round(np.sum(p)) != total:
            raise RuntimeError("Count mismatch!")
        kl[n] = entropy(p, q)
    best = len(kl) - 1 - int(np.argmin(kl[::-1]))
    return torch.tensor(calib_bin_edges[best * stride + start_bin].item())


def _tensor_quant_legacy(inputs, amax, num_bits, unsigned, narrow_range=True):
    """tensor_quant.py:607-645 for a scalar amax > 2^-24 (the only case the histogram search reaches)."""
    max_bound = torch.tensor((2.0 ** (num_bits - 1 + int(unsigned))) - 1.0)
    min_bound = 0 if unsigned else (-max_bound if narrow_range else -max_bound - 1)
    scale = max_bound / amax.float()
    return torch.clamp((inputs.float() * scale).round_(), min_bound, max_bound) / scale


def compute_amax_mse(calib_hist, calib_bin_edges, num_bits, unsigned, stride=1, start_bin=128, intended=False):
    """histogram.py:285-325."""
    if calib_bin_edges is None and calib_hist is None:
        return None
    counts = torch.from_numpy(calib_hist[:]).float()
    edges = torch.from_numpy(calib_bin_edges[:]).float()
    centers = (edges[1:] + edges[:-1]) / 2
    mses, arguments = [], []
    for i in range(start_bin, len(centers), stride):
        amax = centers[i]
        if isinstance(num_bits, int) and num_bits >= 0:
            if intended:
                q = _tensor_quant_legacy(centers, amax, num_bits, unsigned)
            else:  # the call as written: bias <- num_bits, num_bits <- unsigned
                q = _tensor_quant_legacy(centers - num_bits, amax, int(unsigned), False) + num_bits
        elif num_bits == (4, 3):
            if not intended:
                raise TypeError("scaled_e4m3() missing 1 required positional argument: 'M'")
            q = fp8_qdq(centers, amax)
        else:
            raise TypeError("Invalid num_bits. num_bits must be a positive integer or tuple (4,3).")
        mses.append(((q - centers) ** 2 * counts).mean())
        arguments.append(i)
    argmin = np.argmin(mses)
    return centers[arguments[argmin]]


def compute_amax_percentile(calib_hist, calib_bin_edges, percentile):
    """histogram.py:328-343."""
    if percentile < 0 or percentile > 100:
        raise ValueError("Invalid percentile. Must be in range 0 <= percentile <= 100.")
    if calib_bin_edges is None and calib_hist is None:
        return None
    total = calib_hist.sum()
    cdf = np.cumsum(calib_hist / total)
    idx = np.searchsorted(cdf, percentile / 100)
    return torch.tensor(calib_bin_edges[idx].item())
