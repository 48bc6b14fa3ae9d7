"""Tensor-level wrappers over the C ABI: validate, pick pointers/sizes/stream, call.  torch is only
plumbing here (device memory, current stream); all arithmetic happens in libmodelopt_b200.so."""

from __future__ import annotations

import ctypes as C
import math
import os

import torch

from . import _lib
from ._lib import MO_BF16, MO_DIV_IEEE, MO_DIV_TRITON, MO_F16, MO_F32, check

_DT = {torch.float32: MO_F32, torch.float16: MO_F16, torch.bfloat16: MO_BF16}

# kernels enqueued through the C ABI since import (bench.py reports the delta as `gpu_launches`)
LAUNCHES = [0]


def _launched(rc, what, n=1):
    check(rc, what)
    LAUNCHES[0] += n


_CPU_NAMES = ("ieee", "cpu")


def default_numerics() -> int:
    """Numerics mode (include/modelopt_b200.h MO_NUMERICS_*): 'gpu' / 'triton' = what the reference executes
    on a GPU (default), 'cpu' / 'ieee' = what it executes on CPU tensors (the oracle)."""
    return MO_DIV_IEEE if os.environ.get("MODELOPT_B200_NUMERICS", "gpu").lower() in _CPU_NAMES else MO_DIV_TRITON


_OVERRIDE: list = []


class numerics:
    """Context manager: run the enclosed calls in 'gpu' or 'cpu' numerics (see include/modelopt_b200.h)."""

    def __init__(self, mode):
        self.mode = mode

    def __enter__(self):
        _OVERRIDE.append(self.mode)

    def __exit__(self, *a):
        _OVERRIDE.pop()


def div_scalar(t: torch.Tensor, c: float) -> torch.Tensor:
    """``t / c`` as the reference's torch op computes it: CUDA tensors multiply by the fp32 reciprocal of a Python
    scalar (GPU numerics); CPU tensors divide (CPU numerics: reproduced here with a tensor / tensor division)."""
    if _mode(None) == MO_DIV_IEEE:
        return t / torch.tensor(c, dtype=t.dtype, device=t.device)
    return t / c


def _mode(numerics) -> int:
    if numerics is None and _OVERRIDE:
        numerics = _OVERRIDE[-1]
    if numerics is None:
        return default_numerics()
    if isinstance(numerics, str):
        return MO_DIV_IEEE if numerics.lower() in _CPU_NAMES else MO_DIV_TRITON
    return int(numerics)


def _code(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError(f"modelopt_b200: unsupported dtype {t.dtype} (float32/float16/bfloat16 only)") from None


def _cuda(t: torch.Tensor, what: str):
    if not t.is_cuda:
        raise RuntimeError(
            f"modelopt_b200.{what}: expected a CUDA tensor, got {t.device}. This package is the B200 CUDA path "
            "only; there is no CPU fallback."
        )


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream(t: torch.Tensor) -> int:
    """cudaStream_t of torch's current stream on t's device (the raw-handle query is ~10 us cheaper per call than building
    a torch.cuda.Stream object)."""
    if _raw_stream is not None and t.device.index is not None:
        return _raw_stream(t.device.index)
    return torch.cuda.current_stream(t.device).cuda_stream


class _dev:
    """Make t's device current for the duration of the call (cheap no-op when it already is)."""

    def __init__(self, t):
        self.idx = t.device.index
        self.prev = None

    def __enter__(self):
        cur = torch.cuda.current_device()
        if self.idx is not None and self.idx != cur:
            self.prev = cur
            torch.cuda.set_device(self.idx)

    def __exit__(self, *a):
        if self.prev is not None:
            torch.cuda.set_device(self.prev)


# MODELOPT_B200_NO_CACHE=1 turns off every cache that is keyed on a tensor's (address, version counter): the quantizer's
# fp32 amax copy and the modules' fake-quantized / packed weight caches.  A write through `.data` or a raw pointer does
# not move the version counter, so code that mutates amax or weights that way needs this switch (or
# `invalidate_weight_cache()`); the reference re-quantizes on every forward.
NO_CACHE = os.environ.get("MODELOPT_B200_NO_CACHE", "0") not in ("", "0")


def _f32_vec(a: torch.Tensor, like: torch.Tensor) -> torch.Tensor:
    """amax -> flat fp32 on like.device (reference: amax.to(float), tensor_quant_gpu.cu:83)."""
    a = a.detach()
    if a.device != like.device:
        a = a.to(like.device)
    if a.dtype != torch.float32:
        a = a.to(torch.float32)
    return a.contiguous().reshape(-1)


def _axis_of(x: torch.Tensor, amax: torch.Tensor):
    """(axis_size, inner) for a scalar or single-axis amax (tensor_quant.py:105-111); else ValueError."""
    n = amax.numel()
    if n == 1:
        return 1, x.numel()
    shape = list(amax.shape)
    if amax.dim() == x.dim() and shape.count(1) == x.dim() - 1:
        axis = shape.index(n)
    elif amax.dim() == 1:
        axis = list(x.shape).index(n)  # 1-D amax: first dim of that size (reference squeezes first)
    else:
        raise ValueError(f"amax of shape {tuple(amax.shape)} is not a scalar or a single-axis vector")
    if x.shape[axis] != n:
        raise ValueError(f"amax has {n} elements but input dim {axis} has {x.shape[axis]}")
    return n, math.prod(x.shape[axis + 1:])


def _via_f32(fn, x, *a, **k):
    """float64 input: the reference kernels compute on (float)x and cast back (tensor_quant_gpu.cu:72)."""
    return fn(x.to(torch.float32), *a, **k).to(torch.float64)


# ------------------------------------------------------------------------------------------------
def _prescale_t(prescale, x):
    """pre_quant_scale as a flat contiguous vector in x.dtype on x.device (no copy when it already is one)."""
    p = prescale.detach()
    if p.device != x.device or p.dtype != x.dtype:
        p = p.to(device=x.device, dtype=x.dtype)
    return p.contiguous().reshape(-1)


def _prescale_vec(prescale, x, f32=False):
    """(pre_quant_scale vector, cols): in x.dtype for the streaming kernels (their ABI takes the tensor's dtype), as the
    fp32 copy of its x.dtype values with f32=True (the GEMM prologue); None passes through."""
    if prescale is None:
        return None, 1
    cols = prescale.numel()
    if cols == 0 or x.numel() % cols != 0 or (x.shape[-1] % cols != 0 and cols % x.shape[-1] != 0):
        raise ValueError(f"prescale has {cols} elements, which does not tile the rows of a {tuple(x.shape)} tensor")
    return (_f32_vec(prescale.to(x.dtype), x) if f32 else _prescale_t(prescale, x)), cols


def qdq_int(x, amax, num_bits=8, unsigned=False, narrow_range=True, out=None, status=None, prescale=None):
    """`prescale` (optional, x.shape[-1] values): quantize (x * prescale) rounded to x.dtype, in the same pass."""
    _cuda(x, "qdq_int")
    if x.dtype == torch.float64:
        return _via_f32(qdq_int, x, amax, num_bits, unsigned, narrow_range, None, status, prescale)
    x = x.contiguous()
    axis_size, inner = _axis_of(x, amax)
    am = _f32_vec(amax, x)
    ps, cols = _prescale_vec(prescale, x)
    y = torch.empty_like(x) if out is None else out
    with _dev(x):
        _launched(_lib.load().mo_qdq_int_pqs(x.data_ptr(), y.data_ptr(), x.numel(), _code(x), am.data_ptr(), axis_size,
                                             inner, int(num_bits), int(bool(unsigned)), int(bool(narrow_range)),
                                             status.data_ptr() if status is not None else None,
                                             ps.data_ptr() if ps is not None else None, cols, _stream(x)),
                  "mo_qdq_int")
    return y


def qdq_fp8(x, amax=None, out=None, numerics=None, prescale=None):
    _cuda(x, "qdq_fp8")
    if x.dtype == torch.float64:
        return _via_f32(qdq_fp8, x, amax, None, numerics, prescale)
    x = x.contiguous()
    ps, cols = _prescale_vec(prescale, x)
    y = torch.empty_like(x) if out is None else out
    if amax is None:
        axis_size, inner, ap, am = 1, x.numel(), None, None
    else:
        axis_size, inner = _axis_of(x, amax)
        am = _f32_vec(amax, x)
        ap = am.data_ptr()
    with _dev(x):
        _launched(_lib.load().mo_qdq_fp8_pqs(x.data_ptr(), y.data_ptr(), x.numel(), _code(x), ap, axis_size, inner,
                                             _mode(numerics), ps.data_ptr() if ps is not None else None, cols,
                                             _stream(x)), "mo_qdq_fp8")
    return y


def qdq_nvfp4_dynamic(x, global_amax, block=16, numerics=None, out=None):
    _cuda(x, "qdq_nvfp4_dynamic")
    if x.numel() == 0:
        return x.clone()
    x = x.contiguous()
    cols = x.shape[-1]
    rows = x.numel() // cols
    ga = _f32_vec(global_amax, x)
    if ga.numel() != 1:
        ga = ga.amax().reshape(1)  # tensor_quant.py:173-174
    y = torch.empty_like(x) if out is None else out
    with _dev(x):
        _launched(_lib.load().mo_qdq_nvfp4_dynamic(x.data_ptr(), y.data_ptr(), rows, cols, _code(x), ga.data_ptr(),
                                               int(block), _mode(numerics), _stream(x)), "mo_qdq_nvfp4_dynamic")
    return y


def fp4_scales(amax, global_amax=None, quantize_block_scales=True, fp8_max_for_normalization=448.0, numerics=None):
    _cuda(amax, "fp4_scales")
    am = _f32_vec(amax, amax)
    scale = torch.empty_like(am)
    ga = _f32_vec(global_amax, amax) if global_amax is not None else None
    scratch = torch.empty(1, dtype=torch.int32, device=amax.device) if (ga is None and quantize_block_scales) else None
    with _dev(amax):
        _launched(_lib.load().mo_fp4_scales(am.data_ptr(), scale.data_ptr(), am.numel(),
                                        ga.data_ptr() if ga is not None else None, int(bool(quantize_block_scales)),
                                        float(fp8_max_for_normalization), _mode(numerics),
                                        scratch.data_ptr() if scratch is not None else None, _stream(amax)),
              "mo_fp4_scales")
    return scale.reshape(amax.shape)


def qdq_nvfp4_static(x, amax, global_amax=None, quantize_block_scales=True, fp8_max_for_normalization=448.0,
                     out_dtype=None, numerics=None):
    _cuda(x, "qdq_nvfp4_static")
    nb = amax.numel()
    if nb == 0 or x.numel() % nb != 0:
        raise ValueError(
            f"x.numel() ({x.numel()}) is not divisible by amax.numel() ({nb}); "
            "they must satisfy x.numel() == NUM_FP4_BLOCKS * BLOCK_SIZE."
        )
    block = x.numel() // nb
    out_dtype = out_dtype or x.dtype
    x = x.contiguous()
    y = torch.empty(x.shape, dtype=out_dtype, device=x.device)
    if (out_dtype == x.dtype and block in (16, 32, 64) and x.data_ptr() % 32 == 0
            and (global_amax is not None or not quantize_block_scales)):
        # one launch: the block scale is derived from the block amax inside the QDQ kernel
        am = _f32_vec(amax, x)
        ga = _f32_vec(global_amax, x) if global_amax is not None else None
        with _dev(x):
            rc = _lib.load().mo_qdq_nvfp4_static_amax(x.data_ptr(), y.data_ptr(), nb, block, _code(x), am.data_ptr(),
                                                      ga.data_ptr() if ga is not None else None,
                                                      int(bool(quantize_block_scales)), float(fp8_max_for_normalization),
                                                      _mode(numerics), _stream(x))
        if rc != _lib.MO_ERR_ALIGN:
            _launched(rc, "mo_qdq_nvfp4_static_amax")
            return y
    scale = fp4_scales(amax, global_amax, quantize_block_scales, fp8_max_for_normalization, numerics)
    with _dev(x):
        _launched(_lib.load().mo_qdq_nvfp4_static(x.data_ptr(), y.data_ptr(), nb, block, _code(x), _code(y),
                                              scale.data_ptr(), _mode(numerics), _stream(x)), "mo_qdq_nvfp4_static")
    return y


def fp4_cast(x, out_dtype=None):
    _cuda(x, "fp4_cast")
    x = x.contiguous()
    y = torch.empty(x.shape, dtype=out_dtype or x.dtype, device=x.device)
    with _dev(x):
        _launched(_lib.load().mo_fp4_cast(x.data_ptr(), y.data_ptr(), x.numel(), _code(x), _code(y), _stream(x)),
              "mo_fp4_cast")
    return y


# ------------------------------------------------------------------------------------------------
def amax_into(x, outer, axis_size, inner, bits):
    """Merge max|x| over the [outer, axis_size, inner] view into `bits` (int32 view of fp32 amax)."""
    _cuda(x, "amax_into")
    assert bits.numel() == axis_size and bits.element_size() == 4 and bits.is_contiguous()
    x = x.contiguous()
    with _dev(x):
        _launched(_lib.load().mo_amax(x.data_ptr(), outer, axis_size, inner, _code(x), bits.data_ptr(), _stream(x)),
              "mo_amax")


def _collapse_keep_axes(shape, keep):
    """View a contiguous tensor as [outer, axis, inner] keeping the (contiguous run of) dims in `keep`."""
    keep = sorted(keep)
    if not keep:
        return 1, 1, math.prod(shape)
    # drop size-1 dims from consideration: they are free to sit on either side
    eff = [d for d in keep if shape[d] != 1]
    if not eff:
        return 1, 1, math.prod(shape)
    lo, hi = eff[0], eff[-1]
    for d in range(lo, hi + 1):
        if d not in keep and shape[d] != 1:
            return None  # kept dims are not adjacent
    return math.prod(shape[:lo]), math.prod(shape[lo:hi + 1]), math.prod(shape[hi + 1:])


def reduce_amax(x, axis=None, keepdims=True, squeeze_scalar=True):
    """reduce_amax (core_utils.py:146-183) in one read; `axis` = dims to REDUCE; result in x.dtype."""
    _cuda(x, "reduce_amax")
    if x.dtype in (torch.float8_e4m3fn, torch.float8_e5m2):
        x = x.to(torch.get_default_dtype())
    if x.dtype == torch.float64:
        return reduce_amax(x.float(), axis, keepdims, squeeze_scalar).double()  # exact unless |x| > fp32 range
    nd = x.dim()
    if axis is None:
        red = list(range(nd))
    else:
        red = [axis] if isinstance(axis, int) else list(axis)
        red = sorted({a % nd for a in red}) if nd else []
    keep = [d for d in range(nd) if d not in red]
    x = x.contiguous()
    view = _collapse_keep_axes(list(x.shape), keep)
    if view is None:
        # kept dims not adjacent: move them together (one extra copy; rare, e.g. axis=(0, 2) of a 3-D tensor)
        perm = keep + red
        xp = x.permute(perm).contiguous()
        out = reduce_amax(xp, list(range(len(keep), nd)), keepdims=False, squeeze_scalar=False)
        if keepdims:
            out = out.reshape([x.shape[d] if d in keep else 1 for d in range(nd)])
        if squeeze_scalar and out.numel() == 1:
            out = out.squeeze()
        return out
    outer, axis_size, inner = view
    bits = torch.zeros(axis_size, dtype=torch.int32, device=x.device)
    if x.numel():
        amax_into(x, outer, axis_size, inner, bits)
    out = bits.view(torch.float32).to(x.dtype)
    if axis is None:
        return out.reshape(())
    if keepdims:
        out = out.reshape([x.shape[d] if d in keep else 1 for d in range(nd)])
    else:
        out = out.reshape([x.shape[d] for d in keep])
    if squeeze_scalar and out.numel() == 1:
        out = out.squeeze()
    return out


# ------------------------------------------------------------------------------------------------
def selftest_e2m1() -> int:
    import ctypes as C
    n = C.c_uint64(0)
    check(_lib.load().mo_selftest_e2m1(C.byref(n)), "mo_selftest_e2m1")
    return n.value


def selftest_exact_div(q_max=65536, n_scales=4096, seed=0) -> int:
    import ctypes as C
    n = C.c_uint64(0)
    check(_lib.load().mo_selftest_exact_div(q_max, n_scales, seed, C.byref(n)), "mo_selftest_exact_div")
    return n.value


# ------------------------------------------------------------------------------------------------
# HOST-buffer entry points (x / y are CPU tensors, ideally pinned): H2D -> kernel -> D2H inside the call.
# ------------------------------------------------------------------------------------------------
def _host(t: torch.Tensor, what: str):
    if t.is_cuda or not t.is_contiguous():
        raise RuntimeError(f"modelopt_b200.{what}: expected a contiguous CPU tensor")


def _chunks(x):
    return max(1, -(-x.numel() * x.element_size() // (32 << 20)))


def amax_host(x: torch.Tensor) -> float:
    """max|x| of a host tensor (streams it through the GPU)."""
    import ctypes as C

    _host(x, "amax_host")
    out = C.c_float(0.0)
    _launched(_lib.load().mo_amax_host(x.data_ptr(), x.numel(), _code(x), C.byref(out)), "mo_amax_host",
              n=_chunks(x))
    return out.value


def calib_qdq_host_batch(xs, running_amax, kind="nvfp4", block=16, numerics=None, outs=None, download=True):
    """Collect + fake-quant of a batch of tensors in one pipelined call: running_amax[i] (list of floats, updated in
    place) = max(running_amax[i], max|x_i|) and y_i = fake-quant(x_i) under it.  A HOST tensor crosses the bus once per
    direction (the upload of x_{i+1} overlaps the fake-quant + download of y_i); a CUDA tensor among `xs` (weights that
    live on the GPU) is read in place.  ``download=False`` (or a ``None`` entry in ``outs``) is the calibration-only
    form: the fake-quantized tensor stays on the device and only the amax comes back.  Returns the list of outputs."""
    n = len(xs)
    assert len(running_amax) == n
    for t in xs:
        if not t.is_cuda:
            _host(t, "calib_qdq_host_batch")
        elif not t.is_contiguous():
            raise ValueError("modelopt_b200.calib_qdq_host_batch: expected contiguous tensors")
    if outs is None:
        ys = [torch.empty_like(t) if download else None for t in xs]
    else:
        ys = list(outs)
    code = _code(xs[0])
    assert all(_code(t) == code for t in xs), "one dtype per batch"
    xp = (C.c_void_p * n)(*[t.data_ptr() for t in xs])
    yp = (C.c_void_p * n)(*[t.data_ptr() if t is not None else None for t in ys])
    cols = [t.shape[-1] if t.dim() else 1 for t in xs]
    rows = (C.c_int64 * n)(*[t.numel() // max(c, 1) for t, c in zip(xs, cols)])
    colsa = (C.c_int64 * n)(*cols)
    am = (C.c_float * n)(*[float(a) for a in running_amax])
    _launched(_lib.load().mo_calib_qdq_host_batch(n, xp, yp, rows, colsa, code, am, 0 if kind == "nvfp4" else 1, int(block),
                                                  _mode(numerics)), "mo_calib_qdq_host_batch", n=3 * n)
    for i in range(n):
        running_amax[i] = am[i]
    return ys


def qdq_fp8_host(x: torch.Tensor, amax: float, out: torch.Tensor | None = None, numerics=None) -> torch.Tensor:
    _host(x, "qdq_fp8_host")
    y = torch.empty_like(x) if out is None else out
    _launched(_lib.load().mo_qdq_fp8_host(x.data_ptr(), y.data_ptr(), x.numel(), _code(x), float(amax),
                                          _mode(numerics)), "mo_qdq_fp8_host", n=_chunks(x))
    return y


def qdq_int_host(x: torch.Tensor, amax: float, num_bits=8, unsigned=False, narrow_range=True,
                 out: torch.Tensor | None = None) -> torch.Tensor:
    _host(x, "qdq_int_host")
    y = torch.empty_like(x) if out is None else out
    _launched(_lib.load().mo_qdq_int_host(x.data_ptr(), y.data_ptr(), x.numel(), _code(x), float(amax),
                                          int(num_bits), int(bool(unsigned)), int(bool(narrow_range))),
              "mo_qdq_int_host", n=_chunks(x))
    return y


def qdq_nvfp4_dynamic_host(x: torch.Tensor, global_amax: float, block=16, numerics=None,
                           out: torch.Tensor | None = None) -> torch.Tensor:
    _host(x, "qdq_nvfp4_dynamic_host")
    y = torch.empty_like(x) if out is None else out
    cols = x.shape[-1]
    _launched(_lib.load().mo_qdq_nvfp4_dynamic_host(x.data_ptr(), y.data_ptr(), x.numel() // cols, cols, _code(x),
                                                    float(global_amax), int(block), _mode(numerics)),
              "mo_qdq_nvfp4_dynamic_host", n=_chunks(x))
    return y


# ------------------------------------------------------------------------------------------------
# Histogram collection and amax searches (csrc/hist.cu, csrc/hist_search.cu)
# ------------------------------------------------------------------------------------------------
def hist_into(x, range_max, counts, amax_bits=None, skip_zeros=False):
    """Add the histogram of |x| over [0, range_max] (torch.histc binning, len(counts) bins) to `counts`
    (int64 device tensor, exact) and merge max|x| into `amax_bits` (int32 view of an fp32 word) in the same
    read.  `range_max` is a device fp32 scalar tensor."""
    _cuda(x, "hist_into")
    assert counts.dtype == torch.int64 and counts.is_contiguous() and counts.device == x.device
    assert range_max.dtype == torch.float32 and range_max.numel() == 1 and range_max.device == x.device
    x = x.contiguous()
    with _dev(x):
        _launched(_lib.load().mo_hist(x.data_ptr(), x.numel(), _code(x), range_max.data_ptr(), counts.numel(),
                                      counts.data_ptr(), amax_bits.data_ptr() if amax_bits is not None else None,
                                      int(bool(skip_zeros)), _stream(x)), "mo_hist")


def hist_collect_begin(x, state, counts, smem_bins_hint, skip_zeros, first):
    """First half of a sync-free HistogramCalibrator.collect (mo_hist_collect_begin): first batch = max|x| into
    state[MO_HIST_BATCH_AMAX]; later batches = bin |x| under the state's current range and fold max|x| in the same read."""
    _cuda(x, "hist_collect_begin")
    assert state.dtype == torch.int32 and state.numel() >= 16 and state.device == x.device
    assert counts.dtype == torch.int64 and counts.is_contiguous() and counts.device == x.device and x.is_contiguous()
    with _dev(x):
        _launched(_lib.load().mo_hist_collect_begin(x.data_ptr(), x.numel(), _code(x), state.data_ptr(), counts.data_ptr(),
                                                    counts.numel(), int(smem_bins_hint), int(bool(skip_zeros)),
                                                    int(bool(first)), _stream(x)), "mo_hist_collect_begin")


def hist_collect_end(x, state, counts, smem_bins_hint, skip_zeros, first, init_bins, numerics=None):
    """Second half (mo_hist_collect_end): first batch = fix the range and bin; later batches = the reference's range-growth
    decision on the device, plus the undo / redo passes, which return at once unless the range grew."""
    _cuda(x, "hist_collect_end")
    with _dev(x):
        _launched(_lib.load().mo_hist_collect_end(x.data_ptr(), x.numel(), _code(x), state.data_ptr(), counts.data_ptr(),
                                                  counts.numel(), int(smem_bins_hint), int(bool(skip_zeros)),
                                                  int(bool(first)), int(init_bins), _mode(numerics), _stream(x)),
                  "mo_hist_collect_end", n=2)


def hist_collect(x, state, counts, smem_bins_hint, skip_zeros, first, init_bins, numerics=None):
    """begin + end in one C call (no data-parallel exchange between the halves)."""
    _cuda(x, "hist_collect")
    assert state.dtype == torch.int32 and state.numel() >= 16 and state.device == x.device
    assert counts.dtype == torch.int64 and counts.is_contiguous() and counts.device == x.device and x.is_contiguous()
    with _dev(x):
        _launched(_lib.load().mo_hist_collect(x.data_ptr(), x.numel(), _code(x), state.data_ptr(), counts.data_ptr(),
                                              counts.numel(), int(smem_bins_hint), int(bool(skip_zeros)), int(bool(first)),
                                              int(init_bins), _mode(numerics), _stream(x)), "mo_hist_collect", n=3)


def hist_percentile_index(counts, percentile: float) -> torch.Tensor:
    """Device int32 scalar: searchsorted(cumsum(counts / counts.sum()), percentile / 100) (fp64, numpy order)."""
    _cuda(counts, "hist_percentile_index")
    if percentile < 0 or percentile > 100:
        raise ValueError("Invalid percentile. Must be in range 0 <= percentile <= 100.")
    idx = torch.empty(1, dtype=torch.int32, device=counts.device)
    with _dev(counts):
        _launched(_lib.load().mo_hist_percentile(counts.data_ptr(), counts.numel(), float(percentile), idx.data_ptr(),
                                                 _stream(counts)), "mo_hist_percentile")
    return idx


def hist_mse_index(counts, edges, num_bits, unsigned, stride=1, start_bin=128, as_written=True):
    """Device int32 scalar k: the chosen candidate is bin start_bin + k*stride (amax = its centre)."""
    _cuda(counts, "hist_mse_index")
    fp8 = num_bits == (4, 3)
    if not fp8 and not (isinstance(num_bits, int) and num_bits >= 0):
        raise TypeError("Invalid num_bits. num_bits must be a positive integer or tuple (4,3).")
    n = counts.numel()
    edges = edges.to(device=counts.device, dtype=torch.float32).contiguous()
    assert edges.numel() == n + 1
    ncand = max(0, -(-(n - start_bin) // stride))
    if ncand == 0:
        raise ValueError("attempt to get argmin of an empty sequence")
    loss = torch.empty(ncand, dtype=torch.float64, device=counts.device)
    idx = torch.empty(1, dtype=torch.int32, device=counts.device)
    with _dev(counts):
        _launched(_lib.load().mo_hist_mse(counts.data_ptr(), edges.data_ptr(), n, 8 if fp8 else int(num_bits),
                                          int(bool(unsigned)), int(fp8), int(bool(as_written)), int(stride),
                                          int(start_bin), loss.data_ptr(), idx.data_ptr(), _stream(counts)),
                  "mo_hist_mse", n=2)
    return idx, loss


def hist_entropy_index(counts, num_bits, unsigned, stride=1, start_bin=128):
    """Device int32 scalar k: the chosen amax is edges[start_bin + k*stride] (LAST minimum of the KL search)."""
    _cuda(counts, "hist_entropy_index")
    n = counts.numel()
    lib = _lib.load()
    nbytes = lib.mo_hist_entropy_scratch_bytes(n, int(stride), int(start_bin))
    if nbytes <= 0 or start_bin > n:
        raise ValueError("attempt to get argmin of an empty sequence")
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=counts.device)
    idx = torch.empty(1, dtype=torch.int32, device=counts.device)
    with _dev(counts):
        _launched(lib.mo_hist_entropy(counts.data_ptr(), n, int(num_bits), int(bool(unsigned)), int(stride),
                                      int(start_bin), scratch.data_ptr(), idx.data_ptr(), _stream(counts)),
                  "mo_hist_entropy", n=3)
    return idx


def hist_entropy_index_static(counts, num_bits, unsigned, out, stride=1, start_bin=128):
    """`hist_entropy_index` for callers that cannot allocate (CUDA graph capture, serving loops): the scratch is a fixed
    device array inside the library (`mo_hist_entropy_static`; at most `mo_hist_entropy_static_max_bins()` = 65536
    bins), and the result goes to the caller's `out` (device int32[1]).  Nothing is allocated here, by torch or by the
    library.  Calls that overlap on the GPU share the scratch and race: keep them on one stream."""
    _cuda(counts, "hist_entropy_index_static")
    _cuda(out, "hist_entropy_index_static")
    if out.dtype != torch.int32 or out.numel() < 1:
        raise TypeError("out must be a device int32 tensor with at least one element")
    with _dev(counts):
        _launched(_lib.load().mo_hist_entropy_static(counts.data_ptr(), counts.numel(), int(num_bits),
                                                     int(bool(unsigned)), int(stride), int(start_bin), out.data_ptr(),
                                                     _stream(counts)),
                  "mo_hist_entropy_static", n=3)
    return out


# ------------------------------------------------------------------------------------------------
# NVFP4 packer (csrc/pack_nvfp4.cu)
# ------------------------------------------------------------------------------------------------
def nvfp4_pack(x, block, sf2, sf=None):
    """x [..., N] (N % block == 0) -> (codes uint8 [..., N/2], sf float8_e4m3fn [..., N/block]); sf2: fp32 scalar
    tensor.  With `sf` given (float8_e4m3fn), those block scales are used."""
    _cuda(x, "nvfp4_pack")
    x = x.contiguous()
    n = x.shape[-1]
    assert n % block == 0 and block % 2 == 0
    nb = x.numel() // block
    codes = torch.empty((*x.shape[:-1], n // 2), dtype=torch.uint8, device=x.device)
    sf2 = _f32_vec(sf2, x)
    if sf is None:
        sf_out = torch.empty((*x.shape[:-1], n // block), dtype=torch.float8_e4m3fn, device=x.device)
        sf_in_ptr = None
    else:
        sf_out = sf.to(device=x.device, dtype=torch.float8_e4m3fn).contiguous()
        assert sf_out.numel() == nb
        sf_in_ptr = sf_out.data_ptr()
    with _dev(x):
        _launched(_lib.load().mo_nvfp4_pack(x.data_ptr(), nb, int(block), _code(x), sf2.data_ptr(), sf_in_ptr,
                                            codes.data_ptr(), None if sf is not None else sf_out.data_ptr(),
                                            _stream(x)), "mo_nvfp4_pack")
    return codes, sf_out


def nvfp4_unpack(codes, sf, sf2, block, dtype, fast=False):
    """Inverse of nvfp4_pack: [..., N/2] uint8 -> [..., N] `dtype`."""
    _cuda(codes, "nvfp4_unpack")
    codes = codes.contiguous()
    sf = sf.contiguous()
    n = codes.shape[-1] * 2
    nb = codes.numel() * 2 // block
    assert sf.numel() == nb and sf.element_size() == 1
    y = torch.empty((*codes.shape[:-1], n), dtype=dtype, device=codes.device)
    sf2 = _f32_vec(sf2, codes)
    with _dev(codes):
        _launched(_lib.load().mo_nvfp4_unpack(codes.data_ptr(), sf.data_ptr(), sf2.data_ptr(), nb, int(block), _code(y),
                                              int(bool(fast)), y.data_ptr(), _stream(codes)), "mo_nvfp4_unpack")
    return y


# ------------------------------------------------------------------------------------------------
# MX formats (csrc/qdq_mx.cu) and the integer / FP8 / NF4 packers (csrc/pack_int.cu)
# ------------------------------------------------------------------------------------------------
MX_FORMATS = {(4, 3): "E4M3", (5, 2): "E5M2", (3, 2): "E3M2", (2, 3): "E2M3", 8: "INT8", (8, 0): "E8M0",
              (2, 1): "E2M1", (1, 2): "E1M2", (0, 3): "E0M3", (3, 0): "E3M0"}  # tensor_quant.py:30-41


def qdq_mx(x, block, num_bits, scale_bits=(8, 0), global_amax=None, numerics=None, out=None):
    """``fused_amax_convert``: dynamic block fake-quant of the last dim to an MX element format under E8M0 (or a
    low-precision format's) block scales.  num_bits / scale_bits use the reference's (E, M) tuples or 8 for INT8."""
    _cuda(x, "qdq_mx")
    if x.dtype == torch.float64:
        return _via_f32(qdq_mx, x, block, num_bits, scale_bits, global_amax, numerics)
    if num_bits not in MX_FORMATS or num_bits == (8, 0) or scale_bits not in MX_FORMATS:
        raise NotImplementedError(f"Unsupported num_bits: {num_bits}, scale_bits: {scale_bits} for dynamic block quantization.")
    if int(block) not in (8, 16, 32):
        raise ValueError("Blocksize for fused call must be one of {8, 16, 32}")
    x = x.contiguous()
    y = torch.empty_like(x) if out is None else out
    cols = x.shape[-1] if x.dim() else 1
    rows = x.numel() // max(cols, 1)
    ga = None
    if global_amax is not None and scale_bits != (8, 0):
        ga = _f32_vec(global_amax.amax() if global_amax.numel() != 1 else global_amax, x)
    with _dev(x):
        _launched(_lib.load().mo_qdq_mx(x.data_ptr(), y.data_ptr(), rows, cols, _code(x), int(block),
                                        _lib.MO_FMT[MX_FORMATS[num_bits]], _lib.MO_FMT[MX_FORMATS[scale_bits]],
                                        None if ga is None else ga.data_ptr(), _mode(numerics), _stream(x)), "mo_qdq_mx")
    return y


def mxfp8_pack(x, sf=None, scales_only=False, numerics=None):
    """x [..., N] -> (q float8_e4m3fn [..., N], sf uint8 [..., ceil(N/32)]); `sf` given: quantize with it."""
    _cuda(x, "mxfp8_pack")
    x = x.contiguous()
    cols = x.shape[-1]
    rows = x.numel() // max(cols, 1)
    nb = (cols + 31) // 32
    q = None if scales_only else torch.empty(x.shape, dtype=torch.float8_e4m3fn, device=x.device)
    if sf is None:
        sf_t = torch.empty((*x.shape[:-1], nb), dtype=torch.uint8, device=x.device)
    else:
        sf_t = sf.to(device=x.device).contiguous()
        assert sf_t.dtype == torch.uint8 and sf_t.numel() == rows * nb
    with _dev(x):
        _launched(_lib.load().mo_mxfp8_pack(x.data_ptr(), None if q is None else q.data_ptr(), sf_t.data_ptr(), rows, cols,
                                            _code(x), int(sf is not None), _mode(numerics), _stream(x)), "mo_mxfp8_pack")
    return q, sf_t


def mxfp4_pack(x, block=32, numerics=None):
    """x (numel % block == 0) -> (codes uint8 [..., N/2], sf uint8 [numel/block, 1])."""
    _cuda(x, "mxfp4_pack")
    x = x.contiguous()
    assert x.numel() % block == 0 and block % 2 == 0 and x.shape[-1] % 2 == 0
    nb = x.numel() // block
    q = torch.empty((*x.shape[:-1], x.shape[-1] // 2), dtype=torch.uint8, device=x.device)
    sf = torch.empty((nb, 1), dtype=torch.uint8, device=x.device)
    with _dev(x):
        _launched(_lib.load().mo_mxfp4_pack(x.data_ptr(), q.data_ptr(), sf.data_ptr(), nb, int(block), _code(x),
                                            _mode(numerics), _stream(x)), "mo_mxfp4_pack")
    return q, sf


def mx_unpack(q, sf, block, dtype, fp4=False):
    """MXFP8 (q = e4m3 [..., N]) / MXFP4 (q = packed uint8 [..., N/2]) -> [..., N] `dtype`."""
    _cuda(q, "mx_unpack")
    q = q.contiguous()
    sf = sf.to(q.device).contiguous()
    cols = q.shape[-1] * (2 if fp4 else 1)
    rows = q.numel() // max(q.shape[-1], 1)
    assert sf.dtype == torch.uint8 and sf.numel() == rows * ((cols + block - 1) // block)
    y = torch.empty((*q.shape[:-1], cols), dtype=dtype, device=q.device)
    with _dev(q):
        _launched(_lib.load().mo_mx_unpack(q.data_ptr(), sf.data_ptr(), y.data_ptr(), rows, cols, int(block), int(fp4),
                                           _code(y), _stream(q)), "mo_mx_unpack")
    return y


def _pack4(fn, what, x, block, numerics):
    _cuda(x, what)
    x = x.contiguous()
    assert x.numel() % block == 0 and block % 2 == 0
    nb = x.numel() // block
    scales = torch.empty((nb, 1), dtype=x.dtype, device=x.device)
    q = torch.empty(x.numel() // 2, dtype=torch.uint8, device=x.device)
    with _dev(x):
        _launched(getattr(_lib.load(), fn)(x.data_ptr(), nb, int(block), _code(x), scales.data_ptr(), q.data_ptr(),
                                           _mode(numerics), _stream(x)), fn)
    return q, scales


def int4_pack(x, block, numerics=None):
    """flat x (numel % block == 0) -> (uint8 [numel/2], scales [numel/block, 1] = 7 / block amax in x.dtype)."""
    return _pack4("mo_int4_pack", "int4_pack", x, block, numerics)


def nf4_pack(x, block, numerics=None):
    """flat x -> (uint8 [numel/2] NF4 indices, scales [numel/block, 1] = block amax in x.dtype)."""
    return _pack4("mo_nf4_pack", "nf4_pack", x, block, numerics)


def int4_unpack(q, scales, block):
    _cuda(q, "int4_unpack")
    q = q.contiguous().reshape(-1)
    scales = scales.to(q.device).contiguous()
    nb = q.numel() * 2 // block
    assert scales.numel() == nb
    y = torch.empty(q.numel() * 2, dtype=scales.dtype, device=q.device)
    with _dev(q):
        _launched(_lib.load().mo_int4_unpack(q.data_ptr(), scales.data_ptr(), nb, int(block), _code(scales), y.data_ptr(),
                                             _stream(q)), "mo_int4_unpack")
    return y


def nf4_unpack(q, scales, block, bf16_form):
    """scales: dequantized per-block scales (any float dtype).  bf16_form: the reference's CUDA kernel (bf16 out)."""
    _cuda(q, "nf4_unpack")
    q = q.contiguous().reshape(-1)
    nb = q.numel() * 2 // block
    assert scales.numel() == nb
    out_dtype = torch.bfloat16 if (bf16_form or scales.dtype == torch.bfloat16) else torch.float32
    s32 = _f32_vec(scales, q)
    y = torch.empty(q.numel() * 2, dtype=out_dtype, device=q.device)
    with _dev(q):
        _launched(_lib.load().mo_nf4_unpack(q.data_ptr(), s32.data_ptr(), nb, int(block), _code(y), int(bool(bf16_form)),
                                            y.data_ptr(), _stream(q)), "mo_nf4_unpack")
    return y


def _scale_map(shape, scale_shape):
    """dims {O, A, I, bo, ba, bi, so, sa, si} for a tensor of `shape` whose scales (broadcast-shaped `scale_shape`,
    same rank, each dim 1, the full size, or size / block) vary along at most three collapsed groups."""
    shape, scale_shape = list(shape), list(scale_shape)
    assert len(shape) == len(scale_shape), (shape, scale_shape)
    # per dim: (size, block, n_scales)
    dims = []
    for n, m in zip(shape, scale_shape):
        assert m >= 1 and n % m == 0, f"scale shape {scale_shape} does not tile {shape}"
        dims.append((n, n // m, m))
    # merge neighbouring dims that are both unscaled (m == 1) or both fully scaled (block == 1)
    merged = []
    for n, b, m in dims:
        if n == 1:
            continue
        if merged and ((merged[-1][2] == 1 and m == 1) or (merged[-1][1] == 1 and b == 1)):
            pn, pb, pm = merged[-1]
            merged[-1] = (pn * n, pn * n if m == 1 else 1, pm * m)
        else:
            merged.append((n, b, m))
    if len(merged) > 3:
        raise NotImplementedError(f"modelopt_b200: scale layout {scale_shape} for {shape} needs more than 3 index groups")
    while len(merged) < 3:
        merged.insert(0, (1, 1, 1))
    (o, bo, mo_), (a, ba, ma), (i, bi, mi) = merged
    return [o, a, i, bo, ba, bi, ma * mi, mi, 1]


def q8_pack(x, scales, kind):
    """INT8 (kind 'int8') / FP8 E4M3 (kind 'fp8') codes of x / scales; scales broadcast-shaped like the reference's
    (per-tensor, per-axis keepdims, or N-D blocks [.., n_i / block_i, ..])."""
    _cuda(x, "q8_pack")
    x = x.contiguous()
    # 'fp8_sat' (`_to_fp8`, compiled: fp32 quotient of an fp32 scale, clamped): the scale stays fp32
    scales = scales.to(device=x.device, dtype=torch.float32 if kind == "fp8_sat" else x.dtype).contiguous()
    ss = list(scales.shape)
    if scales.numel() == 1:
        ss = [1] * x.dim()
    elif len(ss) < x.dim():
        ss = [1] * (x.dim() - len(ss)) + ss
    dims = (C.c_int64 * 9)(*_scale_map(x.shape, ss))
    q = torch.empty(x.shape, dtype=torch.int8 if kind == "int8" else torch.float8_e4m3fn, device=x.device)
    with _dev(x):
        _launched(_lib.load().mo_q8_pack(x.data_ptr(), q.data_ptr(), scales.data_ptr(), dims, _code(x),
                                         {"int8": 0, "fp8": 1, "fp8_sat": 2}[kind], _stream(x)), "mo_q8_pack")
    return q


# rows from which linear_fp8 runs on CTA pairs (csrc/gemm_2cta.cu, the kernel of the bf16 pairs with kind::f8f6f4): bit-identical, 1-7 % faster on the timed shapes
FP8_PAIR = {"min_rows": 512}


def linear_fp8(a, b, scale_a, scale_b, bias=None, out_dtype=torch.bfloat16, out=None, pair=None):
    """(a [.., K] e4m3) @ (b [N, K] e4m3)^T * scale_a * scale_b (+ bias) on tcgen05 kind::f8f6f4 (``torch._scaled_mm``
    with per-tensor scales, backends/fp8_per_tensor_gemm.py:36-50)."""
    _cuda(a, "linear_fp8")
    assert a.dtype == torch.float8_e4m3fn and b.dtype == torch.float8_e4m3fn and b.dim() == 2
    K = a.shape[-1]
    a2 = a.reshape(-1, K).contiguous()
    b = b.contiguous()
    M, N = a2.shape[0], b.shape[0]
    assert b.shape[1] == K
    y = torch.empty((M, N), dtype=out_dtype, device=a.device) if out is None else out
    sa, sb = _f32_vec(scale_a, a2), _f32_vec(scale_b, a2)
    if bias is not None:
        bias = bias.to(device=a.device, dtype=out_dtype).contiguous()
    use_pair = pair if pair is not None else (FP8_PAIR["min_rows"] is not None and M >= FP8_PAIR["min_rows"])
    fn = _lib.load().mo_linear_fp8_pair if use_pair else _lib.load().mo_linear_fp8
    with _dev(a2):
        _launched(fn(a2.data_ptr(), b.data_ptr(), sa.data_ptr(), sb.data_ptr(),
                                            None if bias is None else bias.data_ptr(), y.data_ptr(), M, N, K, _code(y),
                                            _stream(a2)), "mo_linear_fp8")
    return y.reshape(*a.shape[:-1], N)


def q8_unpack(q, scales, dtype, kind):
    _cuda(q, "q8_unpack")
    q = q.contiguous()
    scales = scales.to(q.device).contiguous()
    ss = list(scales.shape)
    if scales.numel() == 1:
        ss = [1] * q.dim()
    elif len(ss) < q.dim():
        ss = [1] * (q.dim() - len(ss)) + ss
    dims = (C.c_int64 * 9)(*_scale_map(q.shape, ss))
    y = torch.empty(q.shape, dtype=dtype, device=q.device)
    with _dev(q):
        _launched(_lib.load().mo_q8_unpack(q.data_ptr(), scales.data_ptr(), y.data_ptr(), dims, _code(scales), _code(y),
                                           0 if kind == "int8" else 1, _stream(q)), "mo_q8_unpack")
    return y


# ------------------------------------------------------------------------------------------------
# Scale searches (csrc/sweep.cu)
# ------------------------------------------------------------------------------------------------
def nvfp4_fp8_scale_sweep(x, global_amax, block_size=16, numerics=None):
    """Per-block amax (fp32 [n_blocks]) whose FP8-representable scale minimises the NVFP4 squared error."""
    _cuda(x, "nvfp4_fp8_scale_sweep")
    if not isinstance(block_size, int) or block_size <= 0:
        raise ValueError(f"block_size must be a positive int, got {block_size!r}.")
    if x.numel() % block_size != 0:
        raise ValueError(f"x.numel() ({x.numel()}) is not divisible by block_size ({block_size}).")
    x = x.contiguous()
    nb = x.numel() // block_size
    ga = _f32_vec(global_amax, x)
    best = torch.empty(nb, dtype=torch.float32, device=x.device)
    with _dev(x):
        _launched(_lib.load().mo_nvfp4_fp8_scale_sweep(x.data_ptr(), nb, block_size, _code(x), ga.data_ptr(),
                                                       best.data_ptr(), _mode(numerics), _stream(x)),
                  "mo_nvfp4_fp8_scale_sweep")
    return best


def block_hessian_into(x, hessian, block=16):
    """hessian [cin/block, block, block] fp32 += per-block X^T X of x [..., cin] (one read of x)."""
    _cuda(x, "block_hessian_into")
    cin = x.shape[-1]
    x = x.reshape(-1, cin).contiguous()
    tokens = x.shape[0]
    assert hessian.dtype == torch.float32 and hessian.is_contiguous() and hessian.shape == (cin // block, block, block)
    lib = _lib.load()
    scratch = torch.empty(max(int(lib.mo_block_hessian_scratch_bytes(tokens, cin)), 4), dtype=torch.uint8, device=x.device)
    with _dev(x):
        _launched(lib.mo_block_hessian(x.data_ptr(), tokens, cin, _code(x), int(block), hessian.data_ptr(),
                                       scratch.data_ptr(), _stream(x)), "mo_block_hessian", 2)
    return hessian


def nvfp4_fp8_scale_sweep_hessian(x, global_amax, hessian, block_size=16, numerics=None):
    """Per-block amax (fp32 [n_blocks]) whose FP8 scale minimises the Hessian-weighted NVFP4 error dw^T H dw; x is the
    weight in the blocked [n_blocks, block] layout, row-major over (cout, cin // block)."""
    _cuda(x, "nvfp4_fp8_scale_sweep_hessian")
    if hessian.dim() != 3 or hessian.shape[1] != block_size or hessian.shape[2] != block_size:
        raise ValueError(f"hessian must have shape [n_cin_blocks, {block_size}, {block_size}], got {tuple(hessian.shape)}.")
    x = x.contiguous()
    n_blocks = x.numel() // block_size
    n_cin = hessian.shape[0]
    if n_blocks % n_cin != 0:
        raise ValueError(f"n_blocks ({n_blocks}) is not divisible by n_cin_blocks ({n_cin}).")
    h = hessian.to(device=x.device, dtype=torch.float32).contiguous()
    ga = _f32_vec(global_amax, x)
    best = torch.empty(n_blocks, dtype=torch.float32, device=x.device)
    with _dev(x):
        _launched(_lib.load().mo_nvfp4_fp8_scale_sweep_hessian(x.data_ptr(), n_blocks // n_cin, n_cin, int(block_size),
                                                               _code(x), h.data_ptr(), ga.data_ptr(), best.data_ptr(),
                                                               _mode(numerics), _stream(x)),
                  "mo_nvfp4_fp8_scale_sweep_hessian")
    return best


def gptq_block_nvfp4(w_block, block_amax, global_scale, h_inv_blk, quant_block_size, block_start, numerics=None):
    """``gptq_fused_block_scalar``: (qw_block, err_block) of one GPTQ block, fp32 [rows, block]."""
    _cuda(w_block, "gptq_block_nvfp4")
    w_block = w_block.float().contiguous()
    rows, block = w_block.shape
    am = block_amax.to(device=w_block.device, dtype=torch.float32).contiguous()
    h = h_inv_blk.to(device=w_block.device, dtype=torch.float32).contiguous()
    assert h.shape == (block, block) and am.shape[0] == rows
    qw, err = torch.empty_like(w_block), torch.empty_like(w_block)
    with _dev(w_block):
        _launched(_lib.load().mo_gptq_block_nvfp4(w_block.data_ptr(), qw.data_ptr(), err.data_ptr(), am.data_ptr(),
                                                  float(global_scale), h.data_ptr(), rows, block, am.shape[1],
                                                  int(quant_block_size), int(block_start), _mode(numerics),
                                                  _stream(w_block)), "mo_gptq_block_nvfp4")
    return qw, err


def mse_sweep_into(x, cand_amax, losses, num_bits=8, unsigned=False, narrow_range=False, numerics=None):
    """losses[k, a] += sum (x - fq(x, cand_amax[k, a]))^2.  `cand_amax` is [K] (per-tensor) or [K, *amax_shape] for a
    single-axis amax of `x`; `losses` is a float64 [K, axis_size] device tensor."""
    _cuda(x, "mse_sweep_into")
    x = x.contiguous()
    K = cand_amax.shape[0]
    axis_size, inner = _axis_of(x, cand_amax[0])
    cand = cand_amax.detach().to(device=x.device, dtype=torch.float32).contiguous()
    fp8 = num_bits == (4, 3)
    assert losses.dtype == torch.float64 and losses.is_contiguous() and losses.numel() == K * axis_size
    outer = x.numel() // (axis_size * inner)
    with _dev(x):
        _launched(_lib.load().mo_mse_sweep(x.data_ptr(), outer, axis_size, inner, _code(x), cand.data_ptr(), K,
                                           8 if fp8 else int(num_bits), int(bool(unsigned)),
                                           int(bool(narrow_range)), int(fp8), _mode(numerics), losses.data_ptr(),
                                           _stream(x)), "mo_mse_sweep")


# ------------------------------------------------------------------------------------------------
# SmoothQuant / AWQ-lite support (csrc/smooth.cu, csrc/qdq_ew.cu)
# ------------------------------------------------------------------------------------------------
def qdq_int_block_dynamic(x, block, num_bits=4, unsigned=False, narrow_range=False, prescale=None, out=None,
                          return_amax=False):
    """INT-N fake quant with the amax of each `block` consecutive last-dim elements taken from the data, after an
    optional per-column `prescale` multiply (a tensor of x.shape[-1] values, rounded like `x * prescale` in x.dtype).
    ONE pass.  Returns y (and the fp32 [n_blocks] amax when asked)."""
    _cuda(x, "qdq_int_block_dynamic")
    x = x.contiguous()
    cols = x.shape[-1]
    if cols % block != 0:
        raise ValueError(f"last dim {cols} is not a multiple of block {block}")
    y = torch.empty_like(x) if out is None else out
    assert y.is_contiguous() and y.shape == x.shape and y.dtype == x.dtype
    nb = x.numel() // block
    ps = None
    if prescale is not None:
        if prescale.numel() != cols:
            raise ValueError(f"prescale has {prescale.numel()} elements, last dim is {cols}")
        ps = _prescale_t(prescale, x)
    am = torch.empty(nb, dtype=torch.float32, device=x.device) if return_amax else None
    with _dev(x):
        _launched(_lib.load().mo_qdq_int_block_dynamic(x.data_ptr(), y.data_ptr(), nb, block, _code(x),
                                                       ps.data_ptr() if ps is not None else None, cols,
                                                       int(num_bits), int(bool(unsigned)), int(bool(narrow_range)),
                                                       am.data_ptr() if am is not None else None, _stream(x)),
                  "mo_qdq_int_block_dynamic")
    return (y, am) if return_amax else y


def scale_cols(x, scale, out=None, amax_bits=None, amax_mode="none"):
    """out = (x * scale[None, :]) rounded to x.dtype over the [-1, x.shape[-1]] view (out may be x); optionally the
    amax of the result in the same pass, merged into `amax_bits` (int32 bit patterns: [1] 'tensor' or [rows] 'row')."""
    _cuda(x, "scale_cols")
    assert x.is_contiguous()
    cols = x.shape[-1]
    rows = x.numel() // cols if cols else 0
    y = torch.empty_like(x) if out is None else out
    assert y.is_contiguous() and y.shape == x.shape and y.dtype == x.dtype
    s = _f32_vec(scale, x)
    if s.numel() != cols:
        raise ValueError(f"scale has {s.numel()} elements, last dim is {cols}")
    mode = {"none": 0, "tensor": 1, "row": 2}[amax_mode]
    if mode:
        assert amax_bits is not None and amax_bits.dtype == torch.int32 and amax_bits.is_contiguous()
        assert amax_bits.numel() == (1 if mode == 1 else rows)
    with _dev(x):
        _launched(_lib.load().mo_scale_cols_amax(x.data_ptr(), y.data_ptr(), rows, cols, _code(x), s.data_ptr(),
                                                 amax_bits.data_ptr() if mode else None, mode, _stream(x)),
                  "mo_scale_cols_amax")
    return y


def _round_like(v32: torch.Tensor, dtype) -> torch.Tensor:
    return v32.to(dtype).to(torch.float32)


def _col_splits(rows: int, cols: int, per_vec: int) -> int:
    strips = max(1, -(-cols // (32 * per_vec)))
    return int(max(1, min(-(-rows // 8), -(-4 * 148 // strips), 512)))


def awq_weight_scale(weight, block_size=None):
    """AWQ get_weight_scale (model_calib.py:1453-1469): mean over rows of |w| / (blockmax + tiny), fp32 [cols]."""
    _cuda(weight, "awq_weight_scale")
    w = weight.contiguous().reshape(-1, weight.shape[-1])
    rows, cols = w.shape
    block = int(block_size) if block_size else cols
    per_vec = 8 if w.dtype == torch.float32 else 16
    splits = _col_splits(rows, cols, per_vec)
    partial = torch.empty(splits, cols, dtype=torch.float32, device=w.device)
    with _dev(w):
        _launched(_lib.load().mo_awq_weight_scale(w.data_ptr(), rows, cols, _code(w), block, partial.data_ptr(),
                                                  splits, _stream(w)), "mo_awq_weight_scale")
    return _round_like(partial.sum(0) / rows, w.dtype)


def abs_col_sum(x):
    """sum over rows of |x| per last-dim column, fp32 [cols] (AWQ get_act_scale numerator)."""
    _cuda(x, "abs_col_sum")
    x2 = x.contiguous().reshape(-1, x.shape[-1])
    rows, cols = x2.shape
    per_vec = 8 if x2.dtype == torch.float32 else 16
    splits = _col_splits(rows, cols, per_vec)
    partial = torch.empty(splits, cols, dtype=torch.float32, device=x2.device)
    with _dev(x2):
        _launched(_lib.load().mo_abs_col_sum(x2.data_ptr(), rows, cols, _code(x2), partial.data_ptr(), splits,
                                             _stream(x2)), "mo_abs_col_sum")
    return partial.sum(0)


def awq_act_scale(x):
    """AWQ get_act_scale (model_calib.py:1471-1472): x.abs().view(-1, C).mean(0).to(float32)."""
    rows = x.numel() // x.shape[-1]
    return _round_like(abs_col_sum(x) / rows, x.dtype)


def sq_diff_mean_into(a, b, acc):
    """acc (float64 scalar tensor on the device) += mean((a - b)^2), the difference rounded to a.dtype first."""
    _cuda(a, "sq_diff_mean_into")
    assert a.shape == b.shape and a.dtype == b.dtype and acc.dtype == torch.float64 and acc.numel() == 1
    a, b = a.contiguous(), b.contiguous()
    with _dev(a):
        _launched(_lib.load().mo_sq_diff_mean(a.data_ptr(), b.data_ptr(), a.numel(), _code(a), acc.data_ptr(),
                                              _stream(a)), "mo_sq_diff_mean")
    return acc


AWQ_CLIP_MAX_BLOCK = 256
_AWQ_CLIP_RATIOS_PER_LAUNCH = 12


def awq_clip_loss_into(weight, tokens, w_amax, ratios, loss, block, num_bits=4, unsigned=False, narrow_range=False,
                       amax_dtype=torch.float32):
    """loss[k, r, nb] += mean_t (sum_{j in block nb} tokens[t, j] * (QDQ_k(weight)[r, j] - weight[r, j]))^2 for every
    clip ratio k (amax = w_amax * ratios[k]): the blocked branch of `_clip_search` (model_calib.py:1800-1858), all
    ratios in one pass over the weight.  ``tokens`` is a 2-D view whose rows may be strided (``inputs[0::s]``);
    ``w_amax`` is passed in fp32 and ``amax_dtype`` names the dtype the amax buffer really has (the candidate
    ``w_amax * ratio`` is rounded to it, as the reference's tensor product is)."""
    _cuda(weight, "awq_clip_loss_into")
    co, ci = weight.shape
    assert tokens.dim() == 2 and tokens.shape[1] == ci and tokens.stride(1) == 1 and tokens.dtype == weight.dtype
    assert weight.is_contiguous() and 1 <= block <= AWQ_CLIP_MAX_BLOCK
    nb = -(-ci // block)
    assert w_amax.dtype == torch.float32 and w_amax.numel() == co * nb and w_amax.is_contiguous()
    assert loss.dtype == torch.float32 and tuple(loss.shape) == (len(ratios), co, nb) and loss.is_contiguous()
    lib = _lib.load()
    with _dev(weight):
        for k0 in range(0, len(ratios), _AWQ_CLIP_RATIOS_PER_LAUNCH):
            part = [float(r) for r in ratios[k0:k0 + _AWQ_CLIP_RATIOS_PER_LAUNCH]]
            arr = (C.c_float * len(part))(*part)
            _launched(lib.mo_awq_clip_loss(weight.data_ptr(), tokens.data_ptr(), co, ci, block, tokens.shape[0],
                                           tokens.stride(0), _code(weight), w_amax.data_ptr(),
                                           _DT[amax_dtype], arr, len(part), num_bits,
                                           int(unsigned), int(narrow_range), loss[k0:].data_ptr(), _stream(weight)),
                      "mo_awq_clip_loss")
    return loss


# ------------------------------------------------------------------------------------------------
# Quant-aware linear (csrc/gemm.cu): tcgen05 GEMM
# ------------------------------------------------------------------------------------------------
def linear_supported(x, weight, bias=None) -> bool:
    """Shapes / dtypes the tcgen05 kernel takes (bf16 or f16, K and N multiples of 8, 16-byte aligned rows)."""
    if not (x.is_cuda and weight.is_cuda and x.dtype == weight.dtype and x.dtype in (torch.bfloat16, torch.float16)
            and weight.dim() == 2 and x.dim() >= 1 and x.shape[-1] == weight.shape[1]):
        return False
    N, K = weight.shape
    if K < 8 or K % 8 or N % 8 or (bias is not None and (bias.dtype != x.dtype or bias.numel() != N)):
        return False
    # TMA descriptors and the 16-byte epilogue accesses need 16-byte aligned bases: a contiguous view with an odd storage
    # offset (flat-parameter slices) takes the caller's eager path instead of failing with MO_ERR_ARG.  A non-contiguous
    # operand is copied by .contiguous() in the wrappers, and fresh allocations are aligned.
    for t in (x, weight, bias):
        if t is not None and t.is_contiguous() and t.data_ptr() % 16:
            return False
    return True


# rows from which linear / linear_loss_into run on CTA pairs (tcgen05 cta_group::2, csrc/gemm_2cta.cu); None = never.
# Bit-identical to the single-CTA kernel; 1-5 % faster from 2048 rows up (profiles/r01_gemm_pair.txt); below 512 rows a
# 256-row pair tile would be mostly padding.
LINEAR_PAIR = {"min_rows": 512}


def _pair_fn(M, pair):
    use = pair if pair is not None else (LINEAR_PAIR["min_rows"] is not None and M >= LINEAR_PAIR["min_rows"])
    return _lib.load().mo_linear_pair if use else _lib.load().mo_linear


# Which GEMM a quantized linear's PLAIN product (no fused prologue / loss epilogue / real-quant operands) runs on:
#   "tcgen05" = this library's kernels (mo_linear / mo_linear_pair), "cublas" = torch's F.linear, "auto" = whichever was
#   measured faster for the shape on B200.  Inside a Llama-3-8B forward cuBLAS's nvjet 256x256 pair kernel sustains ~1600
#   TFLOP/s against ~1330 for mo_linear_pair (profiles/r02_model_profile.txt), so "auto" keeps a plain GEMM on the library
#   the reference itself calls (quant_linear.py:38-53 -> F.linear); the tcgen05 kernels carry what cuBLAS cannot do
#   (loss epilogue, QDQ prologue, NVFP4 / FP8 real-quant operands).  Override: MODELOPT_B200_PLAIN_GEMM=tcgen05|cublas|auto.
PLAIN_GEMM = {"backend": os.environ.get("MODELOPT_B200_PLAIN_GEMM", "auto").lower()}


def plain_linear(x, weight, bias=None):
    """The quantized linear's plain product, on the backend PLAIN_GEMM selects (operands already fake-quantized)."""
    be = PLAIN_GEMM["backend"]
    if be == "tcgen05":
        return linear(x, weight, bias)
    return torch.nn.functional.linear(x, weight, bias)


def linear(x, weight, bias=None, out=None, pair=None):
    """x[..., K] @ weight[N, K]^T (+ bias) -> [..., N]; fp32 accumulation in TMEM, output in x.dtype."""
    _cuda(x, "linear")
    if not linear_supported(x, weight, bias):
        raise ValueError(f"modelopt_b200.linear: unsupported operands x={tuple(x.shape)} {x.dtype}, "
                         f"weight={tuple(weight.shape)} {weight.dtype} (bf16/f16, K and N multiples of 8)")
    x2 = x.contiguous().reshape(-1, x.shape[-1])
    w = weight.contiguous()
    b = bias.contiguous() if bias is not None else None
    M, K = x2.shape
    N = w.shape[0]
    y = torch.empty(*x.shape[:-1], N, dtype=x.dtype, device=x.device) if out is None else out
    assert y.is_contiguous() and y.numel() == M * N and y.dtype == x.dtype
    with _dev(x):
        _launched(_pair_fn(M, pair)(x2.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None,
                                    y.data_ptr(), M, N, K, _code(x2), None, None, _stream(x)), "mo_linear")
    return y


def linear_loss_into(x, weight, bias, ref, acc, pair=None):
    """acc (float64 device scalar) += mean((linear(x, weight, bias) - ref)^2) without writing the output."""
    _cuda(x, "linear_loss_into")
    if not linear_supported(x, weight, bias):
        raise ValueError("modelopt_b200.linear_loss_into: unsupported operands")
    x2 = x.contiguous().reshape(-1, x.shape[-1])
    w = weight.contiguous()
    b = bias.contiguous() if bias is not None else None
    r = ref.contiguous()
    M, K = x2.shape
    N = w.shape[0]
    assert r.numel() == M * N and r.dtype == x.dtype and acc.dtype == torch.float64 and acc.numel() == 1
    with _dev(x):
        _launched(_pair_fn(M, pair)(x2.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None, None,
                                    M, N, K, _code(x2), r.data_ptr(), acc.data_ptr(), _stream(x)), "mo_linear")
    return acc


# Which input formats take the fused-prologue GEMM, and up to how many rows of x.  Measured on B200
# (profiles/r01_gemm_check.txt): at M = 16384 the prologue warps cannot keep up with the tensor pipe (1.13 ms vs 0.44 ms
# for QDQ kernel + GEMM), so large batches use two kernels; the fused form is kept for small M where one launch and no
# round trip of xq through HBM is what counts.
FUSED_PROLOGUE = {"fp8": True, "int": True, "max_rows": 256}


def linear_qdq(x, weight, bias, x_amax, num_bits=8, narrow_range=False, prescale=None, numerics=None, out=None):
    """linear(fake_quant(x * prescale, x_amax), weight, bias) in ONE kernel: the per-tensor input fake-quant
    (num_bits == (4, 3): FP8 E4M3, else signed INT-num_bits) runs in the GEMM's operand-A prologue."""
    _cuda(x, "linear_qdq")
    if not linear_supported(x, weight, bias):
        raise ValueError("modelopt_b200.linear_qdq: unsupported operands")
    x2 = x.contiguous().reshape(-1, x.shape[-1])
    w = weight.contiguous()
    b = bias.contiguous() if bias is not None else None
    M, K = x2.shape
    N = w.shape[0]
    am = _f32_vec(x_amax, x2)
    if am.numel() != 1:
        raise ValueError("linear_qdq: per-tensor input amax only")
    ps, _ = _prescale_vec(prescale, x2, f32=True)
    fp8 = num_bits == (4, 3)
    y = torch.empty(*x.shape[:-1], N, dtype=x.dtype, device=x.device) if out is None else out
    with _dev(x):
        _launched(_lib.load().mo_linear_qdq(x2.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None,
                                            y.data_ptr(), M, N, K, _code(x2), am.data_ptr(),
                                            ps.data_ptr() if ps is not None else None, 8 if fp8 else int(num_bits),
                                            int(fp8), int(bool(narrow_range)), _mode(numerics), _stream(x)),
                  "mo_linear_qdq")
    return y


# ------------------------------------------------------------------------------------------------
# NVFP4 W4A4 linear on the block-scaled tensor-core path (csrc/gemm_nvfp4.cu)
# ------------------------------------------------------------------------------------------------
def nvfp4_sf_swizzle(sf, row_pad=128):
    """E4M3 block scales [rows, kblocks] -> uint8 buffer in the 128 x 4 tile order tcgen05 reads (zero padded)."""
    _cuda(sf, "nvfp4_sf_swizzle")
    assert sf.element_size() == 1
    sf2d = sf.contiguous().reshape(-1, sf.shape[-1])
    rows, kblocks = sf2d.shape
    lib = _lib.load()
    nbytes = lib.mo_nvfp4_sf_swizzle_bytes(rows, kblocks, int(row_pad))
    if nbytes < 0:
        raise ValueError("nvfp4_sf_swizzle: row_pad must be 128 or 256")
    out = torch.empty(nbytes, dtype=torch.uint8, device=sf.device)
    with _dev(sf):
        _launched(lib.mo_nvfp4_sf_swizzle(sf2d.data_ptr(), rows, kblocks, int(row_pad), out.data_ptr(), _stream(sf)),
                  "mo_nvfp4_sf_swizzle")
    return out


def nvfp4_quantize_operand(x, global_amax, row_pad):
    """x [rows, K] -> (codes [rows, K/2] uint8, swizzled scale buffer, fp32 global scale [1]) for linear_nvfp4;
    the numbers are NVFP4QTensor.quantize's (global scale = amax / (6 * 448))."""
    _cuda(x, "nvfp4_quantize_operand")
    gs = div_scalar(_f32_vec(global_amax, x), 6.0 * 448.0)
    x2 = x.contiguous().reshape(-1, x.shape[-1])
    rows, cols = x2.shape
    if cols % 64:
        codes, sf = nvfp4_pack(x2, 16, gs)
        return codes, nvfp4_sf_swizzle(sf, row_pad), gs
    lib = _lib.load()
    nbytes = lib.mo_nvfp4_sf_swizzle_bytes(rows, cols // 16, int(row_pad))
    alloc = torch.empty if rows % row_pad == 0 else torch.zeros
    sfw = alloc(nbytes, dtype=torch.uint8, device=x.device)
    codes = torch.empty(rows, cols // 2, dtype=torch.uint8, device=x.device)
    with _dev(x):
        _launched(lib.mo_nvfp4_pack_gemm(x2.data_ptr(), rows, cols, _code(x2), gs.data_ptr(), codes.data_ptr(),
                                         sfw.data_ptr(), _stream(x)), "mo_nvfp4_pack_gemm")
    return codes, sfw, gs


# Which NVFP4 GEMM runs: "auto" = the CTA-pair kernel (256 x 256 pair tile) when there are more than 128 rows, the single-CTA
# kernel (128 x 256) otherwise; "pair" / "single" force one (MODELOPT_B200_NVFP4_GEMM).
NVFP4_GEMM = {"form": os.environ.get("MODELOPT_B200_NVFP4_GEMM", "auto")}


def linear_nvfp4(a_codes, a_sf, a_gs, b_codes, b_sf, b_gs, bias=None, out_dtype=torch.bfloat16, out=None, pair=None):
    """out = (deq(a) @ deq(b)^T) + bias from packed NVFP4 operands (see nvfp4_quantize_operand), fp32 accumulation on
    tcgen05.mma.kind::mxf4nvf4.block_scale.  pair: None = NVFP4_GEMM's choice, True / False = force the CTA-pair /
    single-CTA kernel (same results)."""
    _cuda(a_codes, "linear_nvfp4")
    a2 = a_codes.contiguous().reshape(-1, a_codes.shape[-1])
    b2 = b_codes.contiguous()
    M, K = a2.shape[0], a2.shape[1] * 2
    N = b2.shape[0]
    if b2.shape[1] * 2 != K or K % 256 or N % 8:
        raise ValueError(f"linear_nvfp4: K={K} must be a multiple of 256 (both operands), N={N} of 8")
    lib = _lib.load()
    if a_sf.numel() != lib.mo_nvfp4_sf_swizzle_bytes(M, K // 16, 128) or \
            b_sf.numel() != lib.mo_nvfp4_sf_swizzle_bytes(N, K // 16, 256):
        raise ValueError("linear_nvfp4: scale buffers must come from nvfp4_sf_swizzle(row_pad=128 / 256)")
    b = bias.contiguous() if bias is not None else None
    if b is not None and (b.dtype != out_dtype or b.numel() != N):
        raise ValueError("linear_nvfp4: bias must have the output dtype and N elements")
    y = torch.empty(*a_codes.shape[:-1], N, dtype=out_dtype, device=a_codes.device) if out is None else out
    code = _DT[out_dtype]
    if pair is None:
        form = NVFP4_GEMM["form"]
        pair = form == "pair" or (form == "auto" and M > 128)
    name = "mo_linear_nvfp4_pair" if pair else "mo_linear_nvfp4"
    with _dev(a_codes):
        _launched(getattr(lib, name)(a2.data_ptr(), a_sf.data_ptr(), a_gs.data_ptr(), b2.data_ptr(), b_sf.data_ptr(),
                                     b_gs.data_ptr(), b.data_ptr() if b is not None else None, y.data_ptr(), M, N, K,
                                     code, _stream(a_codes)), name)
    return y
