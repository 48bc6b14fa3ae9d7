"""GPU: out-of-bounds write detection with guard bands (compute-sanitizer is closed on this GPU pool —
profiles/r02_sanitizer_unavailable.txt — so the kernels' stores are fenced in here instead): every output is a window of a
larger buffer whose head and tail hold a sentinel; after the call the sentinels must be untouched and the window must
equal the result computed into a stand-alone tensor.  Ragged sizes on purpose: vector tails, TMA-clipped tiles, the
scalar fallbacks of misaligned pointers."""

import pytest
import torch

pytestmark = pytest.mark.gpu

GUARD = 4096  # elements on each side


def _window(shape, dtype, fill):
    n = 1
    for s in shape:
        n *= s
    buf = torch.full((n + 2 * GUARD,), fill, dtype=dtype, device="cuda")
    return buf, buf[GUARD:GUARD + n].view(shape)


def _intact(buf, n, fill):
    head, tail = buf[:GUARD], buf[GUARD + n:]
    if isinstance(fill, float) and fill != fill:
        return bool(torch.isnan(head.float()).all()) and bool(torch.isnan(tail.float()).all())
    return bool((head == fill).all()) and bool((tail == fill).all())


@pytest.mark.parametrize("M,N,K", [(1, 8, 8), (300, 264, 144), (129, 520, 72), (513, 1032, 264), (1000, 8, 4096)])
@pytest.mark.parametrize("pair", [False, True])
def test_linear_stores_stay_inside_the_output(M, N, K, pair):
    from modelopt_b200 import ops

    torch.manual_seed(M + N + K)
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
    b = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    want = ops.linear(x, w, b, pair=pair)
    buf, out = _window((M, N), torch.bfloat16, 768.0)
    ops.linear(x, w, b, out=out, pair=pair)
    torch.cuda.synchronize()
    assert _intact(buf, M * N, 768.0), "the kernel wrote outside its output"
    assert torch.equal(out, want)
    ref = x.float() @ w.float().t() + b.float()
    assert (out.float() - ref).abs().max().item() <= 2.0 ** -6 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("shape", [(1, 16), (7, 48), (333, 80), (1025, 4112), (3, 5, 272)])
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float32])
def test_streaming_qdq_stores_stay_inside_the_output(shape, dt):
    from modelopt_b200 import ops

    torch.manual_seed(sum(shape))
    x = (torch.randn(shape, device="cuda") * 3).to(dt)
    am = ops.reduce_amax(x)
    for name, fn in (("fp8", lambda o: ops.qdq_fp8(x, am, out=o)),
                     ("nvfp4", lambda o: ops.qdq_nvfp4_dynamic(x, am.float(), 16, None, out=o))):
        want = fn(None)
        buf, out = _window(shape, dt, 512.0)
        fn(out)
        torch.cuda.synchronize()
        assert _intact(buf, x.numel(), 512.0), name
        assert torch.equal(out, want), name
    # a misaligned input / output pair takes the kernels' scalar path
    flat = torch.zeros(x.numel() + 3, dtype=dt, device="cuda")
    xm = flat[1:1 + x.numel()].view(shape)
    xm.copy_(x)
    buf, out = _window((x.numel() + 1,), dt, 512.0)
    om = out[1:].view(shape)
    ops.qdq_fp8(xm, am, out=om)
    torch.cuda.synchronize()
    assert _intact(buf, x.numel() + 1, 512.0) and float(out[0]) == 512.0
    assert torch.equal(om, ops.qdq_fp8(x, am))


@pytest.mark.parametrize("n", [1, 17, 4097, 1 << 20])
def test_amax_and_histogram_write_only_their_words(n):
    from modelopt_b200 import ops
    from modelopt_b200.calib import HistogramCalibrator

    torch.manual_seed(n)
    x = torch.randn(n, device="cuda", dtype=torch.bfloat16)
    words = torch.full((64,), 0x11111111, dtype=torch.int32, device="cuda")
    words[32] = 0
    ops.amax_into(x, 1, 1, n, words[32:33])
    torch.cuda.synchronize()
    assert bool((words[:32] == 0x11111111).all()) and bool((words[33:] == 0x11111111).all())
    assert float(words[32:33].view(torch.float32)) == float(x.float().abs().max())
    c = HistogramCalibrator(8, None, False, num_bins=128, headroom=2)
    c.collect(x)
    c.collect(x * 1.9)   # grows inside the capacity
    assert int(c._counts.sum()) == 2 * n and c._buf.numel() == 256
    assert int(c._buf[c._num_bins:].sum()) == 0  # bins beyond the current count stay empty


def test_histogram_capacity_overflow_is_reported_not_written():
    from modelopt_b200.calib import HistogramCalibrator

    x = torch.randn(4096, device="cuda", dtype=torch.bfloat16)
    c = HistogramCalibrator(8, None, False, num_bins=128, headroom=2)
    c.collect(x)
    before = c._counts.clone()
    c.collect(x * 50)    # would need ~6400 bins: more than the 256 of capacity
    with pytest.raises(RuntimeError, match="headroom"):
        _ = c._counts
    assert torch.equal(c._buf[:128], before)  # the frozen state kept the last good histogram
    c2 = HistogramCalibrator(8, None, False)
    bad = x.clone()
    bad[7] = float("inf")
    c2.collect(bad)
    with pytest.raises(RuntimeError, match="not finite"):
        c2.compute_amax("percentile")
