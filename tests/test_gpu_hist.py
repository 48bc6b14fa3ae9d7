"""GPU parity: histogram collection and the percentile / mse / entropy searches (through the C ABI) against
the oracle (the reference's algorithm; torch.histc on the same device) and the reference golden vectors."""

import pytest
import torch

from conftest import from_np, load_golden
from oracle import histogram as ohist

pytestmark = pytest.mark.gpu

DT = {"f32": torch.float32, "bf16": torch.bfloat16}
CASES = [("grow", 3, ((8, False), (4, False))), ("same", 3, ((8, False), (4, False))),
         ("pos", 2, ((8, False), (4, False), (8, True)))]


@pytest.fixture(scope="module")
def calib():
    from modelopt_b200 import calib as _c

    return _c


@pytest.mark.parametrize("dn", list(DT))
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_golden_histogram_and_searches(calib, case, dn):
    """Counts, edges and every search result of the REAL reference (CPU run, tests/golden/hist_cpu.npz)."""
    g = load_golden("hist_cpu.npz")
    cn, nb, cfgs = case
    for bits, uns in cfgs:
        c = calib.HistogramCalibrator(bits, None, uns)
        c.numerics = "cpu"  # the golden run was on CPU tensors: grown edges come from the CPU torch.arange
        for i in range(nb):
            c.collect(from_np(g[f"{cn}_{dn}_x{i}"], DT[dn]).cuda())
            want_edges = torch.from_numpy(g[f"{cn}_{dn}_edges{i}"])
            assert torch.equal(c._calib_bin_edges.cpu(), want_edges), (cn, i, "edges")
            assert torch.equal(c._calib_hist.cpu(), torch.from_numpy(g[f"{cn}_{dn}_hist{i}"])), (cn, i, "counts")
        k = f"{cn}_{dn}_b{bits}_u{int(uns)}"
        for pct in (99.0, 99.9, 99.99, 100.0):
            assert float(c.compute_amax("percentile", percentile=pct)) == float(g[f"{k}_pct{pct}"]), pct
        assert float(c.compute_amax("mse")) == float(g[f"{k}_mse"])
        assert float(c.compute_amax("mse", stride=4, start_bin=256)) == float(g[f"{k}_mse_s4"])
        assert float(c.compute_amax("entropy")) == float(g[f"{k}_entropy"])
        assert float(c.compute_amax("entropy", stride=3, start_bin=200)) == float(g[f"{k}_entropy_s3"])


@pytest.mark.parametrize("dt", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape", [(17,), (1000, 33), (64, 4096), (3, 5, 257)])
def test_collect_vs_oracle_on_device(calib, dt, shape):
    """The oracle's collect is device-agnostic torch (torch.histc): run it on the GPU beside the product."""
    g = torch.Generator().manual_seed(sum(shape))
    xs = [(torch.randn(shape, generator=g) * s).to(dt).cuda() for s in (1.0, 0.5, 2.0, 2.0, 7.0)]
    xs[2].view(-1)[0] = 0.0
    for skip in (False, True):
        c, o = calib.HistogramCalibrator(8, None, False, skip_zeros=skip), ohist.HistogramCalibrator(8, None, False, skip_zeros=skip)
        for i, x in enumerate(xs):
            c.collect(x)
            o.collect(x)
            assert torch.equal(c._calib_bin_edges.cpu(), o._calib_bin_edges.cpu()), i
            assert torch.equal(c._calib_hist, o._calib_hist), i
        assert int(c._counts.sum()) == sum(int((x != 0).sum()) if skip else x.numel() for x in xs)


def test_edge_values_land_in_the_reference_bins(calib):
    """Values exactly on bin edges, the range maximum, zeros, a non-power-of-two bin count."""
    for nbins in (2048, 1000, 37):
        edges = torch.linspace(0, 7.3, nbins + 1)
        x = torch.cat([edges, edges * 0.999999, edges * 1.000001, torch.tensor([7.3, 0.0, -7.3])]).cuda()
        c = calib.HistogramCalibrator(8, None, False, num_bins=nbins)
        c.collect(x)
        want = torch.histc(x.abs(), bins=nbins, min=0, max=float(x.abs().max()))
        assert torch.equal(c._calib_hist, want), nbins
    z = torch.zeros(100, device="cuda")
    c = calib.HistogramCalibrator(8, None, False)
    c.collect(z)
    assert torch.equal(c._calib_hist, torch.histc(z, bins=2048, min=0, max=0))


def test_counts_are_exact_beyond_fp32_saturation(calib):
    """2^25 identical values: torch.histc's fp32 accumulator would stop at 2^24, the product does not."""
    x = torch.full((1 << 25,), 1.0, device="cuda", dtype=torch.bfloat16)
    x[0] = 2.0
    c = calib.HistogramCalibrator(8, None, False)
    c.collect(x)
    assert int(c._counts.sum()) == 1 << 25
    assert int(c._counts[1024]) == (1 << 25) - 1


def test_searches_vs_oracle_on_large_histogram(calib):
    """A 5k-bin histogram from bf16 activations: percentile / mse (as written and intended) / entropy indices."""
    g = torch.Generator().manual_seed(5)
    xs = [(torch.randn(256, 4096, generator=g) * s).bfloat16() for s in (1.0, 2.5)]
    for bits, uns in ((8, False), (4, False)):
        c, o = calib.HistogramCalibrator(bits, None, uns), ohist.HistogramCalibrator(bits, None, uns)
        for x in xs:
            c.collect(x.cuda())
            o.collect(x)
        assert torch.equal(c._calib_hist.cpu(), o._calib_hist)
        for pct in (99.0, 99.99):
            assert float(c.compute_amax("percentile", percentile=pct)) == float(o.compute_amax("percentile", percentile=pct))
        assert float(c.compute_amax("mse")) == float(o.compute_amax("mse"))
        assert float(c.compute_amax("mse", intended_mse=True)) == float(o.compute_amax("mse", intended_mse=True))
        assert float(c.compute_amax("entropy", stride=7)) == float(o.compute_amax("entropy", stride=7))


def test_fp8_mse_reference_error_and_intended_path(calib):
    c = calib.HistogramCalibrator((4, 3), None, False)
    c.collect(torch.randn(64, 512, device="cuda"))
    with pytest.raises(TypeError):
        c.compute_amax("mse")
    a = float(c.compute_amax("mse", intended_mse=True))
    o = ohist.HistogramCalibrator((4, 3), None, False)
    o._calib_hist, o._calib_bin_edges = c._calib_hist.cpu(), c._calib_bin_edges.cpu()
    assert a == float(o.compute_amax("mse", intended_mse=True))


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("skip_zeros", [False, True])
@pytest.mark.parametrize("nbins", [2048, 3001])
def test_whole_equals_chunked_and_single_bin_extremes(dt, skip_zeros, nbins):
    """Size-independent property: the same data counted whole and in chunks gives identical exact counts, including
    ReLU-like data (half zeros), 2^24 elements in ONE bin (fp32 counters would stop being exact there) and every element
    equal to the range maximum."""
    from modelopt_b200 import ops

    torch.manual_seed(0)
    x = torch.randn(3000, 4096, device="cuda") * 2
    x[x < 0] = 0  # half of the elements are exactly zero
    x[7, 9] = 11.0
    x = x.to(dt)
    rmax = x.float().abs().amax().reshape(1)
    whole = torch.zeros(nbins, dtype=torch.int64, device="cuda")
    ops.hist_into(x, rmax, whole, skip_zeros=skip_zeros)
    parts = torch.zeros(nbins, dtype=torch.int64, device="cuda")
    flat = x.reshape(-1)
    step = (1 << 20) - 4096
    for off in range(0, flat.numel(), step):
        ops.hist_into(flat[off:off + step], rmax, parts, skip_zeros=skip_zeros)
    assert torch.equal(whole, parts)
    expect = flat.numel() - (int((flat == 0).sum()) if skip_zeros else 0)
    assert int(whole.sum()) == expect
    # all elements in ONE bin: 2^24 zeros -> bin 0 exactly (16-bit halves must never carry)
    z = torch.zeros(1 << 24, device="cuda", dtype=dt)
    c = torch.zeros(nbins, dtype=torch.int64, device="cuda")
    ops.hist_into(z, torch.ones(1, device="cuda"), c)
    assert int(c[0]) == 1 << 24 and int(c[1:].sum()) == 0
    # all elements equal to the range maximum -> last bin
    o = torch.ones(1 << 21, device="cuda", dtype=dt)
    c.zero_()
    ops.hist_into(o, torch.ones(1, device="cuda"), c)
    assert int(c[-1]) == 1 << 21 and int(c[:-1].sum()) == 0


def test_collect_is_capturable_in_a_cuda_graph(calib):
    """The sync-free collect enqueues kernels only (its asynchronous state read-back is skipped while the stream is being
    captured), so a captured collect - including a batch that grows the range - replays to the counts of the eager path."""
    torch.manual_seed(5)
    x0 = torch.randn(64, 1024, device="cuda", dtype=torch.bfloat16)
    xs = [torch.randn(64, 1024, device="cuda", dtype=torch.bfloat16) * s for s in (0.5, 3.0, 1.0)]   # the second one grows
    eager, graphed = calib.HistogramCalibrator(8, None, False), calib.HistogramCalibrator(8, None, False)
    eager.collect(x0)
    graphed.collect(x0)
    for x in xs:
        eager.collect(x)
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for x in xs:
                graphed.collect(x)
        g.replay()
    torch.cuda.synchronize()
    assert torch.equal(eager._calib_hist, graphed._calib_hist)
    assert torch.equal(eager._calib_bin_edges, graphed._calib_bin_edges)
    graphed.collect(xs[0])          # and the calibrator keeps working eagerly afterwards (read-back included)
    eager.collect(xs[0])
    assert torch.equal(eager._calib_hist, graphed._calib_hist)


def test_finish_stats_collection_batches_the_searches():
    """finish_stats_collection enqueues every calibrator's entropy search on a few streams and reads the indices back once;
    the amax each quantizer ends up with is the one a lone compute_amax call returns; the two-phase percentile search
    (launch_search / finish_search) equals compute_amax as well."""
    import modelopt_b200 as mtq
    from modelopt_b200 import model_calib
    from torch import nn

    torch.manual_seed(4)
    m = nn.Sequential(*[nn.Linear(256, 256) for _ in range(12)]).to("cuda", torch.bfloat16)
    mtq.replace_quant_module(m)
    mtq.set_quantizer_by_cfg(m, [{"quantizer_name": "*", "enable": False},
                                 {"quantizer_name": "*input_quantizer", "cfg": {"num_bits": 8, "calibrator": "histogram"}}])
    model_calib.enable_stats_collection(m)
    with torch.no_grad():
        for s in (1.0, 3.0, 0.5):
            m(torch.randn(64, 256, device="cuda", dtype=torch.bfloat16) * s)
    model_calib.finish_stats_collection(m, "entropy")
    for lin in m:
        q = lin.input_quantizer
        cal = q._calibrator
        assert cal._memo is not None
        cal._memo = None
        lone = cal.compute_amax("entropy")
        assert float(q._amax) == float(lone) and float(lone) > 0
        cal._memo = None
        k = int(cal.launch_search("percentile", percentile=99.9))
        two_phase = cal.finish_search("percentile", k, percentile=99.9)
        assert float(cal.compute_amax("percentile", percentile=99.9)) == float(two_phase)   # served from the memo
        cal._memo = None
        assert float(cal.compute_amax("percentile", percentile=99.9)) == float(two_phase)   # and computed afresh
