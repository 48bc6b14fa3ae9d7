"""CPU: oracle/histogram.py reproduces the REAL reference's HistogramCalibrator (tests/golden/hist_cpu.npz,
written by oracle/gen_golden.py from /root/reference): counts, edges and the entropy / mse / percentile amax."""

import pytest
import torch

from conftest import from_np, load_golden
from oracle import histogram as ohist

CASES = [("grow", 3, ((8, False), (4, False))), ("same", 3, ((8, False), (4, False))),
         ("pos", 2, ((8, False), (4, False), (8, True)))]
DT = {"f32": torch.float32, "bf16": torch.bfloat16}


@pytest.mark.parametrize("dn", list(DT))
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_histogram_oracle_matches_reference(case, dn):
    g = load_golden("hist_cpu.npz")
    cn, nb, cfgs = case
    for bits, uns in cfgs:
        c = ohist.HistogramCalibrator(bits, None, uns)
        for i in range(nb):
            c.collect(from_np(g[f"{cn}_{dn}_x{i}"], DT[dn]))
            assert torch.equal(c._calib_hist, torch.from_numpy(g[f"{cn}_{dn}_hist{i}"])), (cn, i)
            assert torch.equal(c._calib_bin_edges, torch.from_numpy(g[f"{cn}_{dn}_edges{i}"])), (cn, i)
        k = f"{cn}_{dn}_b{bits}_u{int(uns)}"
        for pct in (99.0, 99.9, 99.99, 100.0):
            assert float(c.compute_amax("percentile", percentile=pct)) == float(g[f"{k}_pct{pct}"]), pct
        assert float(c.compute_amax("mse")) == float(g[f"{k}_mse"])
        assert float(c.compute_amax("mse", stride=4, start_bin=256)) == float(g[f"{k}_mse_s4"])
        if dn == "f32" or bits == 8:  # keep the CPU suite short: entropy is O(bins^2)
            assert float(c.compute_amax("entropy")) == float(g[f"{k}_entropy"])
        assert float(c.compute_amax("entropy", stride=3, start_bin=200)) == float(g[f"{k}_entropy_s3"])


def test_mse_reference_bug_is_reproduced_and_flagged():
    """the as-written mse search returns a meaningless amax; `intended` does not."""
    g = load_golden("hist_cpu.npz")
    c = ohist.HistogramCalibrator(8, None, False)
    for i in range(3):
        c.collect(from_np(g[f"grow_f32_x{i}"], torch.float32))
    buggy, fixed = float(c.compute_amax("mse")), float(c.compute_amax("mse", intended_mse=True))
    pct = float(c.compute_amax("percentile", percentile=99.99))
    assert buggy < 0.2 * pct and 0.5 * pct < fixed < 2.0 * pct
    with pytest.raises(TypeError):
        ohist.HistogramCalibrator((4, 3), None, False).__class__.compute_amax(
            _with_hist(c, (4, 3)), "mse")


def _with_hist(c, bits):
    o = ohist.HistogramCalibrator(bits, None, False)
    o._calib_hist, o._calib_bin_edges = c._calib_hist, c._calib_bin_edges
    return o
