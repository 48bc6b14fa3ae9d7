"""CPU: the sweep oracle (oracle/mse_sweep.py) against the REAL reference's MseCalibrator run on CPU
(tests/golden/mse_cpu.npz) and self-consistency of the NVFP4 FP8-scale sweep restatement."""

import pytest
import torch

from conftest import load_golden
from oracle import fake_quant as ofq
from oracle import mse_sweep as osw

FMTS = {"int8": lambda x, a: ofq.int_qdq(x, a, 8, False, True), "int4": lambda x, a: ofq.int_qdq(x, a, 4, False, True),
        "fp8": lambda x, a: ofq.fp8_qdq(x, a)}


@pytest.mark.parametrize("fmt", list(FMTS))
@pytest.mark.parametrize("an", ["tensor", "row"])
def test_mse_sweep_matches_reference(fmt, an):
    g = load_golden("mse_cpu.npz")
    w = torch.from_numpy(g["w"])
    amax = torch.from_numpy(g[f"{fmt}_{an}_amax0"])
    red = None if an == "tensor" else [1]
    _, l1 = osw.mse_sweep(w, amax, FMTS[fmt], red)
    _, l2 = osw.mse_sweep(w * 0.5, amax, FMTS[fmt], red)
    losses = l1 + l2
    want = torch.from_numpy(g[f"{fmt}_{an}_losses"])
    assert torch.allclose(losses.reshape(want.shape), want, rtol=1e-6, atol=0)
    best = torch.argmin(losses, dim=0)
    pick = osw.mse_candidates()[best]
    got = amax * (pick.view_as(amax) if pick.ndim else pick)
    assert torch.equal(got, torch.from_numpy(g[f"{fmt}_{an}_best"]))


def test_candidates():
    c = osw.mse_candidates()
    assert len(c) == 39 and c[0] == 0.25 and c[-1] == 4.0
    f = osw.fp8_scale_candidates()
    assert len(f) == 126 and f[-1] == 1.0 and f[0] == 2.0 ** -9 / 448.0 and bool((f[1:] > f[:-1]).all())


def test_nvfp4_sweep_picks_the_minimum():
    torch.manual_seed(3)
    x = torch.randn(256, 16) * torch.rand(256, 1) * 4
    x[3] = 0
    ga = x.abs().max()
    best, losses = osw.nvfp4_fp8_sweep(x, ga, return_losses=True)
    assert best.shape == (256,) and best[3] == ga * osw.fp8_scale_candidates()[0]  # all-zero block: first candidate
    chosen = osw.nvfp4_block_loss(x, best)
    assert torch.all(chosen <= losses.double().min(dim=0).values * (1 + 1e-5) + 1e-30)
    # the search never does worse than the max-calibrated per-block amax rounded to E4M3 (it is a candidate)
    bam = x.abs().amax(dim=1)
    assert torch.all(chosen <= osw.nvfp4_block_loss(x, ofq.compute_fp4_scales(bam, ga)[0] * 6.0) * (1 + 1e-5) + 1e-30)
