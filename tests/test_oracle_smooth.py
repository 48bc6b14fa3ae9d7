"""CPU: the SmoothQuant / AWQ-lite oracle (oracle/smooth.py) against what the REAL reference's mtq.quantize produced
on CPU (tests/golden/quantize_cpu.npz)."""

import pytest
import torch

from conftest import load_golden
from oracle import smooth as osm


def _model_io(g):
    w0, b0, w2 = (torch.from_numpy(g[k]) for k in ("w_0.weight", "w_0.bias", "w_2.weight"))
    batches = [torch.from_numpy(g[f"batch{i}"]) for i in range(3)]
    return w0, b0, w2, batches


def test_smoothquant_first_linear():
    g = load_golden("quantize_cpu.npz")
    w0, _, _, batches = _model_io(g)
    act_amax = torch.stack([b.abs().reshape(-1, b.shape[-1]).amax(0) for b in batches]).amax(0)
    scale = osm.smooth_scale(act_amax, w0, alpha=1.0)
    key = "INT8_SMOOTHQUANT_CFG|0.{}"
    assert torch.equal(scale, torch.from_numpy(g[key.format("input_quantizer|_pre_quant_scale")]))
    w_new = osm.fold_weight(w0, 1.0 / scale)
    assert torch.equal(w_new, torch.from_numpy(g["INT8_SMOOTHQUANT_CFG|weight_after|0.weight"]))
    assert torch.equal(w_new.abs().amax(dim=1, keepdim=True),
                       torch.from_numpy(g[key.format("weight_quantizer|_amax")]))
    assert torch.equal((act_amax * scale).amax(), torch.from_numpy(g[key.format("input_quantizer|_amax")]))


def test_awq_lite_first_linear():
    g = load_golden("quantize_cpu.npz")
    w0, b0, _, batches = _model_io(g)
    alphas, losses, best, best_scale = osm.awq_lite_linear(w0, b0, batches, 128, 4, 0.1)
    assert len(alphas) == 11 and all(l > 0 for l in losses)
    pqs = 1.0 / best_scale
    assert torch.equal(pqs, torch.from_numpy(g["INT4_AWQ_CFG|0.input_quantizer|_pre_quant_scale"]))
    w_new = osm.fold_weight(w0, 1.0 / pqs)
    assert torch.equal(w_new, torch.from_numpy(g["INT4_AWQ_CFG|weight_after|0.weight"]))
    amax = w_new.view(-1, 128).abs().amax(dim=1, keepdim=True)
    assert torch.equal(amax, torch.from_numpy(g["INT4_AWQ_CFG|0.weight_quantizer|_amax"]))


def _clip_inputs(g, case, mod):
    """Sampled tokens per batch for linear `mod` of the tiny model during awq_clip (weight quantizers are off while the
    calibration data flows, so linear 2 sees relu(linear 0) unquantized)."""
    import torch.nn.functional as F

    w0, b0 = torch.from_numpy(g["w_0.weight"]), torch.from_numpy(g["w_0.bias"])
    out = []
    for i in range(2):
        x = torch.from_numpy(g[f"batch{i}"])
        if mod == "2":
            x = F.relu(F.linear(x, w0, b0))
        out.append(x.reshape(-1, x.shape[-1]))
    return out


@pytest.mark.parametrize("case,max_tokens,lo", [("int4_clip_t64", 64, 0.5), ("int4_clip_t100", 100, 0.5),
                                                ("int4_clip_step", 64, 0.3)])
@pytest.mark.parametrize("mod", ["0", "2"])
def test_awq_clip_block_search(case, max_tokens, lo, mod):
    g = load_golden("awq_clip_cpu.npz")
    w = torch.from_numpy(g[f"w_{mod}.weight"])
    w_amax = torch.from_numpy(g[f"{case}|{mod}|w_amax"])
    ratios = osm.awq_clip_ratios(lo, 0.05)
    assert ratios == [float(r) for r in g[f"{case}|{mod}|ratios"]]
    assert torch.equal(w_amax.view(-1), w.reshape(-1, 128).abs().amax(1))
    total, ntok = 0, 0
    for x in _clip_inputs(g, case, mod):
        tok = x[0::max(1, x.shape[0] // max_tokens)]
        ntok += tok.shape[0]
        total = total + osm.awq_clip_block_losses(w, tok, w_amax.view(w.shape[0], -1), ratios, 128, 4)
    assert ntok == int(g[f"{case}|{mod}|num_tokens"])
    want = torch.from_numpy(g[f"{case}|{mod}|loss"])
    torch.testing.assert_close(total, want, rtol=1e-5, atol=1e-9)
    best = osm.awq_clip_select(total, w_amax, ratios)
    assert torch.equal(best, torch.from_numpy(g[f"{case}|{mod}.weight_quantizer|_amax"]))
