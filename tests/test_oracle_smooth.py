"""CPU: the SmoothQuant / AWQ-lite oracle (oracle/smooth.py) against what the REAL reference's mtq.quantize produced
on CPU (tests/golden/quantize_cpu.npz)."""

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
