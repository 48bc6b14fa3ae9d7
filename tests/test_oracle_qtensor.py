"""CPU: the oracle's packer restatement equals the reference's own CPU outputs (tests/golden/qtensor_cpu.npz)."""

import os

import numpy as np
import pytest
import torch

from oracle import qtensor as oq

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "qtensor_cpu.npz"))
DT = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}


def T(key, dtype=None):
    a = torch.from_numpy(G[key])
    if dtype in (torch.bfloat16, torch.float16) and a.dtype == torch.int16:
        return a.view(dtype)
    return a


def same(a, b):
    a, b = a.double(), b.double()
    return bool(torch.all((a == b) | (torch.isnan(a) & torch.isnan(b))))


@pytest.mark.parametrize("dn", list(DT))
@pytest.mark.parametrize("block", [128, 32])
def test_int4_nf4(dn, block):
    dt = DT[dn]
    x = T(f"xnz_{dn}", dt)
    q, s = oq.int4_quantize(x, block)
    assert torch.equal(q, T(f"int4_q_{dn}_{block}")) and same(s, T(f"int4_s_{dn}_{block}", dt))
    assert same(oq.int4_dequantize(q, s, block, x.shape, dt), T(f"int4_dq_{dn}_{block}", dt))
    sbs = x.numel() // block
    q, s = oq.nf4_quantize(x, block, sbs)
    assert torch.equal(q, T(f"nf4_q_{dn}_{block}")) and same(s, T(f"nf4_s_{dn}_{block}", dt))
    if dn != "f16":
        sq, ds, zp = oq.nf4_double_quantization(s.float(), sbs, 8)
        assert torch.equal(sq, T(f"nf4_sq_{dn}_{block}")) and same(ds, T(f"nf4_ds_{dn}_{block}"))
        assert same(oq.nf4_dequantize(q, sq, ds, zp, block, x.shape, dt), T(f"nf4_dq_{dn}_{block}", dt))


CASES = {"pt": {}, "ax0": {"axis": 0}, "ax1": {"axis": -1}, "b128": {"block_sizes": {-1: 128}},
         "b2d": {"block_sizes": {-1: 64, -2: 8}}}


@pytest.mark.parametrize("dn", list(DT))
@pytest.mark.parametrize("kind", ["int8", "fp8"])
@pytest.mark.parametrize("tag", list(CASES) + ["rag"])
def test_scaled8(dn, kind, tag):
    dt = DT[dn]
    x = T(f"xnz_{dn}", dt)
    kw = CASES.get(tag, {"block_sizes": {-1: 128}})
    if tag == "rag":
        if kind == "int8":
            pytest.skip("INT8QTensor has no padding path")
        x = x[:, :200].contiguous()
    q, s = oq.scaled8_quantize(x, kind, **kw)
    assert torch.equal(q.view(torch.uint8), T(f"{kind}_q_{dn}_{tag}")) and same(s, T(f"{kind}_s_{dn}_{tag}", dt))
    dq = oq.scaled8_dequantize(q, s, dt, x.shape, kw.get("block_sizes"), kind)
    assert same(dq, T(f"{kind}_dq_{dn}_{tag}", dt))


@pytest.mark.parametrize("dn", list(DT))
def test_mx_packers(dn):
    dt = DT[dn]
    x = T(f"x_{dn}", dt)
    q, s = oq.mxfp4_quantize(x, 32)
    assert torch.equal(q, T(f"mxfp4_q_{dn}")) and torch.equal(s, T(f"mxfp4_s_{dn}"))
    assert same(oq.mxfp4_dequantize(q, s, 32, dt), T(f"mxfp4_dq_{dn}", dt))
    for suffix, xx in (("", x), ("_rag", x[:, :200].contiguous())):
        q, s = oq.mxfp8_quantize(xx)
        assert torch.equal(q.view(torch.uint8), T(f"mxfp8_q_{dn}{suffix}")) and torch.equal(s, T(f"mxfp8_s_{dn}{suffix}"))
        assert same(oq.mxfp8_dequantize(q, s, dt), T(f"mxfp8_dq_{dn}{suffix}", dt))
    assert torch.equal(oq.mxfp8_quantize(x)[1], T(f"mxfp8_sf_{dn}"))
