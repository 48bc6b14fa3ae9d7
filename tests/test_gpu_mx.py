"""GPU parity of the MX dynamic block fake-quant (csrc/qdq_mx.cu, `fused_amax_convert` replacement) through the C ABI:
MO_NUMERICS_CPU vs the oracle (oracle/mx.py), MO_NUMERICS_GPU vs the reference's CUDA extension captured on a B200
(tests/golden/mx_b200.npz), presets through mtq.quantize, and size-independent properties at full sizes."""

import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import modelopt_b200 as mtq  # noqa: E402
from modelopt_b200 import config as qcfg  # noqa: E402
from modelopt_b200 import ops  # noqa: E402
from modelopt_b200.tensor_quant import dynamic_block_quant  # noqa: E402
from oracle import mx as omx  # noqa: E402

GOLD = os.path.join(os.path.dirname(__file__), "golden", "mx_b200.npz")
GB = np.load(GOLD) if os.path.exists(GOLD) else None
DT = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}
ELEM = {"E4M3": (4, 3), "E5M2": (5, 2), "INT8": 8, "E0M3": (0, 3), "E1M2": (1, 2), "E3M0": (3, 0), "E2M1": (2, 1),
        "E3M2": (3, 2), "E2M3": (2, 3)}


def T(key, dtype=None):
    a = torch.from_numpy(GB[key])
    if dtype in (torch.bfloat16, torch.float16) and a.dtype == torch.int16:
        return a.view(dtype)
    return a


def same(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return bool(torch.all((a == b) | (torch.isnan(a) & torch.isnan(b))))


def _input(dt):
    torch.manual_seed(11)
    x = torch.randn(48, 160) * torch.rand(48, 1) * 6
    x[0] = 0
    x[1] *= 1e-7
    x[2, 7] = 5e4
    x[3, :32] = torch.linspace(-7.5, 7.5, 32)
    x[4, :8] = torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5.0, 6.0])
    return x.to(dt)


@pytest.mark.parametrize("dn", list(DT))
@pytest.mark.parametrize("fmt", list(ELEM))
@pytest.mark.parametrize("block", [32, 16, 8])
def test_e8m0_vs_oracle(dn, fmt, block):
    x = _input(DT[dn])
    got = ops.qdq_mx(x.cuda(), block, ELEM[fmt], (8, 0), numerics="cpu")
    assert same(got, omx.mx_fake_quant(x, block, fmt)), (fmt, block)
    # ragged last dim (zero-padded last block) takes the generic kernel
    xr = x[:, :100].contiguous()
    assert same(ops.qdq_mx(xr.cuda(), block, ELEM[fmt], (8, 0), numerics="cpu"), omx.mx_fake_quant(xr, block, fmt))


@pytest.mark.parametrize("dn", list(DT))
@pytest.mark.parametrize("fmt,sfmt", [("E2M1", "E4M3"), ("E4M3", "E5M2"), ("INT8", "E4M3"), ("E3M2", "E4M3")])
@pytest.mark.parametrize("use_ga", [False, True])
def test_low_precision_scales_vs_oracle(dn, fmt, sfmt, use_ga):
    x = _input(DT[dn])
    ga = x.float().abs().amax().reshape(1) if use_ga else None
    got = ops.qdq_mx(x.cuda(), 16, ELEM[fmt], ELEM[sfmt], None if ga is None else ga.cuda(), numerics="cpu")
    assert same(got, omx.mx_fake_quant(x, 16, fmt, sfmt, ga))


@pytest.mark.skipif(GB is None, reason="tests/golden/mx_b200.npz not captured yet")
@pytest.mark.parametrize("dn", list(DT))
def test_vs_reference_extension_on_b200(dn):
    """MO_NUMERICS_GPU == the reference's --use_fast_math CUDA extension, bit for bit."""
    dt = DT[dn]
    x = T(f"x_{dn}", dt).cuda()
    n = 0
    for key in GB.files:
        if key.startswith(f"y_{dn}_"):
            _, _, fmt, block = key.split("_")
            got = ops.qdq_mx(x, int(block), ELEM[fmt], (8, 0), numerics="gpu")
            assert same(got, T(key, dt)), key
            n += 1
        elif key.startswith(f"ys_{dn}_"):
            _, _, fmt, sfmt, use_ga = key.split("_")
            ga = T(f"ga_{dn}").cuda() if use_ga == "1" else None
            got = ops.qdq_mx(x, 16, ELEM[fmt], ELEM[sfmt], ga, numerics="gpu")
            assert same(got, T(key, dt)), key
            n += 1
    xr = x[:, :100].contiguous()
    assert same(ops.qdq_mx(xr, 32, (4, 3), (8, 0), numerics="gpu"), T(f"yrag_{dn}", dt))
    assert n >= 15


@pytest.mark.skipif(GB is None, reason="tests/golden/mx_b200.npz not captured yet")
@pytest.mark.parametrize("fmt", ["E4M3", "E2M1", "E3M2"])
def test_e8m0_exponents_vs_reference_on_b200(fmt):
    """4216 amax values incl. every format_max * 2^k: the block scale chosen by the fast-math division."""
    a = T("amax_sweep").cuda()
    probe = torch.zeros(a.numel(), 8, device="cuda")
    probe[:, 0] = a
    got = ops.qdq_mx(probe, 8, ELEM[fmt], (8, 0), numerics="gpu")[:, 0]
    assert same(got, T(f"probe_{fmt}"))


@pytest.mark.parametrize("fmt", ["E4M3", "E2M1", "E3M2", "INT8"])
def test_full_size_properties(fmt):
    """[16384, 14336] bf16 (BASELINE.json's activation shape): idempotence, sign symmetry, vector == generic kernel."""
    torch.manual_seed(2)
    x = (torch.randn(16384, 14336, device="cuda") * 3).to(torch.bfloat16)
    y = ops.qdq_mx(x, 32, ELEM[fmt])
    assert torch.equal(ops.qdq_mx(y, 32, ELEM[fmt]), y)
    assert torch.equal(ops.qdq_mx(-x, 32, ELEM[fmt]), -y)
    buf = torch.empty(4096 * 4096 + 8, dtype=torch.bfloat16, device="cuda")
    un = buf[1: 1 + 4096 * 4096].view(4096, 4096)
    un.copy_(x[:4096, :4096])
    assert torch.equal(ops.qdq_mx(un, 32, ELEM[fmt]), ops.qdq_mx(x[:4096, :4096].contiguous(), 32, ELEM[fmt]))
    err = (y.float() - x.float()).abs().max().item()
    assert err <= x.float().abs().max().item() * (0.26 if fmt == "E2M1" else 0.13)
    del x, y, buf
    torch.cuda.empty_cache()


def test_functional_and_errors():
    x = torch.randn(8, 64, device="cuda")
    y = dynamic_block_quant(x, 32, None, None, (4, 3), (8, 0))
    assert y.shape == x.shape and y.dtype == x.dtype
    with pytest.raises(NotImplementedError):
        dynamic_block_quant(x, 32, None, None, (6, 1), (8, 0))
    with pytest.raises(ValueError):
        ops.qdq_mx(x, 64, (4, 3))
    with pytest.raises(RuntimeError):
        ops.qdq_mx(x.cpu(), 32, (4, 3))
    xg = x.clone().requires_grad_()
    dynamic_block_quant(xg, 32, None, None, (2, 1), (8, 0)).sum().backward()
    assert torch.equal(xg.grad, torch.ones_like(x))  # MX formats: pass-through STE (tensor_quantizer.py:289-290)


@pytest.mark.parametrize("preset", ["MXFP8_DEFAULT_CFG", "MXFP6_DEFAULT_CFG", "MXFP4_DEFAULT_CFG", "MXINT8_DEFAULT_CFG",
                                    "W4A8_MXFP4_FP8_CFG"])
def test_mx_presets_through_quantize(preset):
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(256, 512), torch.nn.ReLU(), torch.nn.Linear(512, 256)).cuda().to(torch.bfloat16)
    x = torch.randn(64, 256, device="cuda", dtype=torch.bfloat16)
    ref = model(x)
    mtq.quantize(model, getattr(qcfg, preset), lambda m: m(x))
    out = model(x)
    assert out.shape == ref.shape and torch.isfinite(out.float()).all()
    lin = model[0]
    assert lin.weight_quantizer.is_mx_format and lin.weight_quantizer.amax is None
    # the quantized linear equals F.linear of the separately fake-quantized operands
    wq = ops.qdq_mx(lin.weight.data, 32, lin.weight_quantizer.num_bits)
    want = torch.nn.functional.linear(lin.input_quantizer(x), wq, lin.bias)
    assert (lin(x).float() - want.float()).abs().max() <= 2.0 ** -6 * want.float().abs().max()
    rel = (out.float() - ref.float()).norm() / ref.float().norm()
    assert rel < (0.35 if "FP4" in preset else 0.15), rel
