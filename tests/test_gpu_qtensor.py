"""GPU parity of the real-quant packers (csrc/pack_int.cu, csrc/qdq_mx.cu packers) through the C ABI:
 * MO_NUMERICS_CPU against the reference's own CPU outputs (tests/golden/qtensor_cpu.npz) - bit-exact codes, scales and
   dequantized values;
 * MO_NUMERICS_GPU against the reference's CUDA extension / CUDA-tensor outputs captured on a B200
   (tests/golden/qtensor_b200.npz, oracle/gen_golden_gpu_mx.py);
 * the known-answer vectors of the reference's tests/gpu/torch/quantization/test_qtensor_cuda.py:108-250 through
   TensorQuantizer(fake_quant=False);
 * round-trip / size-independent properties at Llama-3-8B weight sizes."""

import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from modelopt_b200 import ops  # noqa: E402
from modelopt_b200.config import QuantizerAttributeConfig  # noqa: E402
from modelopt_b200.nn import TensorQuantizer  # noqa: E402
from modelopt_b200.qtensor import (FP8QTensor, INT4QTensor, INT8QTensor, MXFP4QTensor, MXFP8QTensor,  # noqa: E402
                                   NF4QTensor)
from oracle import qtensor as oq  # noqa: E402

GOLD = os.path.join(os.path.dirname(__file__), "golden")
G = np.load(os.path.join(GOLD, "qtensor_cpu.npz"))
GB = np.load(os.path.join(GOLD, "qtensor_b200.npz")) if os.path.exists(os.path.join(GOLD, "qtensor_b200.npz")) else None
DT = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}


def T(g, key, dtype=None):
    a = torch.from_numpy(g[key])
    if dtype in (torch.bfloat16, torch.float16) and a.dtype == torch.int16:
        return a.view(dtype)
    return a


def same(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return bool(torch.all((a == b) | (torch.isnan(a) & torch.isnan(b))))


def u8(t):
    return t.view(torch.uint8).cpu()


@pytest.mark.parametrize("dn", list(DT))
@pytest.mark.parametrize("block", [128, 32])
def test_int4_nf4_vs_reference_cpu(dn, block):
    dt = DT[dn]
    x = T(G, f"xnz_{dn}", dt).cuda()
    with ops.numerics("cpu"):
        qt, s = INT4QTensor.quantize(x, block)
        assert torch.equal(u8(qt._quantized_data), T(G, f"int4_q_{dn}_{block}")) and same(s, T(G, f"int4_s_{dn}_{block}", dt))
        assert same(qt.dequantize(scale=s, block_sizes={-1: block}), T(G, f"int4_dq_{dn}_{block}", dt))
        sbs = x.numel() // block
        qt, s = NF4QTensor.quantize(x, block, sbs)
        assert torch.equal(u8(qt._quantized_data), T(G, f"nf4_q_{dn}_{block}")) and same(s, T(G, f"nf4_s_{dn}_{block}", dt))
        if dn != "f16":
            sq, ds, zp = NF4QTensor.double_quantization(s.float(), sbs, 8)
            assert torch.equal(sq.cpu(), T(G, f"nf4_sq_{dn}_{block}"))
            # the second-level vectors are torch ops on the device: compare dequantization with the golden's own scales
            dq = qt.dequantize(scale=T(G, f"nf4_sq_{dn}_{block}").cuda(), double_scale=T(G, f"nf4_ds_{dn}_{block}").cuda(),
                               scale_zeros=T(G, f"nf4_zp_{dn}_{block}").cuda(), block_sizes={-1: block})
            assert same(dq, T(G, f"nf4_dq_{dn}_{block}", dt))


CASES = {"pt": {}, "ax0": {"axis": 0}, "ax1": {"axis": -1}, "b128": {"block_sizes": {-1: 128}},
         "b2d": {"block_sizes": {-1: 64, -2: 8}}}


@pytest.mark.parametrize("dn", list(DT))
@pytest.mark.parametrize("kind", ["int8", "fp8"])
@pytest.mark.parametrize("tag", list(CASES) + ["rag"])
def test_scaled8_vs_reference_cpu(dn, kind, tag):
    dt = DT[dn]
    cls = INT8QTensor if kind == "int8" else FP8QTensor
    x = T(G, f"xnz_{dn}", dt)
    kw = CASES.get(tag, {"block_sizes": {-1: 128}})
    if tag == "rag":
        if kind == "int8":
            pytest.skip("INT8QTensor has no padding path")
        x = x[:, :200].contiguous()
    with ops.numerics("cpu"):
        qt, s = cls.quantize(x.cuda(), **kw)
        assert torch.equal(u8(qt._quantized_data), T(G, f"{kind}_q_{dn}_{tag}")), tag
        assert same(s, T(G, f"{kind}_s_{dn}_{tag}", dt))
        dkw = {"block_sizes": kw["block_sizes"]} if "block_sizes" in kw else {}
        assert same(qt.dequantize(scale=s, **dkw), T(G, f"{kind}_dq_{dn}_{tag}", dt))


@pytest.mark.parametrize("dn", list(DT))
def test_mx_packers_vs_reference_cpu(dn):
    dt = DT[dn]
    x = T(G, f"x_{dn}", dt).cuda()
    with ops.numerics("cpu"):
        qt, s = MXFP4QTensor.quantize(x, 32)
        assert torch.equal(u8(qt._quantized_data), T(G, f"mxfp4_q_{dn}")) and torch.equal(s.cpu(), T(G, f"mxfp4_s_{dn}"))
        assert same(qt.dequantize(scale=s, block_sizes={-1: 32}), T(G, f"mxfp4_dq_{dn}", dt))
        for suffix, xx in (("", x), ("_rag", x[:, :200].contiguous())):
            qt, s = MXFP8QTensor.quantize(xx)
            assert torch.equal(u8(qt._quantized_data), T(G, f"mxfp8_q_{dn}{suffix}"))
            assert torch.equal(s.cpu(), T(G, f"mxfp8_s_{dn}{suffix}"))
            assert same(qt.dequantize(scale=s), T(G, f"mxfp8_dq_{dn}{suffix}", dt))
            if not suffix:  # quantize_with_scale == quantize with the same scales
                assert torch.equal(u8(MXFP8QTensor.quantize_with_scale(xx, s)), u8(qt._quantized_data))
        assert torch.equal(MXFP8QTensor.get_weights_scaling_factor(x).cpu(), T(G, f"mxfp8_sf_{dn}"))


@pytest.mark.skipif(GB is None, reason="tests/golden/qtensor_b200.npz not captured yet")
@pytest.mark.parametrize("dn", list(DT))
def test_packers_vs_reference_on_b200(dn):
    """GPU numerics: the reference's CUDA kernels (INT4 / NF4, fp32 only) and its eager ops on CUDA tensors."""
    dt = DT[dn]
    x = T(G, f"x_{dn}", dt).cuda()
    nz = T(G, f"xnz_{dn}", dt).cuda()
    for block in (128, 32):
        if f"int4_q_{dn}_{block}" in GB.files:
            qt, s = INT4QTensor.quantize(nz, block)
            assert torch.equal(u8(qt._quantized_data), T(GB, f"int4_q_{dn}_{block}")), ("int4", dn, block)
            assert same(s, T(GB, f"int4_s_{dn}_{block}", dt))
            assert same(qt.dequantize(scale=s, block_sizes={-1: block}), T(GB, f"int4_dq_{dn}_{block}", dt))
        if f"nf4_q_{dn}_{block}" in GB.files:
            qt, s = NF4QTensor.quantize(nz, block, nz.numel() // block)
            assert torch.equal(u8(qt._quantized_data), T(GB, f"nf4_q_{dn}_{block}")), ("nf4", dn, block)
            if f"nf4_dq_{dn}_{block}" in GB.files:
                dq = qt.dequantize(scale=T(GB, f"nf4_sq_{dn}_{block}").cuda(), double_scale=T(GB, f"nf4_ds_{dn}_{block}").cuda(),
                                   scale_zeros=T(GB, f"nf4_zp_{dn}_{block}").cuda(), block_sizes={-1: block})
                assert same(dq, T(GB, f"nf4_dq_{dn}_{block}", dt)), ("nf4 dq", dn, block)
    for kind, cls in (("int8", INT8QTensor), ("fp8", FP8QTensor)):
        for tag in ("pt", "ax0", "b128", "b2d"):
            qt, s = cls.quantize(nz, **CASES[tag])
            assert torch.equal(u8(qt._quantized_data), T(GB, f"{kind}_q_{dn}_{tag}")), (kind, tag)
            assert same(s, T(GB, f"{kind}_s_{dn}_{tag}", dt))
            dkw = {"block_sizes": CASES[tag]["block_sizes"]} if "block_sizes" in CASES[tag] else {}
            assert same(qt.dequantize(scale=s, **dkw), T(GB, f"{kind}_dq_{dn}_{tag}", dt))
    qt, s = MXFP4QTensor.quantize(x, 32)
    assert torch.equal(u8(qt._quantized_data), T(GB, f"mxfp4_q_{dn}")) and torch.equal(s.cpu(), T(GB, f"mxfp4_s_{dn}"))
    qt, s = MXFP8QTensor.quantize(x)
    assert torch.equal(u8(qt._quantized_data), T(GB, f"mxfp8_q_{dn}")) and torch.equal(s.cpu(), T(GB, f"mxfp8_s_{dn}"))
    assert same(qt.dequantize(scale=s), T(GB, f"mxfp8_dq_{dn}", dt))


bf = torch.bfloat16
KNOWN = [  # (num_bits, block_sizes, axis, input, output) - the reference's test_qtensor_accuracy fixtures
    (4, {-1: 2, "scale_bits": 8, "scale_block_sizes": {-1: 4}}, None, [[0, 1, 2, 3, 4, 5, 6, 7]],
     [[0.0000, 1.0156, 2.1719, 3.0156, 3.6094, 5.0000, 5.0625, 7.0000]]),
    (4, {-1: 2, "scale_bits": 8, "scale_block_sizes": {-1: 4}}, None, [[0, 1, 2, 3, 4, 5, 7]],
     [[0.0000, 1.0156, 2.1719, 3.0156, 3.6094, 5.0000, 7.0000]]),
    (4, {-1: 4}, None, [[0, 1, 2, 3, 4, 5, 6, 7]], [[0.0000, 0.8516, 2.1406, 2.9844, 4.0000, 5.0000, 6.0000, 7.0000]]),
    (4, {-1: 4}, None, [[0, 1, 2, 3, 4, 5, 6, 7, 3, 3]],
     [[0.0000, 0.8516, 2.1406, 2.9844, 4.0000, 5.0000, 6.0000, 7.0000, 2.9844, 2.9844]]),
    (8, None, 0, [[0, 1, 2, 3, 4, 5, 6, 7]], [[0.0000, 0.9922, 1.9844, 2.9844, 3.9688, 4.9688, 5.9688, 7.0000]]),
    (8, {-1: 2, -2: 2}, None, [[0, 1, 2, 3], [4, 5, 6, 7]], [[0.0000, 1.0234, 1.9844, 2.9844], [4.0000, 5.0000, 5.9688, 7.0000]]),
    ((4, 3), {-1: 2, -2: 2}, None, [[0, 1, 2, 3], [4, 5, 6, 7]], [[0.0000, 0.9844, 2.0000, 3.0000], [3.9375, 5.0000, 6.0000, 7.0000]]),
    ((4, 3), {-1: 2, -2: 2}, None, [[0, 1, 3], [4, 5, 7]], [[0.0000, 0.9844, 3.0000], [3.9375, 5.0000, 7.0000]]),
    ((4, 3), {-1: 2}, None, [[0, 1, 2, 3], [4, 5, 6, 7]], [[0.0000, 1.0000, 1.9219, 3.0000], [3.9375, 5.0000, 6.0000, 7.0000]]),
    ((4, 3), None, 0, [[0, 1, 2, 3], [4, 5, 6, 7]], [[0.0000, 0.9609, 1.9219, 3.0000], [4.0000, 5.0000, 6.0000, 7.0000]]),
    ((4, 3), None, None, [[0, 1, 2, 3], [4, 5, 6, 7]], [[0, 1, 2, 3], [4, 5, 6, 7]]),
]


@pytest.mark.parametrize("case", range(len(KNOWN)))
@pytest.mark.parametrize("mode", ["gpu", "cpu"])
def test_reference_known_answers(case, mode):
    num_bits, block_sizes, axis, xin, want = KNOWN[case]
    with ops.numerics(mode):
        q = TensorQuantizer(QuantizerAttributeConfig(num_bits=num_bits, block_sizes=block_sizes, fake_quant=False, axis=axis)).cuda()
        x = torch.tensor(xin, dtype=bf).cuda()
        packed = q(x)
        out = q(packed)
    assert out.shape == x.shape and out.dtype == bf
    assert torch.allclose(out.cpu(), torch.tensor(want, dtype=bf)), (out, want)


def test_mxfp4_real_vs_fake_quant():
    """test_qtensor_accuracy's MXFP4 case: real quant -> dequant stays close to the fake-quant of the same format."""
    torch.manual_seed(0)
    x = torch.randn(512, 512, device="cuda")
    bs = {-1: 32, "type": "dynamic", "scale_bits": (8, 0)}
    q = TensorQuantizer(QuantizerAttributeConfig(num_bits=(2, 1), block_sizes=bs, fake_quant=False)).cuda()
    deq = q(q(x))
    fq = TensorQuantizer(QuantizerAttributeConfig(num_bits=(2, 1), block_sizes=bs, fake_quant=True)).cuda()(x)
    assert torch.allclose(fq, deq, rtol=1e-1, atol=1e-1)


@pytest.mark.parametrize("shape", [(4096, 14336), (1024, 4096)])
def test_full_size_round_trips(shape):
    """Llama-3-8B weight shapes: pack -> unpack -> pack is a fixed point, and the vector and generic kernels agree (an
    unaligned view forces the generic path)."""
    torch.manual_seed(3)
    w = (torch.randn(shape, device="cuda") * 0.02).to(bf)
    qt, s = INT4QTensor.quantize(w, 128)
    dq = qt.dequantize(scale=s, block_sizes={-1: 128})
    assert (dq.float() - w.float()).abs().max() <= w.float().abs().max() / 7 * 0.51 + 1e-3
    buf = torch.empty(w.numel() + 8, dtype=bf, device="cuda")
    un = buf[1: 1 + w.numel()]
    un.copy_(w.reshape(-1))
    qg, sg = ops.int4_pack(un, 128)
    assert torch.equal(qg, qt._quantized_data.reshape(-1)) and torch.equal(sg, s)
    qn, sn = ops.nf4_pack(w.reshape(-1), 64)
    qng, sng = ops.nf4_pack(un, 64)
    assert torch.equal(qn, qng) and torch.equal(sn, sng)
    q8, s8 = MXFP8QTensor.quantize(w)
    d8 = q8.dequantize(scale=s8)  # re-quantizing the dequantized tensor under the SAME scales is a fixed point
    assert torch.equal(MXFP8QTensor.quantize_with_scale(d8, s8).view(torch.uint8), q8._quantized_data.view(torch.uint8))
    assert torch.all(MXFP8QTensor.get_weights_scaling_factor(d8) <= s8)
    for cls, kw in ((INT8QTensor, {"axis": 0}), (FP8QTensor, {"block_sizes": {-1: 128, -2: 128}})):
        qt8, sc = cls.quantize(w, **kw)
        d = qt8.dequantize(scale=sc, **({"block_sizes": kw["block_sizes"]} if "block_sizes" in kw else {}))
        rel = (d.float() - w.float()).abs().max() / w.float().abs().max()
        assert rel < (0.01 if cls is INT8QTensor else 0.07)
    with ops.numerics("cpu"):  # the oracle on a slice the CPU finishes quickly
        sl = w[:64].contiguous()
        q, sc = INT4QTensor.quantize(sl, 128)
        oq_q, oq_s = oq.int4_quantize(sl.cpu(), 128)
        assert torch.equal(q._quantized_data.cpu(), oq_q) and same(sc, oq_s)


@pytest.mark.parametrize("preset,ratio,tol", [("INT4_BLOCKWISE_WEIGHT_ONLY_CFG", 0.3, 0.25), ("FP8_DEFAULT_CFG", 0.55, 0.08),
                                               ("NVFP4_DEFAULT_CFG", 0.32, 0.3), ("INT8_DEFAULT_CFG", 0.55, 0.05),
                                               ("MXFP8_DEFAULT_CFG", 0.55, 0.08), ("MXFP4_DEFAULT_CFG", 0.3, 0.3)])
def test_compress_end_to_end(preset, ratio, tol):
    """mtq.quantize -> mtq.compress (compress.py:179-222): weights become packed QTensorWrapper parameters (memory as
    test_qtensor_cuda.py:57-64 checks), and the compressed forward equals the fake-quant forward to the rounding of the
    dequantized weight."""
    import modelopt_b200 as mtq
    from modelopt_b200 import config as qcfg
    from modelopt_b200.qtensor import QTensorWrapper

    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(512, 1024), torch.nn.GELU(), torch.nn.Linear(1024, 512)).cuda().to(bf)
    x = torch.randn(128, 512, device="cuda", dtype=bf)
    mtq.quantize(model, getattr(qcfg, preset), lambda m: m(x))
    with torch.no_grad():
        fake = model(x)
    before = sum(p.numel() * p.element_size() for p in model.parameters() if p.dim() == 2)
    mtq.compress(model)
    assert mtq.is_real_quantized(model)
    for lin in (model[0], model[2]):
        w = lin._parameters["weight"]
        assert isinstance(w, QTensorWrapper) and w.metadata["shape"] == (lin.out_features, lin.in_features)
        assert lin.weight_quantizer._dequantize and hasattr(lin.weight_quantizer, "_scale")
    after = sum(p.numel() * p.element_size() for p in model.parameters() if isinstance(p, QTensorWrapper))
    assert after / before <= ratio, (after, before)
    with torch.no_grad():
        real = model(x)
    rel = (real.float() - fake.float()).norm() / fake.float().norm()
    assert rel < tol, rel
    # a partial config leaves the excluded layer fake-quantized
    model2 = torch.nn.Sequential(torch.nn.Linear(512, 1024), torch.nn.GELU(), torch.nn.Linear(1024, 512)).cuda().to(bf)
    mtq.quantize(model2, getattr(qcfg, preset), lambda m: m(x))
    mtq.compress(model2, {"default": True, "*2.weight_quantizer": False})
    assert isinstance(model2[0]._parameters["weight"], QTensorWrapper)
    assert not isinstance(model2[2]._parameters["weight"], QTensorWrapper)
    assert torch.isfinite(model2(x).float()).all()
