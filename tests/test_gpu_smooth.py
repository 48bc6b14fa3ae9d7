"""GPU parity: SmoothQuant / AWQ-lite kernels (through the C ABI) vs the oracle, and mtq.quantize with
INT8_SMOOTHQUANT_CFG / INT4_AWQ_CFG vs the REAL reference's CPU results (tests/golden/quantize_cpu.npz)."""

import copy

import pytest
import torch
from torch import nn

from conftest import DTYPES, bits_equal, load_golden
from oracle import fake_quant as ofq
from oracle import smooth as osm

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("shape", [(96, 256), (7, 40), (33, 1000), (2, 3, 64)])
def test_scale_cols_and_amax(dn, shape):
    from modelopt_b200 import ops

    torch.manual_seed(1)
    dt = DTYPES[dn]
    x = (torch.randn(shape) * 3).to(dt)
    s = torch.rand(shape[-1]) * 4 + 0.01
    want = (x.reshape(-1, shape[-1]) * s[None, :]).to(dt).reshape(shape)   # fold_weight: fp32 product -> dtype
    rows = x.numel() // shape[-1]
    bits_r = torch.zeros(rows, dtype=torch.int32, device="cuda")
    bits_t = torch.zeros(1, dtype=torch.int32, device="cuda")
    y = ops.scale_cols(x.cuda(), s.cuda(), amax_bits=bits_r, amax_mode="row")
    assert bits_equal(y.cpu(), want)
    assert torch.equal(bits_r.view(torch.float32).cpu(), want.reshape(rows, -1).float().abs().amax(1))
    xc = x.cuda()
    ops.scale_cols(xc, s.cuda(), out=xc, amax_bits=bits_t, amax_mode="tensor")  # in place
    assert bits_equal(xc.cpu(), want)
    assert bits_t.view(torch.float32).item() == want.float().abs().max().item()
    if len(shape) == 2:
        assert bits_equal(y.cpu(), osm.fold_weight(x, s))


@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("block,cols", [(128, 512), (16, 64), (32, 96), (24, 48), (256, 256)])
@pytest.mark.parametrize("pre", [False, True])
def test_int_block_dynamic(dn, block, cols, pre):
    from modelopt_b200 import ops

    torch.manual_seed(block + cols)
    dt = DTYPES[dn]
    x = (torch.randn(48, cols) * torch.rand(48, 1) * 5).to(dt)
    x[3] = 0
    s = (torch.rand(cols) * 2 + 0.1).to(dt) if pre else None
    for bits in (4, 8):
        want = osm.int_block_dynamic_qdq(x, block, bits, prescale=s)
        y, am = ops.qdq_int_block_dynamic(x.cuda(), block, bits, prescale=None if s is None else s.cuda(),
                                          return_amax=True)
        assert bits_equal(y.cpu(), want), (dn, block, bits)
        p = x if s is None else x * s
        assert torch.equal(am.cpu(), p.reshape(-1, block).float().abs().amax(1))


@pytest.mark.parametrize("dn", list(DTYPES))
def test_qdq_with_prescale(dn):
    from modelopt_b200 import ops

    torch.manual_seed(9)
    dt = DTYPES[dn]
    x = (torch.randn(4, 33, 256) * 2).to(dt)
    s = (torch.rand(256) * 2 + 0.1).to(dt)
    p = x * s
    am = p.float().abs().max() * 0.8
    y = ops.qdq_int(x.cuda(), am.cuda(), 8, False, False, prescale=s.cuda())
    assert bits_equal(y.cpu(), ofq.int_qdq(p, am, 8, False, False))
    y = ops.qdq_fp8(x.cuda(), am.cuda(), numerics="cpu", prescale=s.cuda())
    assert bits_equal(y.cpu(), ofq.fp8_qdq(p, am))
    # ragged last dim -> scalar path
    xr, sr = x[..., :250].contiguous(), s[:250].contiguous()
    pr = xr * sr
    y = ops.qdq_int(xr.cuda(), am.cuda(), 8, False, False, prescale=sr.cuda())
    assert bits_equal(y.cpu(), ofq.int_qdq(pr, am, 8, False, False))
    # block amax stored: [n,128] view with the prescale indexed by the original column
    bam = p.reshape(-1, 128).float().abs().amax(1, keepdim=True)
    y = ops.qdq_int(x.cuda().view(-1, 128), bam.cuda(), 4, False, False, prescale=s.cuda())
    assert bits_equal(y.cpu(), ofq.int_qdq(p.reshape(-1, 128), bam, 4, False, False))


@pytest.mark.parametrize("dn", list(DTYPES))
def test_awq_scales(dn):
    from modelopt_b200 import model_calib, ops

    torch.manual_seed(11)
    dt = DTYPES[dn]
    w = (torch.randn(300, 512) * torch.rand(300, 1)).to(dt)
    w[:, 128:256] = 0
    x = (torch.randn(5, 77, 512) * 3).to(dt)
    tol = dict(rtol=2e-6, atol=0) if dt == torch.float32 else dict(rtol=2.0 ** -7, atol=0)
    for block in (128, 16, 512, None):
        got = model_calib._awq_weight_scale(w.cuda(), block).cpu()  # one kernel when the block fits a warp
        want = osm.awq_weight_scale(w, block)
        torch.testing.assert_close(got, want, **tol)
        assert (got != want).float().mean() <= (0.5 if dt == torch.float32 else 0.02)
    got, want = ops.awq_act_scale(x.cuda()).cpu(), osm.awq_act_scale(x)
    torch.testing.assert_close(got, want, **tol)
    acc = torch.zeros(2, dtype=torch.float64, device="cuda")
    a, b = x.cuda(), (x * 1.01 + 0.01).to(dt).cuda()
    ops.sq_diff_mean_into(a, b, acc[1:2])
    ops.sq_diff_mean_into(a, b, acc[1:2])
    want = 2 * (a - b).float().pow(2).double().mean().item()
    assert acc[0].item() == 0 and abs(acc[1].item() - want) <= 1e-6 * want


def test_tensor_quantizer_fused_pre_quant_scale_equals_two_pass():
    from modelopt_b200 import ops
    from modelopt_b200.config import QuantizerAttributeConfig
    from modelopt_b200.nn import TensorQuantizer

    torch.manual_seed(2)
    x = torch.randn(64, 512, device="cuda", dtype=torch.bfloat16)
    pqs = (torch.rand(512, device="cuda") + 0.5).to(torch.bfloat16)
    cfgs = [({"num_bits": 8, "axis": None}, True), ({"num_bits": (4, 3), "axis": None}, True),
            ({"num_bits": 4, "block_sizes": {-1: 128}}, False), ({"num_bits": 4, "block_sizes": {-1: 128}}, True)]
    for cfg, with_amax in cfgs:
        q = TensorQuantizer(QuantizerAttributeConfig(**cfg)).cuda()
        q._enable_pre_quant_scale = True
        q.pre_quant_scale = pqs
        if with_amax:
            p = x * pqs
            q.amax = p.view(-1, 128).abs().amax(1, keepdim=True) if "block_sizes" in cfg else p.abs().max()
        n0 = ops.LAUNCHES[0]
        with torch.no_grad():
            fused = q(x)
        assert ops.LAUNCHES[0] - n0 == 1, cfg
        xg = x.clone().requires_grad_(True)  # gradient needed -> the general (multiply, then quantize) path
        two = q(xg).detach()
        assert bits_equal(fused, two), cfg


def _tiny(g):
    m = nn.Sequential(nn.Linear(256, 512), nn.ReLU(), nn.Linear(512, 256, bias=False))
    m.load_state_dict({k[2:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("w_")})
    return m.cuda()


def test_smoothquant_vs_reference_golden():
    import modelopt_b200 as mtq
    from modelopt_b200 import ops

    g = load_golden("quantize_cpu.npz")
    m = _tiny(g)
    batches = [torch.from_numpy(g[f"batch{i}"]).cuda() for i in range(3)]
    name = "INT8_SMOOTHQUANT_CFG"
    with ops.numerics("cpu"):
        mtq.quantize(m, copy.deepcopy(getattr(mtq, name)), lambda model: [model(b) for b in batches])
        out = m(torch.from_numpy(g["test"]).cuda())
    for n, mod in m.named_modules():
        for buf in ("_amax", "_pre_quant_scale"):
            key = f"{name}|{n}|{buf}"
            if key in g.files:
                want = torch.from_numpy(g[key])
                got = getattr(mod, buf).float().cpu()
                assert got.shape == want.shape and torch.equal(got, want), key
    for k in ("0.weight", "2.weight"):
        assert torch.equal(m.state_dict()[k].cpu(), torch.from_numpy(g[f"{name}|weight_after|{k}"])), k
    torch.testing.assert_close(out.cpu(), torch.from_numpy(g[f"{name}|out"]), rtol=2e-2, atol=2e-2)


def test_awq_lite_vs_reference_golden():
    import modelopt_b200 as mtq
    from modelopt_b200 import ops

    g = load_golden("quantize_cpu.npz")
    m = _tiny(g)
    batches = [torch.from_numpy(g[f"batch{i}"]).cuda() for i in range(3)]
    name = "INT4_AWQ_CFG"
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        with ops.numerics("cpu"):
            mtq.quantize(m, copy.deepcopy(getattr(mtq, name)), lambda model: [model(b) for b in batches])
            out = m(torch.from_numpy(g["test"]).cuda())
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    seen = 0
    for n, mod in m.named_modules():
        for buf in ("_amax", "_pre_quant_scale"):
            key = f"{name}|{n}|{buf}"
            if key in g.files:
                want = torch.from_numpy(g[key])
                got = getattr(mod, buf).float().cpu()
                assert got.shape == want.shape, key
                torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-7, msg=key)  # CUDA pow vs CPU pow: ulps
                seen += 1
        assert not hasattr(mod, "awq_lite")
    assert seen == 4
    for k in ("0.weight", "2.weight"):
        torch.testing.assert_close(m.state_dict()[k].cpu(), torch.from_numpy(g[f"{name}|weight_after|{k}"]),
                                   rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(out.cpu(), torch.from_numpy(g[f"{name}|out"]), rtol=5e-2, atol=5e-2)


@pytest.mark.parametrize("name", ["INT8_SMOOTHQUANT_CFG", "INT4_AWQ_CFG"])
def test_disable_pre_quant_scale_and_resmooth_vs_reference_golden(name):
    """model_calib.py:1169-1204 against the reference's CPU run (tests/golden/resmooth_cpu.npz, oracle/gen_golden_resmooth.py):
    folded weights bit for bit, re-calibrated weight amax, input amax, and the scale kept-but-disabled / deleted."""
    import modelopt_b200 as mtq
    from modelopt_b200 import ops
    from modelopt_b200.model_calib import disable_pre_quant_scale_and_resmooth

    q, g = load_golden("quantize_cpu.npz"), load_golden("resmooth_cpu.npz")
    m = _tiny(q)
    batches = [torch.from_numpy(q[f"batch{i}"]).cuda() for i in range(3)]
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        with ops.numerics("cpu"):
            mtq.quantize(m, copy.deepcopy(getattr(mtq, name)), lambda model: [model(b) for b in batches])
            for idx, delete in ((0, False), (2, True)):
                disable_pre_quant_scale_and_resmooth(m[idx], delete_pre_quant_scale=delete)
            out = m(torch.from_numpy(q["test"]).cuda())
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    seen = 0
    for n, mod in m.named_modules():
        if f"{name}|{n}|has_pqs" not in g.files:
            continue
        assert hasattr(mod, "_pre_quant_scale") == bool(g[f"{name}|{n}|has_pqs"]), n
        assert bool(mod._enable_pre_quant_scale) == bool(g[f"{name}|{n}|pqs_enabled"]), n
        if f"{name}|{n}|_amax" in g.files:
            want = torch.from_numpy(g[f"{name}|{n}|_amax"])
            got = mod._amax.float().cpu()
            assert got.shape == want.shape, n
            if name == "INT4_AWQ_CFG":  # the AWQ scale itself carries CUDA-vs-CPU pow() ulps (test_awq_lite_vs_reference_golden)
                torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-6)
            else:
                assert torch.equal(got, want), n
            seen += 1
    assert seen >= 2
    for k in ("0.weight", "2.weight"):
        want, got = torch.from_numpy(g[f"{name}|weight|{k}"]), m.state_dict()[k].cpu()
        if name == "INT4_AWQ_CFG":
            torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-6)
        else:
            assert torch.equal(got, want), k
    torch.testing.assert_close(out.cpu(), torch.from_numpy(g[f"{name}|out"]), rtol=2e-2, atol=2e-2)
