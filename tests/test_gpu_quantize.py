"""GPU: mtq.quantize(model, cfg, forward_loop) through the product's mirror of the reference API, against what the
REAL reference produced for the same model / data on CPU (tests/golden/quantize_cpu.npz, oracle/gen_golden.py)."""

import copy

import pytest
import torch
from torch import nn

from conftest import load_golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mtq():
    import modelopt_b200

    return modelopt_b200


def build(g, dtype=torch.float32):
    m = nn.Sequential(nn.Linear(256, 512), nn.ReLU(), nn.Linear(512, 256, bias=False))
    m.load_state_dict({k[2:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("w_")})
    return m.to(dtype).cuda()


def run(mtq, g, name, dtype=torch.float32):
    m = build(g, dtype)
    batches = [torch.from_numpy(g[f"batch{i}"]).to(dtype).cuda() for i in range(3)]

    def loop(model):
        for b in batches:
            model(b)

    return mtq.quantize(m, copy.deepcopy(getattr(mtq, name)), loop)


@pytest.mark.parametrize("name", ["INT8_DEFAULT_CFG", "FP8_DEFAULT_CFG", "INT4_BLOCKWISE_WEIGHT_ONLY_CFG",
                                  "NVFP4_DEFAULT_CFG"])
def test_max_calibration_matches_reference(mtq, name):
    g = load_golden("quantize_cpu.npz")
    m = run(mtq, g, name)
    seen = 0
    for n, mod in m.named_modules():
        if not isinstance(mod, mtq.TensorQuantizer):
            continue
        assert bool(g[f"{name}|{n}|enabled"]) == mod.is_enabled, n
        key = f"{name}|{n}|_amax"
        assert (key in g.files) == hasattr(mod, "_amax"), n
        if key in g.files:
            want = torch.from_numpy(g[key])
            got = mod._amax.float().cpu()
            assert got.shape == want.shape, (n, got.shape, want.shape)
            if n.startswith("0.") or "weight" in n:
                assert torch.equal(got, want), n          # weights and the first input: bit-exact
            else:
                torch.testing.assert_close(got, want, rtol=1e-5, atol=0)  # behind a GEMM (cuBLAS vs CPU sum order)
            seen += 1
    assert seen >= 2
    if f"{name}|out" in g.files:
        test = torch.from_numpy(g["test"]).cuda()
        with torch.no_grad():
            out = m(test).cpu()
        want = torch.from_numpy(g[f"{name}|out"])
        # a GEMM sits between the quantizers: values on a rounding boundary may move by one quantization step
        err = (out - want).abs()
        assert float(err.mean()) < 2e-3 * float(want.abs().mean()) + 1e-6, float(err.mean())
        assert float((err > 0.05 * want.abs().max()).float().mean()) < 1e-3


def test_quantizer_state_machine_and_summary(mtq, capsys):
    g = load_golden("quantize_cpu.npz")
    m = run(mtq, g, "FP8_DEFAULT_CFG", torch.bfloat16)
    q = m[0].input_quantizer
    assert q._if_quant and not q._if_calib and q.amax.numel() == 1 and q.amax.dtype == torch.bfloat16
    mtq.print_quant_summary(m)
    assert "6 TensorQuantizers found" in capsys.readouterr().out
    x = torch.randn(2, 16, 256, device="cuda", dtype=torch.bfloat16)
    y1 = m(x)
    mtq.disable_quantizer(m, "*input_quantizer")
    y2 = m(x)
    mtq.enable_quantizer(m, "*input_quantizer")
    assert torch.equal(m(x), y1) and not torch.equal(y1, y2)
    # fold_weight: weight <- fake-quant(weight), quantizer off, output unchanged
    mtq.fold_weight(m)
    assert not m[0].weight_quantizer.is_enabled
    torch.testing.assert_close(m(x), y1, rtol=2e-2, atol=2e-2)


def test_weight_collect_is_skipped_when_unchanged(mtq):
    """The reference re-collects every weight on every forward; the product skips it while the weight is unchanged."""
    from modelopt_b200 import ops

    g = load_golden("quantize_cpu.npz")
    m = build(g)
    mtq.replace_quant_module(m)
    mtq.set_quantizer_by_cfg(m, mtq.FP8_DEFAULT_CFG["quant_cfg"])
    mtq.enable_stats_collection(m)
    x = torch.randn(4, 256, device="cuda")
    m(x)
    n0 = ops.LAUNCHES[0]
    m(x)
    per_forward = ops.LAUNCHES[0] - n0
    assert per_forward == 2  # two input-quantizer collects; no weight re-collect
    with torch.no_grad():
        m[0].weight.mul_(2.0)  # bumps the version -> collected again
    n0 = ops.LAUNCHES[0]
    m(x)
    assert ops.LAUNCHES[0] - n0 == 3
    mtq.finish_stats_collection(m)
    assert torch.equal(m[0].weight_quantizer.amax.float(), m[0].weight.detach().abs().max())


def test_custom_backend_hook(mtq):
    """tests/unit/torch/quantization/test_custom_backend.py:33-72 of the reference."""
    calls = []

    def backend(inputs, tq):
        calls.append(tq.backend_extra_args)
        return inputs * 0 + 1

    mtq.register_quant_backend("ones", backend)
    assert mtq.is_registered_quant_backend("ones")
    m = nn.Sequential(nn.Linear(16, 16)).cuda()
    mtq.replace_quant_module(m)
    mtq.set_quantizer_by_cfg(m, [{"quantizer_name": "*", "enable": False},
                                 {"quantizer_name": "*weight_quantizer",
                                  "cfg": {"backend": "ones", "backend_extra_args": {"a": 1}}}])
    y = m(torch.randn(2, 16, device="cuda"))
    assert calls == [{"a": 1}]
    torch.testing.assert_close(y, torch.randn(2, 16, device="cuda").sum(-1, keepdim=True).expand(2, 16) * 0 + y)
    mtq.unregister_quant_backend("ones")
    assert not mtq.is_registered_quant_backend("ones")
    with pytest.raises(KeyError):
        m(torch.randn(2, 16, device="cuda"))


def test_static_nvfp4_weights_promoted_with_shared_global_amax():
    """max_calibrate ends with promote_static_block_weight_quantizers (model_calib.py:143-157, core_utils.py:1077-1150):
    static NVFP4 weight quantizers become StaticBlockScaleQuantizer, q/k/v and gate/up siblings share one global_amax,
    and the weight fake-quant is the two-level static kernel."""
    import copy

    import modelopt_b200 as mtq
    from modelopt_b200 import ops
    from modelopt_b200.nn import StaticBlockScaleQuantizer
    from oracle import fake_quant as ofq

    class Attn(nn.Module):
        def __init__(self):
            super().__init__()
            self.q_proj, self.k_proj, self.v_proj = nn.Linear(64, 64), nn.Linear(64, 32), nn.Linear(64, 32)
            self.o_proj = nn.Linear(64, 64)

        def forward(self, x):
            return self.o_proj(self.q_proj(x) + torch.cat([self.k_proj(x), self.v_proj(x)], -1))

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.attn = Attn()
            self.gate_proj, self.up_proj = nn.Linear(64, 128), nn.Linear(64, 128)

        def forward(self, x):
            return self.attn(x).sum() + (self.gate_proj(x) * self.up_proj(x)).sum()

    torch.manual_seed(3)
    net = Net()
    with torch.no_grad():
        net.attn.k_proj.weight.mul_(3.0)
        net.up_proj.weight.mul_(0.5)
    net = net.cuda()
    w4 = {"num_bits": (2, 1), "block_sizes": {-1: 16, "type": "static", "scale_bits": (4, 3)}}
    cfg = {"quant_cfg": [{"quantizer_name": "*", "enable": False}, {"quantizer_name": "*weight_quantizer", "cfg": w4}],
           "algorithm": "max"}
    x = torch.randn(8, 64, device="cuda")
    with torch.no_grad():
        mtq.quantize(net, copy.deepcopy(cfg), lambda m: m(x))
    lin = {n: m for n, m in net.named_modules() if isinstance(m, nn.Linear)}
    for n, m in lin.items():
        q = m.weight_quantizer
        assert isinstance(q, StaticBlockScaleQuantizer) and q.is_nvfp4_static, n
        assert q._amax.dtype == torch.float32 and q._global_amax.dtype == torch.float32
        assert torch.equal(q._amax.view(-1).cpu(), m.weight.detach().cpu().reshape(-1, 16).abs().amax(1)), n
    qkv = [lin[f"attn.{n}_proj"].weight_quantizer.global_amax for n in "qkv"]
    want = max(lin[f"attn.{n}_proj"].weight.abs().max() for n in "qkv")
    assert all(torch.equal(g.reshape(()), want.float().reshape(())) for g in qkv)
    assert torch.equal(lin["gate_proj"].weight_quantizer.global_amax.reshape(()),
                       torch.maximum(lin["gate_proj"].weight.abs().max(), lin["up_proj"].weight.abs().max()))
    assert torch.equal(lin["attn.o_proj"].weight_quantizer.global_amax.reshape(()), lin["attn.o_proj"].weight.abs().max())
    for n in ("attn.q_proj", "up_proj"):
        m = lin[n]
        q = m.weight_quantizer
        with ops.numerics("cpu"):  # the oracle divides in IEEE fp32; the default GPU numerics are the Triton kernel's
            got = q(m.weight).detach().cpu()
        wc = m.weight.detach().cpu()
        want_w = ofq.nvfp4_static_qdq(wc.reshape(-1, 16), q._amax.cpu().view(-1, 1), q._global_amax.cpu(), True, 448.0,
                                      None).reshape(wc.shape)
        assert torch.equal(got, want_w), n


def _all_presets():
    from modelopt_b200 import config as qcfg

    return sorted(qcfg.choices)


@pytest.mark.parametrize("preset", _all_presets())
def test_every_preset_quantizes_and_runs(preset):
    """Every preset in config.choices goes through mtq.quantize on a small MLP and produces a finite output close to the
    unquantized one (catches wiring breakage of any format / algorithm combination)."""
    import modelopt_b200 as mtq
    from modelopt_b200 import config as qcfg

    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(256, 512), torch.nn.GELU(), torch.nn.Linear(512, 256)).cuda().to(torch.bfloat16)
    batches = [torch.randn(64, 256, device="cuda", dtype=torch.bfloat16) for _ in range(2)]
    ref = model(batches[0]).float()

    def loop(m):
        for b in batches:
            m(b)

    mtq.quantize(model, getattr(qcfg, preset), loop)
    out = model(batches[0]).float()
    assert out.shape == ref.shape and torch.isfinite(out).all()
    rel = (out - ref).norm() / ref.norm()
    assert rel < 0.5, (preset, rel.item())
