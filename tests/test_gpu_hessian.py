"""GPU parity of the local-Hessian calibration kernels (csrc/hessian.cu) through the C ABI: the block Hessian against an
fp64 restatement and the reference's accumulator captured on a B200, the Hessian-weighted FP8-scale sweep against the
oracle (oracle/hessian.py), the reference's Triton kernel captured on a B200 and the generic error_func sweep, and
mtq.quantize with the local_hessian algorithm end to end."""

import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import modelopt_b200 as mtq  # noqa: E402
from modelopt_b200 import config as qcfg  # noqa: E402
from modelopt_b200 import ops  # noqa: E402
from modelopt_b200.calib import NVFP4MSECalibrator  # noqa: E402
from modelopt_b200.model_calib import _LocalHessianAccumulator  # noqa: E402
from oracle import hessian as oh  # noqa: E402

GOLD = os.path.join(os.path.dirname(__file__), "golden", "hessian_b200.npz")
GB = np.load(GOLD) if os.path.exists(GOLD) else None
DT = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}


def T(key, dtype=None):
    a = torch.from_numpy(GB[key])
    if dtype in (torch.bfloat16, torch.float16) and a.dtype == torch.int16:
        return a.view(dtype)
    return a


@pytest.mark.parametrize("dt", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("tokens,cin", [(1000, 256), (37, 4096), (16384, 1024), (5, 48)])
def test_block_hessian_vs_fp64(dt, tokens, cin):
    torch.manual_seed(0)
    x = (torch.randn(tokens, cin) * (torch.rand(cin) + 0.2)).to(dt)
    h = torch.zeros(cin // 16, 16, 16, device="cuda")
    ops.block_hessian_into(x.cuda(), h)
    want = oh.block_hessian(x.float(), 16)
    err = (h.double().cpu() - want).abs().max() / want.abs().max()
    assert err < 2e-6, err
    assert torch.equal(h, h.transpose(1, 2))  # exactly symmetric: same products, same order
    # accumulation over calls, 3-D activations, deterministic
    h2 = torch.zeros_like(h)
    ops.block_hessian_into(x.cuda(), h2)
    assert torch.equal(h, h2)
    ops.block_hessian_into(x.cuda().reshape(1, tokens, cin), h2)
    assert (h2.double().cpu() - 2 * want).abs().max() / want.abs().max() < 4e-6


@pytest.mark.parametrize("dn", list(DT))
def test_hessian_sweep_vs_oracle(dn):
    dt = DT[dn]
    torch.manual_seed(5)
    cout, cin = 96, 128
    x = torch.randn(500, cin) * (torch.rand(cin) * 3 + 0.1)
    h = (oh.block_hessian(x, 16) / 500).float()
    w = (torch.randn(cout, cin) * 0.03).to(dt)
    w[2] = 0
    ga = w.float().abs().amax()
    got = ops.nvfp4_fp8_scale_sweep_hessian(w.cuda().reshape(-1, 16), ga.cuda(), h.cuda(), numerics="cpu").cpu()
    want, losses = oh.hessian_sweep(w, ga, h, dtype=torch.float64)
    same = got == want
    assert same.float().mean() >= 0.99, same.float().mean()
    # where the choice differs it is a near-tie of the objective (fp32 summation order)
    cand = oh.fp8_scale_candidates() * ga
    for b in torch.nonzero(~same).flatten().tolist():
        k = int(torch.argmin((cand - got[b]).abs()))
        assert losses[k, b] <= losses[:, b].min() * (1 + 1e-4) + 1e-30


@pytest.mark.skipif(GB is None, reason="tests/golden/hessian_b200.npz not captured yet")
@pytest.mark.parametrize("dn", list(DT))
def test_vs_reference_on_b200(dn):
    """The reference's accumulator and Triton Hessian sweep, captured on a B200."""
    dt = DT[dn]
    cout, cin = 192, 256
    acc = _LocalHessianAccumulator(cout, cin, 16)
    acc.accumulate(T(f"x1_{dn}", dt).cuda())
    acc.accumulate(T(f"x2_{dn}", dt).cuda().reshape(7, 11, cin))
    h = acc.normalized_hessian()
    ref_h = T(f"h_{dn}").cuda()
    assert (h - ref_h).abs().max() <= 2e-5 * ref_h.abs().max()  # fp32 GEMM summation order
    w = T(f"w_{dn}", dt).cuda()
    best = ops.nvfp4_fp8_scale_sweep_hessian(w.reshape(-1, 16), T(f"ga_{dn}").cuda(), ref_h).cpu()
    ref = T(f"best_{dn}")
    assert (best == ref).float().mean() >= 0.99, (best == ref).float().mean()


def test_calibrator_paths_agree():
    """NVFP4MSECalibrator: the fused Hessian kernel picks what the generic 126-step error_func sweep picks."""
    torch.manual_seed(1)
    cout, cin = 64, 128
    w = (torch.randn(cout, cin, device="cuda") * 0.02).to(torch.bfloat16)
    acc = _LocalHessianAccumulator(cout, cin, 16)
    acc.accumulate(torch.randn(256, cin, device="cuda", dtype=torch.bfloat16) * 2)
    err = acc.build_error_func(keep_buffer=True)
    h = acc.normalized_hessian()
    blocks = w.reshape(-1, 16)
    bam = blocks.float().abs().amax(dim=-1)
    ga = bam.max()

    def quant_func(x, amax):
        return ops.qdq_nvfp4_static(x, amax, ga, True, 448.0)

    fused = NVFP4MSECalibrator(amax=bam, axis=0, global_amax=ga, quant_func=quant_func, error_func=err, hessian=h)
    fused.collect(blocks)
    generic = NVFP4MSECalibrator(amax=bam, axis=0, global_amax=ga, quant_func=quant_func, error_func=err)
    generic.collect(blocks)
    a, b = fused.compute_amax(), generic.compute_amax()
    assert (a == b).float().mean() >= 0.98, (a == b).float().mean()


def test_local_hessian_quantize_end_to_end():
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(256, 512), torch.nn.GELU(), torch.nn.Linear(512, 256)).cuda().to(torch.bfloat16)
    scale = (torch.rand(256, device="cuda") * 4 + 0.05).to(torch.bfloat16)  # anisotropic inputs: the Hessian matters
    batches = [torch.randn(64, 256, device="cuda", dtype=torch.bfloat16) * scale for _ in range(3)]

    def loop(m):
        for b in batches:
            m(b)

    import copy

    ref_out = model(batches[0]).float()
    m_max, m_hes = copy.deepcopy(model), copy.deepcopy(model)
    cfg_max = copy.deepcopy(qcfg.NVFP4_W4A4_WEIGHT_LOCAL_HESSIAN_CFG)
    cfg_max["algorithm"] = "max"
    mtq.quantize(m_max, cfg_max, loop)
    mtq.quantize(m_hes, qcfg.NVFP4_W4A4_WEIGHT_LOCAL_HESSIAN_CFG, loop)
    q = m_hes[0].weight_quantizer
    assert q.is_nvfp4_static and q._amax.numel() == 512 * 256 // 16
    assert not torch.equal(q._amax, m_max[0].weight_quantizer._amax)  # refined
    # the refined scales do not increase the layer's Hessian-weighted weight error (the objective they minimise)
    w = model[0].weight.data
    h = (oh.block_hessian(torch.cat(batches).float().cpu(), 16) / (3 * 64)).float().cuda()

    def objective(mod):
        dw = (w.float() - mod[0].weight_quantizer(w).float()).view(512, -1, 16)
        return torch.einsum("cnb,nbd,cnd->", dw, h, dw).item()

    assert objective(m_hes) <= objective(m_max) * 1.0001
    e_max = (m_max(batches[0]).float() - ref_out).norm().item()
    e_hes = (m_hes(batches[0]).float() - ref_out).norm().item()
    assert e_hes <= e_max * 1.1, (e_hes, e_max)
