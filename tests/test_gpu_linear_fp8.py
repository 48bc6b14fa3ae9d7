"""GPU checks of the FP8 per-tensor real-quant linear (csrc/gemm_fp8.cu, tcgen05.mma.kind::f8f6f4) through the C ABI:
against the fp64 product of the dequantized operands (tolerance: the bf16 rounding of the output), against
torch._scaled_mm on the same E4M3 operands (the reference backend's GEMM, backends/fp8_per_tensor_gemm.py:36-50), the
`_to_fp8` packer against its torch formula, and the opt-in module path against the fake-quant forward."""

import pytest
import torch

pytestmark = pytest.mark.gpu

import modelopt_b200 as mtq  # noqa: E402
from modelopt_b200 import config as qcfg  # noqa: E402
from modelopt_b200 import ops  # noqa: E402


def _operands(M, N, K, seed=0):
    torch.manual_seed(seed)
    x = (torch.randn(M, K, device="cuda") * 2).to(torch.bfloat16)
    w = (torch.randn(N, K, device="cuda") * 0.05).to(torch.bfloat16)
    sa = (x.float().abs().amax() / 448.0).reshape(1)
    sb = (w.float().abs().amax() / 448.0).reshape(1)
    return x, w, sa, sb


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
def test_to_fp8_pack_matches_formula(dt):
    torch.manual_seed(1)
    x = (torch.randn(333, 256, device="cuda") * 3).to(dt)
    x[0, 0] = 1000.0  # beyond the calibrated amax: clamps to 448
    scale = torch.tensor([0.0123], device="cuda")
    got = ops.q8_pack(x, scale, "fp8_sat")
    want = (x.float() / scale).clamp(-448.0, 448.0).to(torch.float8_e4m3fn)
    assert torch.equal(got.view(torch.uint8), want.view(torch.uint8))


@pytest.mark.parametrize("M,N,K", [(128, 256, 128), (300, 264, 144), (1, 8, 16), (2048, 4096, 4096), (777, 1024, 14336)])
@pytest.mark.parametrize("bias", [False, True])
def test_linear_fp8_vs_fp64(M, N, K, bias):
    x, w, sa, sb = _operands(M, N, K)
    a = ops.q8_pack(x, sa, "fp8_sat")
    b = ops.q8_pack(w, sb, "fp8_sat")
    bv = (torch.randn(N, device="cuda") * 0.1).to(torch.bfloat16) if bias else None
    out = ops.linear_fp8(a, b, sa, sb, bv)
    want = (a.double() @ b.double().t()) * (sa.double() * sb.double())
    if bias:
        want = want + bv.double()
    err = (out.double() - want).abs().max().item()
    assert err <= 2.0 ** -7 * want.abs().max().item() + 1e-6, err
    # the library GEMM the reference backend calls, on the same operands
    ref = torch._scaled_mm(a, b.t(), scale_a=sa, scale_b=sb, bias=bv, out_dtype=torch.bfloat16, use_fast_accum=True) \
        if K % 16 == 0 and N % 16 == 0 and M > 1 else None
    if ref is not None:
        assert (out.float() - ref.float()).abs().max().item() <= 2.0 ** -7 * want.abs().max().item() + 1e-6


def test_errors():
    x, w, sa, sb = _operands(16, 16, 32)
    a, b = ops.q8_pack(x, sa, "fp8_sat"), ops.q8_pack(w, sb, "fp8_sat")
    with pytest.raises(Exception):
        ops.linear_fp8(a[:, :24].contiguous(), b[:, :24].contiguous(), sa, sb)  # K % 16 != 0


def test_module_real_quant_fp8_path():
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(512, 1024), torch.nn.GELU(), torch.nn.Linear(1024, 256)).cuda().to(torch.bfloat16)
    x = torch.randn(200, 512, device="cuda", dtype=torch.bfloat16)
    mtq.quantize(model, qcfg.FP8_DEFAULT_CFG, lambda m: m(x))
    with torch.no_grad():
        fake = model(x)
        for lin in (model[0], model[2]):
            lin.real_quant_gemm = True
        before = ops.LAUNCHES[0]
        real = model(x)
        launches = ops.LAUNCHES[0] - before
    assert launches == 2 * 2 + 2  # per linear: input pack + GEMM; weights packed once (2)
    rel = (real.float() - fake.float()).norm() / fake.float().norm()
    assert rel < 0.02, rel
    with torch.no_grad():
        before = ops.LAUNCHES[0]
        model(x)
        assert ops.LAUNCHES[0] - before == 4  # cached weight codes


def test_gemm_registry_dispatch():
    """enable_real_quant_gemm + gemm_registry (backends/gemm_registry.py): default FP8 backend, a user backend registered
    in front of it, and disable."""
    from modelopt_b200.backends import GEMMRegistry, gemm_registry

    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(256, 512)).cuda().to(torch.bfloat16)
    x = torch.randn(64, 256, device="cuda", dtype=torch.bfloat16)
    mtq.quantize(model, qcfg.FP8_DEFAULT_CFG, lambda m: m(x))
    with torch.no_grad():
        fake = model(x)
        assert not mtq.is_real_quant_gemm_enabled(model)
        mtq.enable_real_quant_gemm(model)
        assert mtq.is_real_quant_gemm_enabled(model) and model[0].real_quant_gemm
        real = model(x)
        assert (real.float() - fake.float()).norm() / fake.float().norm() < 0.02
        calls = []

        def user_gemm(module, input, weight, bias):
            calls.append(1)
            return torch.nn.functional.linear(input, weight, bias)

        reg = gemm_registry._registry
        gemm_registry._registry = []
        try:
            gemm_registry.register(user_gemm, None)
            assert user_gemm in gemm_registry
            model(x)
            assert calls == [1]
            gemm_registry.unregister(user_gemm)
            assert user_gemm not in gemm_registry
        finally:
            gemm_registry._registry = reg
        mtq.disable_real_quant_gemm(model)
        assert torch.equal(model(x), fake)
    r = GEMMRegistry()
    assert r.find_match(model[0], x) is None


@pytest.mark.parametrize("M,N,K", [(256, 256, 128), (300, 264, 144), (1, 8, 16), (1024, 512, 1024), (4096, 4096, 4096)])
def test_cta_pair_kernel_equals_single_cta(M, N, K):
    """mo_linear_fp8_pair (csrc/gemm_2cta.cu with kind::f8f6f4, cta_group::2) == csrc/gemm_fp8.cu bit for bit."""
    x, w, sa, sb = _operands(M, N, K)
    a, b = ops.q8_pack(x, sa, "fp8_sat"), ops.q8_pack(w, sb, "fp8_sat")
    bv = (torch.randn(N, device="cuda") * 0.1).to(torch.bfloat16)
    assert torch.equal(ops.linear_fp8(a, b, sa, sb, bv, pair=False), ops.linear_fp8(a, b, sa, sb, bv, pair=True))
