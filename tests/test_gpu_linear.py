"""GPU parity: the tcgen05 quant-aware linear (through the C ABI) vs a plain PyTorch fp32 reference of the same op
(tolerance: bf16 / f16 output rounding + fp32 accumulation order, stated per test) and vs cuBLAS on the same operands."""

import copy

import pytest
import torch
import torch.nn.functional as F
from torch import nn

pytestmark = pytest.mark.gpu

SHAPES = [(128, 256, 64), (128, 256, 256), (256, 512, 128), (1000, 264, 72), (77, 8, 8), (1, 4096, 1024),
          (300, 520, 4096), (2048, 1024, 2048)]


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("bias", [True, False])
def test_linear_vs_fp32_reference(M, N, K, dt, bias):
    from modelopt_b200 import ops

    torch.manual_seed(M * 3 + N * 5 + K)
    x = torch.randn(M, K, device="cuda", dtype=dt)
    w = torch.randn(N, K, device="cuda", dtype=dt) * 0.05
    b = torch.randn(N, device="cuda", dtype=dt) if bias else None
    y = ops.linear(x, w, b)
    want = x.float() @ w.float().t() + (b.float() if bias else 0.0)
    # one output ulp (2^-8 bf16 / 2^-11 f16, relative) + fp32 accumulation-order noise
    ulp = 2.0 ** -8 if dt == torch.bfloat16 else 2.0 ** -11
    torch.testing.assert_close(y.float(), want, rtol=ulp, atol=ulp * want.abs().max().item() * 0.25)
    ref = F.linear(x, w, b)  # cuBLAS, same operands, also fp32 accumulation
    assert (y != ref).float().mean().item() <= 0.02
    torch.testing.assert_close(y.float(), ref.float(), rtol=2 * ulp, atol=ulp * want.abs().max().item() * 0.25)


def test_linear_batched_input_and_errors():
    from modelopt_b200 import ops

    x = torch.randn(3, 50, 128, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(64, 128, device="cuda", dtype=torch.bfloat16)
    y = ops.linear(x, w)
    assert y.shape == (3, 50, 64)
    torch.testing.assert_close(y.float(), (x.float() @ w.float().t()), rtol=2.0 ** -7, atol=0.1)
    with pytest.raises(ValueError):
        ops.linear(x.float(), w.float())
    with pytest.raises(ValueError):
        ops.linear(x[..., :100].contiguous(), w[:, :100].contiguous())  # K % 8 != 0
    with pytest.raises(RuntimeError):
        ops.linear(x.cpu(), w.cpu())
    assert ops.linear(x[:0], w).shape == (0, 50, 64)


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
def test_linear_fused_loss(dt):
    from modelopt_b200 import ops

    torch.manual_seed(5)
    x = torch.randn(700, 512, device="cuda", dtype=dt)
    w = torch.randn(264, 512, device="cuda", dtype=dt) * 0.05
    b = torch.randn(264, device="cuda", dtype=dt)
    ref = (F.linear(x, w, b).float() * 1.02 + 0.01).to(dt)
    acc = torch.zeros(3, dtype=torch.float64, device="cuda")
    ops.linear_loss_into(x, w, b, ref, acc[1:2])
    ops.linear_loss_into(x, w, b, ref, acc[1:2])
    y = ops.linear(x, w, b)
    want = 2 * (y - ref).float().pow(2).double().mean().item()
    assert acc[0].item() == 0 and acc[2].item() == 0
    assert abs(acc[1].item() - want) <= 1e-6 * want  # fp32 per-thread partial sums


def _quant_linear(cfg_name, dt=torch.bfloat16):
    import modelopt_b200 as mtq

    torch.manual_seed(3)
    m = nn.Sequential(nn.Linear(512, 1024), nn.GELU(), nn.Linear(1024, 256, bias=False)).to("cuda", dt)
    data = [torch.randn(4, 64, 512, device="cuda", dtype=dt) for _ in range(2)]
    mtq.quantize(m, copy.deepcopy(getattr(mtq, cfg_name)), lambda model: [model(b) for b in data])
    return m, data


@pytest.mark.parametrize("cfg", ["INT8_DEFAULT_CFG", "FP8_DEFAULT_CFG", "NVFP4_DEFAULT_CFG",
                                 "INT4_BLOCKWISE_WEIGHT_ONLY_CFG"])
def test_quant_linear_forward_matches_three_pass_path(cfg):
    """no_grad forward (tcgen05, cached weight) vs the reference's structure (QDQ x, QDQ W, F.linear) on the same
    module: the operands are bit-identical, the GEMMs differ by fp32 accumulation order only."""
    from modelopt_b200 import ops

    m, data = _quant_linear(cfg)
    x = data[0]
    with torch.no_grad():
        n0 = ops.LAUNCHES[0]
        y1 = m(x)
        first = ops.LAUNCHES[0] - n0
        n0 = ops.LAUNCHES[0]
        y2 = m(x)
        second = ops.LAUNCHES[0] - n0
    assert torch.equal(y1, y2)
    assert second < first, (first, second)  # the weights were not re-quantized
    yg = m(x.clone().requires_grad_(True)).detach()  # gradient needed -> cuBLAS path, weights re-quantized
    assert (y1 != yg).float().mean().item() <= 0.02
    torch.testing.assert_close(y1.float(), yg.float(), rtol=2.0 ** -6, atol=2.0 ** -6 * yg.float().abs().max().item())


def test_quantized_weight_cache_invalidation():
    m, data = _quant_linear("INT8_DEFAULT_CFG")
    lin = m[0]
    x = data[0]
    with torch.no_grad():
        y0 = m(x)
        w_cached = lin.__dict__["_wq_cache"][1]
        lin.weight_quantizer.amax = lin.weight_quantizer.amax * 0.5   # setter -> buffer version changes
        y1 = m(x)
        assert lin.__dict__["_wq_cache"][1] is not w_cached and not torch.equal(y0, y1)
        w_cached = lin.__dict__["_wq_cache"][1]
        lin.weight.mul_(1.5)                                          # in-place weight update
        y2 = m(x)
        assert lin.__dict__["_wq_cache"][1] is not w_cached and not torch.equal(y1, y2)
        lin.weight_quantizer.disable()
        y3 = m(x)
        want = F.linear(lin.input_quantizer(x), lin.weight, lin.bias)
        torch.testing.assert_close(m[2](m[1](want)).float(), y3.float(), rtol=2.0 ** -6, atol=1e-2)


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (1000, 264, 72), (77, 520, 1024)])
def test_linear_with_fused_input_quantizer_is_bit_identical_to_two_kernels(dt, M, N, K):
    from modelopt_b200 import ops

    torch.manual_seed(M + N + K + 1)
    x = torch.randn(M, K, device="cuda", dtype=dt) * 2
    w = torch.randn(N, K, device="cuda", dtype=dt) * 0.05
    b = torch.randn(N, device="cuda", dtype=dt)
    pqs = (torch.rand(K, device="cuda") + 0.5).to(dt)
    am = x.float().abs().max() * 0.7
    for fmt in ((4, 3), 8, 4):
        for ps in (None, pqs):
            xq = ops.qdq_fp8(x, am, prescale=ps) if fmt == (4, 3) else ops.qdq_int(x, am, fmt, False, False, prescale=ps)
            assert torch.equal(ops.linear_qdq(x, w, b, am, fmt, False, prescale=ps), ops.linear(xq, w, b)), (fmt, ps is None)


def test_quant_linear_small_batch_uses_one_kernel():
    from modelopt_b200 import ops

    m, data = _quant_linear("FP8_DEFAULT_CFG")
    x = data[0][:1, :32]            # 32 rows <= FUSED_PROLOGUE["max_rows"]
    with torch.no_grad():
        m(x)                         # fills the weight caches
        n0 = ops.LAUNCHES[0]
        y = m(x)
        assert ops.LAUNCHES[0] - n0 == 2          # one fused kernel per linear
        old, old_gemm = dict(ops.FUSED_PROLOGUE), dict(ops.PLAIN_GEMM)
        ops.FUSED_PROLOGUE["max_rows"] = 0
        try:
            ops.PLAIN_GEMM["backend"] = "tcgen05"
            n0 = ops.LAUNCHES[0]
            y2 = m(x)
            assert ops.LAUNCHES[0] - n0 == 4      # input QDQ + this library's GEMM per linear
            ops.PLAIN_GEMM["backend"] = "cublas"
            n0 = ops.LAUNCHES[0]
            y3 = m(x)
            assert ops.LAUNCHES[0] - n0 == 2      # input QDQ per linear; the plain product runs in cuBLAS
        finally:
            ops.FUSED_PROLOGUE.update(old)
            ops.PLAIN_GEMM.update(old_gemm)
    assert torch.equal(y, y2)
    assert (y.float() - y3.float()).abs().max().item() <= 2.0 ** -7 * y.float().abs().max().item()


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (300, 264, 144), (1, 8, 8), (1024, 512, 1024), (4096, 4096, 4096)])
def test_cta_pair_kernel_equals_single_cta(dt, M, N, K):
    """csrc/gemm_2cta.cu (tcgen05 cta_group::2, 256 x 256 tile per 2-CTA cluster) == csrc/gemm.cu bit for bit, plain and
    loss modes; ragged M / N / K exercise the TMA zero fill of a CTA whose half tile is partly or wholly out of range."""
    from modelopt_b200 import ops

    torch.manual_seed(0)
    x = torch.randn(M, K, device="cuda", dtype=dt)
    w = (torch.randn(N, K, device="cuda") * 0.05).to(dt)
    b = (torch.randn(N, device="cuda") * 0.1).to(dt)
    one, two = ops.linear(x, w, b, pair=False), ops.linear(x, w, b, pair=True)
    assert torch.equal(one, two)
    ref = torch.randn(M, N, device="cuda", dtype=dt)
    l1 = torch.zeros(1, dtype=torch.float64, device="cuda")
    l2 = torch.zeros(1, dtype=torch.float64, device="cuda")
    ops.linear_loss_into(x, w, b, ref, l1, pair=False)
    ops.linear_loss_into(x, w, b, ref, l2, pair=True)
    assert abs(l1.item() - l2.item()) <= 1e-12 * max(1.0, abs(l1.item()))  # fp64 atomics: order only
    want = (x.float() @ w.float().t() + b.float())
    assert (two.float() - want).abs().max() <= 2.0 ** -7 * want.abs().max() + 1e-6


def test_unaligned_views_fall_back_to_the_eager_linear():
    """A contiguous weight / input view whose storage offset is not 16-byte aligned (flat-parameter slices) is not
    offered to the tcgen05 kernels (`linear_supported` is False) and the module forward still works."""
    import modelopt_b200 as mtq
    from modelopt_b200 import config as qcfg
    from modelopt_b200 import ops

    torch.manual_seed(0)
    flat = (torch.randn(64 * 128 + 3, device="cuda") * 0.05).to(torch.bfloat16)
    w = flat[3:].view(64, 128)                      # 6-byte offset
    x = torch.randn(32, 128, device="cuda", dtype=torch.bfloat16)
    assert w.is_contiguous() and w.data_ptr() % 16 != 0
    assert not ops.linear_supported(x, w)
    assert ops.linear_supported(x, w.clone())
    lin = torch.nn.Linear(128, 64, bias=False, device="cuda", dtype=torch.bfloat16)
    lin.weight = torch.nn.Parameter(w)
    m = torch.nn.Sequential(lin)
    mtq.quantize(m, qcfg.FP8_DEFAULT_CFG, lambda mod: mod(x))
    with torch.no_grad():
        y = m(x)
    ref = torch.nn.Linear(128, 64, bias=False, device="cuda", dtype=torch.bfloat16)
    ref.weight = torch.nn.Parameter(w.clone())
    m2 = torch.nn.Sequential(ref)
    mtq.quantize(m2, qcfg.FP8_DEFAULT_CFG, lambda mod: mod(x))
    with torch.no_grad():
        y2 = m2(x)
    assert (y.float() - y2.float()).abs().max().item() <= 2.0 ** -7 * y2.float().abs().max().item()


def test_weight_cache_invalidation_and_opt_out():
    """A write through `.data` leaves (data_ptr, _version) unchanged: the cached fake-quantized weight is stale until
    `invalidate_weight_cache()`; MODELOPT_B200_NO_CACHE (ops.NO_CACHE) restores the reference's always-requantize
    behaviour, and the debug checksum (quant_module.CHECK_CACHE) turns the stale hit into an error."""
    from modelopt_b200 import ops
    from modelopt_b200.nn import quant_module as qm

    m, data = _quant_linear("FP8_DEFAULT_CFG")
    x = data[0]
    lin = m[0]
    with torch.no_grad():
        y0 = m(x)
        lin.weight_quantizer.reset_amax()  # dynamic amax: the fake-quantized weight follows the weight
        y0 = m(x)
        lin._parameters["weight"].data.mul_(2.0)          # no version bump
        stale = m(x)
        lin.invalidate_weight_cache()
        fresh = m(x)
        assert not torch.equal(stale, fresh)
        lin._parameters["weight"].data.mul_(0.5)
        old = ops.NO_CACHE
        ops.NO_CACHE = True
        try:
            assert torch.equal(m(x), y0)                  # no cache: the `.data` write is seen at once
        finally:
            ops.NO_CACHE = old
        lin.invalidate_weight_cache()
        qm.CHECK_CACHE = True
        try:
            m(x)                                          # fills the cache with a checksum
            lin._parameters["weight"].data.mul_(2.0)
            with pytest.raises(RuntimeError, match="version counter"):
                m(x)
        finally:
            qm.CHECK_CACHE = False
            lin.invalidate_weight_cache()
