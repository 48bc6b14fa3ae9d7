"""GPU parity: the block-scaled NVFP4 linear (tcgen05.mma.kind::mxf4nvf4) through the C ABI.

Oracle: NVFP4 codes / E4M3 scales / global scale from the oracle's restatement of NVFP4QTensor.quantize, dequantized
on the CPU and multiplied in fp64.  The instruction forms the products of the dequantized values exactly and
accumulates in fp32, so the only differences are fp32 accumulation order and the bf16 / f16 rounding of the output:
tolerance 2^-7 relative (bf16) / 2^-10 (f16) of max|out|, stated per assertion."""

import pytest
import torch

from oracle import nvfp4_pack as opack

pytestmark = pytest.mark.gpu


def _oracle_dequant(x):
    """x (CPU) -> fp64 dequantized NVFP4 value of every element (oracle/nvfp4_pack.py)."""
    amax = x.float().abs().max()
    codes, sf, sf2 = opack.quantize(x.float(), 16)
    return opack.dequantize(codes, sf, sf2, 16, torch.float32).double(), amax


@pytest.mark.parametrize("M,N,K", [(128, 256, 256), (200, 264, 512), (77, 8, 256), (1, 512, 1024), (640, 1024, 2048)])
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
def test_nvfp4_linear_vs_cpu_oracle(M, N, K, dt):
    from modelopt_b200 import ops

    torch.manual_seed(M + 3 * N + K)
    x = (torch.randn(M, K) * 2).to(dt)
    w = (torch.randn(N, K) * 0.05).to(dt)
    x[0, :16] = 0                      # an all-zero block
    b = torch.randn(N).to(dt)
    xd, ax = _oracle_dequant(x)
    wd, aw = _oracle_dequant(w)
    want = xd @ wd.t() + b.double()
    with ops.numerics("cpu"):
        xc, xs, xg = ops.nvfp4_quantize_operand(x.cuda(), ax.cuda(), 128)
        wc, ws, wg = ops.nvfp4_quantize_operand(w.cuda(), aw.cuda(), 256)
    tol = (2.0 ** -7 if dt == torch.bfloat16 else 2.0 ** -10) * want.abs().max().item()
    ys = {}
    for pair in (False, True):      # single-CTA kernel (128 x 256 tile) and CTA-pair kernel (256 x 256 pair tile)
        ys[pair] = ops.linear_nvfp4(xc, xs, xg, wc, ws, wg, b.cuda(), out_dtype=dt, pair=pair)
        assert (ys[pair].cpu().double() - want).abs().max().item() <= tol
    # exact products, the same K order, fp32 accumulation in tensor memory: the two kernels agree bit for bit
    assert torch.equal(ys[False], ys[True])


@pytest.mark.parametrize("M,N,K", [(1024, 1536, 1280), (2049, 520, 768), (4096, 4096, 4096)])
def test_nvfp4_pair_kernel_many_tiles(M, N, K):
    """More tiles than CTA pairs / several tiles per pair (accumulators alternate between the two overlapping TMEM
    regions, scale slots and the smem ring wrap many times): pair kernel == single-CTA kernel bit for bit, twice in a row,
    with the single-accumulator form (MO_LAB_F4P=1) agreeing as well."""
    import os

    from modelopt_b200 import ops

    torch.manual_seed(K)
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16) * 2
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) * 0.05
    b = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    a = ops.nvfp4_quantize_operand(x, x.float().abs().max(), 128)
    bb = ops.nvfp4_quantize_operand(w, w.float().abs().max(), 256)
    ref = ops.linear_nvfp4(*a, *bb, b, pair=False)
    y1 = ops.linear_nvfp4(*a, *bb, b, pair=True)
    y2 = ops.linear_nvfp4(*a, *bb, b, pair=True)
    os.environ["MO_LAB_F4P"] = "1"
    try:
        y3 = ops.linear_nvfp4(*a, *bb, b, pair=True)
    finally:
        os.environ["MO_LAB_F4P"] = "0"
    assert torch.equal(ref, y1) and torch.equal(y1, y2) and torch.equal(y1, y3)
    assert ops.NVFP4_GEMM["form"] == "auto"     # more than 128 rows: the pair kernel is what linear_nvfp4 picks


def test_nvfp4_linear_close_to_fake_quant_linear():
    """What the reference simulates (fake-quantized bf16 operands through a bf16 GEMM) against the real W4A4 product.
    Two sources of difference: the bf16 rounding of the dequantized operands (2^-9 per term) and elements that sit on an
    E2M1 rounding boundary, where the fake-quant kernel's div.full.f32 (what the reference's Triton kernel executes) and
    the packer's IEEE division (NVFP4QTensor.quantize) pick neighbouring codes.  Tolerance: 6 % of max|out|, and the
    mean error must stay below 0.5 %."""
    from modelopt_b200 import ops

    torch.manual_seed(0)
    x = torch.randn(512, 1024, device="cuda", dtype=torch.bfloat16) * 2
    w = torch.randn(768, 1024, device="cuda", dtype=torch.bfloat16) * 0.05
    ax, aw = x.float().abs().max(), w.float().abs().max()
    y = ops.linear_nvfp4(*ops.nvfp4_quantize_operand(x, ax, 128), *ops.nvfp4_quantize_operand(w, aw, 256))
    fq = torch.nn.functional.linear(ops.qdq_nvfp4_dynamic(x, ax, 16), ops.qdq_nvfp4_dynamic(w, aw, 16))
    d = (y.float() - fq.float()).abs()
    top = fq.float().abs().max().item()
    assert d.max().item() <= 0.06 * top and d.mean().item() <= 0.005 * top


def test_nvfp4_linear_argument_errors_and_swizzle_layout():
    from modelopt_b200 import ops

    sf = torch.arange(130 * 8, dtype=torch.uint8, device="cuda").reshape(130, 8)
    sw = ops.nvfp4_sf_swizzle(sf, 128).cpu()
    assert sw.numel() == 2 * 2 * 512          # rows padded to 256, 8 k-blocks = 2 tiles of 4
    for r, k in ((0, 0), (5, 3), (33, 4), (127, 7), (129, 2)):
        off = ((r // 128) * 2 + k // 4) * 512 + (r % 32) * 16 + ((r % 128) // 32) * 4 + k % 4
        assert sw[off].item() == sf[r, k].item()
    assert sw[(1 * 2 + 0) * 512 + 2 * 16].item() == 0   # padded row 130
    x = torch.randn(64, 128, device="cuda", dtype=torch.bfloat16)
    a = ops.nvfp4_quantize_operand(x, x.abs().max(), 128)
    with pytest.raises(ValueError):
        ops.linear_nvfp4(*a, *ops.nvfp4_quantize_operand(x, x.abs().max(), 256))   # K = 128 is not a multiple of 256


def test_quant_linear_real_nvfp4_gemm_opt_in():
    """NVFP4_DEFAULT_CFG module with real_quant_gemm = True: pack x, cached packed W, block-scaled GEMM; close to the
    fake-quant forward (same tolerance as above) and 3 kernel launches per linear after the first call."""
    import copy

    import modelopt_b200 as mtq
    from modelopt_b200 import ops
    from torch import nn

    torch.manual_seed(3)
    m = nn.Sequential(nn.Linear(512, 1024), nn.GELU(), nn.Linear(1024, 256, bias=False)).to("cuda", torch.bfloat16)
    data = [torch.randn(4, 64, 512, device="cuda", dtype=torch.bfloat16) for _ in range(2)]
    mtq.quantize(m, copy.deepcopy(mtq.NVFP4_DEFAULT_CFG), lambda model: [model(b) for b in data])
    with torch.no_grad():
        fake = m(data[0])
        for mod in m.modules():
            if hasattr(mod, "real_quant_gemm"):
                mod.real_quant_gemm = True
        real = m(data[0])
        n0 = ops.LAUNCHES[0]
        real2 = m(data[0])
        assert ops.LAUNCHES[0] - n0 == 4      # per linear: pack x (scales written swizzled) + GEMM; weights cached
    assert torch.equal(real, real2) and real.shape == fake.shape
    d = (real.float() - fake.float()).abs()
    top = fake.float().abs().max().item()
    assert d.max().item() <= 0.08 * top and d.mean().item() <= 0.01 * top
