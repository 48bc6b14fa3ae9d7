"""GPU parity: NVFP4 packer / unpacker (through the C ABI) vs the REAL reference's NVFP4QTensor outputs
(tests/golden/nvfp4_pack_cpu.npz) and the oracle.  Codes, E4M3 scale bytes and the global scale: bit-exact."""

import pytest
import torch

from conftest import DTYPES, bits_equal, from_np, load_golden
from oracle import nvfp4_pack as opack

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def Q():
    from modelopt_b200.qtensor import NVFP4QTensor

    return NVFP4QTensor


@pytest.mark.parametrize("dn", list(DTYPES))
def test_golden_pack_and_dequant(Q, dn):
    g = load_golden("nvfp4_pack_cpu.npz")
    dt = DTYPES[dn]
    from modelopt_b200 import ops

    x = from_np(g[f"x_{dn}"], dt).cuda()
    with ops.numerics("cpu"):
        qt, sf, sf2 = Q.quantize(x.clone(), 16)
    assert torch.equal(sf2.cpu(), torch.from_numpy(g[f"sf2_{dn}"]))
    assert torch.equal(sf.cpu().view(torch.uint8), torch.from_numpy(g[f"sf_{dn}"]))
    assert torch.equal(qt._quantized_data.cpu(), torch.from_numpy(g[f"codes_{dn}"]))
    d = qt.dequantize(dtype=dt, scale=sf, double_scale=sf2, block_sizes={-1: 16})
    assert bits_equal(d.cpu(), from_np(g[f"deq_{dn}"], dt))
    # externally supplied global scale
    sf2b = torch.from_numpy(g[f"sf2_ext_{dn}"]).cuda()
    qt, sf, _ = Q.quantize(x.clone(), 16, weights_scaling_factor_2=sf2b)
    assert torch.equal(sf.cpu().view(torch.uint8), torch.from_numpy(g[f"sf_ext_{dn}"]))
    assert torch.equal(qt._quantized_data.cpu(), torch.from_numpy(g[f"codes_ext_{dn}"]))
    # ragged last dim: padded to the block
    xr = x[:, :184].contiguous()
    with ops.numerics("cpu"):
        qt, sf, sf2 = Q.quantize(xr.clone(), 16)
    assert torch.equal(sf2.cpu(), torch.from_numpy(g[f"sf2_rag_{dn}"]))
    assert torch.equal(sf.cpu().view(torch.uint8), torch.from_numpy(g[f"sf_rag_{dn}"]))
    assert torch.equal(qt._quantized_data.cpu(), torch.from_numpy(g[f"codes_rag_{dn}"]))
    assert qt.dequantize(scale=sf, double_scale=sf2, block_sizes={-1: 16}).shape == xr.shape


@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("shape,block", [((1, 16), 16), ((64, 4096), 16), ((3, 5, 64), 16), ((8, 256), 32),
                                         ((16, 96), 8), ((7, 40), 16)])
def test_pack_vs_oracle(Q, dn, shape, block):
    dt = DTYPES[dn]
    g = torch.Generator().manual_seed(shape[-1] + block)
    x = (torch.randn(shape, generator=g) * 3).to(dt)
    x.view(-1)[::37] *= 40
    x.view(-1)[5] = 0.0
    x.view(-1)[6] = -0.0
    if shape[-1] >= 32:
        x.view(-1, shape[-1])[0, :block] = 0          # zero block -> scale 1.0
        x.view(-1, shape[-1])[0, block:2 * block] *= 1e-6  # tiny block -> scale clamps to 2^-9
    from modelopt_b200 import ops

    codes_o, sf_o, sf2_o = opack.quantize(x.clone(), block)
    with ops.numerics("cpu"):  # the oracle divides amax by 6*448 on CPU tensors (true division)
        qt, sf, sf2 = Q.quantize(x.cuda(), block)
    assert torch.equal(sf2.cpu(), sf2_o)
    assert torch.equal(sf.cpu().view(torch.uint8), sf_o.view(torch.uint8))
    assert torch.equal(qt._quantized_data.cpu(), codes_o)
    for fast in (False, True):
        d = qt.dequantize(dtype=dt, fast=fast, scale=sf, double_scale=sf2, block_sizes={-1: block})
        want = opack.dequantize(codes_o, sf_o, sf2_o, block, dt, shape=x.shape, fast=fast)
        assert bits_equal(d.cpu(), want), fast


def test_exact_ties_and_given_scales(Q):
    ties = torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5.0, 6.0, -0.25, -0.75, -1.25, -1.75, -2.5, -3.5, -5.0, -6.0])
    x = ties.repeat(4, 1)
    sf2 = torch.tensor(6.0 / (6.0 * 448.0))
    codes_o, sf_o, _ = opack.quantize(x.clone(), 16, sf2=sf2)
    qt, sf, _ = Q.quantize(x.cuda(), 16, weights_scaling_factor_2=sf2.cuda())
    assert torch.equal(qt._quantized_data.cpu(), codes_o) and torch.equal(sf.cpu().view(torch.uint8), sf_o.view(torch.uint8))
    # given block scales
    given = torch.full((4, 1), 2.0).to(torch.float8_e4m3fn)
    codes_o, _, _ = opack.quantize(x.clone(), 16, sf=given, sf2=sf2)
    qt, sf, _ = Q.quantize(x.cuda(), 16, weights_scaling_factor=given.cuda(), weights_scaling_factor_2=sf2.cuda())
    assert torch.equal(qt._quantized_data.cpu(), codes_o)


def test_roundtrip_property_full_size(Q):
    """Llama-3-8B gate_proj [14336, 4096] bf16: pack -> unpack equals the static fake-quant of the same scales, and
    re-packing the dequantized tensor with the same scales reproduces the codes (idempotence)."""
    g = torch.Generator(device="cuda").manual_seed(3)
    w = torch.randn((14336, 4096), generator=g, device="cuda", dtype=torch.bfloat16) * 0.02
    qt, sf, sf2 = Q.quantize(w, 16)
    d = qt.dequantize(scale=sf, double_scale=sf2, block_sizes={-1: 16})
    qt2, _, _ = Q.quantize(d, 16, weights_scaling_factor=sf, weights_scaling_factor_2=sf2)
    a, b = qt._quantized_data, qt2._quantized_data
    # codes agree except +-0 encodings (a negative value that rounds to zero is -0 once, +0 after the round trip)
    diff = a != b
    assert int(diff.sum()) == int((((a ^ b) & 0x77) == 0)[diff].sum())
    rows = torch.arange(0, 14336, 997)
    co, so, s2o = opack.quantize(w[rows].cpu(), 16, sf2=sf2.cpu())
    assert torch.equal(a[rows].cpu(), co) and torch.equal(sf[rows].cpu().view(torch.uint8), so.view(torch.uint8))
