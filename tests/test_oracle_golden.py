"""CPU: the oracle restatements reproduce the REAL reference's outputs (tests/golden/*.npz,
written by oracle/gen_golden.py from /root/reference) bit for bit.  This is what pins the oracle."""

import pytest
import torch

from conftest import DTYPES, bits_equal, from_np, load_golden
from oracle import amax as oamax
from oracle import fake_quant as ofq
from oracle import nvfp4_pack as opack


@pytest.fixture(scope="module")
def gq():
    return load_golden("qdq_cpu.npz")


@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("bits", [3, 4, 5, 7, 8, 11])
@pytest.mark.parametrize("unsigned", [False, True])
@pytest.mark.parametrize("narrow", [False, True])
def test_int_per_tensor(gq, dn, bits, unsigned, narrow):
    dt = DTYPES[dn]
    x = from_np(gq[f"x_{dn}"], dt)
    x = x.abs() if unsigned else x
    k = f"int_{dn}_b{bits}_u{int(unsigned)}_n{int(narrow)}"
    amax = torch.from_numpy(gq[k + "_amax"])
    y = ofq.int_qdq(x, amax, bits, unsigned, narrow)
    assert bits_equal(y, from_np(gq[k + "_y"], dt))


@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("case", ["axis0", "axis1", "blk128"])
def test_int_per_axis(gq, dn, case):
    dt = DTYPES[dn]
    x = from_np(gq[f"x_{dn}"], dt)
    bits = 8
    if case == "blk128":
        x, bits = x.reshape(-1, 128), 4
    amax = torch.from_numpy(gq[f"int_{dn}_{case}_amax"])
    y = ofq.int_qdq(x, amax, bits, False, False)
    assert bits_equal(y, from_np(gq[f"int_{dn}_{case}_y"], dt))


@pytest.mark.parametrize("dn", list(DTYPES))
def test_int_tiny_amax_conventions(gq, dn):
    dt = DTYPES[dn]
    x = from_np(gq[f"x_{dn}"], dt)
    for nm, am in (("tiny", 1e-26), ("zero", 0.0), ("eps2", 1.5 / (1 << 24))):
        y = ofq.int_qdq(x, torch.tensor(am), 8, False, False)
        assert bits_equal(y, from_np(gq[f"int_{dn}_{nm}_y"], dt)), nm
    # amax == 2^-24 is the one point where the CPU reference (<=) and the CUDA kernel (<) differ
    eps = torch.tensor(1.0 / (1 << 24))
    y_cpu = ofq.int_qdq(x, eps, 8, False, False, cuda_semantics=False)
    assert bits_equal(y_cpu, from_np(gq[f"int_{dn}_eps_y"], dt))
    assert torch.count_nonzero(y_cpu) == 0
    assert torch.count_nonzero(ofq.int_qdq(x, eps, 8, False, False, cuda_semantics=True)) > 0


@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("case", ["max", "clip", "zero", "small", "none", "axis0", "axis1", "axis2"])
def test_fp8(gq, dn, case):
    dt = DTYPES[dn]
    x = from_np(gq[f"x_{dn}"], dt)
    if case.startswith("axis"):
        x = x.reshape(6, 8, 256)
    amax = None if case == "none" else torch.from_numpy(gq[f"fp8_{dn}_{case}_amax"])
    y = ofq.fp8_qdq(x, amax)
    assert bits_equal(y, from_np(gq[f"fp8_{dn}_{case}_y"], dt))


def test_compute_fp4_scales():
    g = load_golden("fp4_scales_cpu.npz")
    amax = torch.from_numpy(g["amax"])
    for nm in ("auto", "given", "small"):
        ga = None if nm == "auto" else torch.from_numpy(g[f"ga_{nm}"])
        for q in (True, False):
            for m in (448.0, 256.0):
                s = ofq.compute_fp4_scales(amax, ga, q, m)
                assert bits_equal(s, torch.from_numpy(g[f"scale_{nm}_q{int(q)}_m{int(m)}"])), (nm, q, m)


@pytest.mark.parametrize("dn", list(DTYPES))
def test_amax_and_max_calibrator(dn):
    g = load_golden("amax_cpu.npz")
    dt = DTYPES[dn]
    xs = [from_np(g[f"x{i}_{dn}"], dt) for i in range(3)]
    assert bits_equal(oamax.reduce_amax(xs[0]).float(), torch.from_numpy(g[f"amax_all_{dn}"]))
    for axis, an in ((0, "0"), (1, "1"), (2, "2"), (-1, "-1"), ((0, 2), "0_2"), (None, None)):
        c = oamax.MaxCalibrator(8, axis, False)
        for x in xs:
            c.collect(x)
        key = f"calib_none_{dn}" if axis is None else f"calib_axis{an}_{dn}"
        assert bits_equal(c.compute_amax().float(), torch.from_numpy(g[key])), axis
    xb = xs[1].reshape(-1, 96)
    assert bits_equal(oamax.reduce_block_amax(xb, {-1: 16}).float(), torch.from_numpy(g[f"blk16_{dn}"]))
    assert bits_equal(oamax.reduce_block_amax(xb, {-1: 32}).float(), torch.from_numpy(g[f"blk32_{dn}"]))


@pytest.mark.parametrize("dn", list(DTYPES))
def test_nvfp4_pack(dn):
    g = load_golden("nvfp4_pack_cpu.npz")
    dt = DTYPES[dn]
    x = from_np(g[f"x_{dn}"], dt)
    codes, sf, sf2 = opack.quantize(x.clone(), 16)
    assert torch.equal(codes, torch.from_numpy(g[f"codes_{dn}"]))
    assert torch.equal(sf.view(torch.uint8), torch.from_numpy(g[f"sf_{dn}"]))
    assert torch.equal(sf2, torch.from_numpy(g[f"sf2_{dn}"]))
    d = opack.dequantize(codes, sf, sf2, 16, dt)
    assert bits_equal(d, from_np(g[f"deq_{dn}"], dt))
    codes, sf, _ = opack.quantize(x.clone(), 16, sf2=torch.from_numpy(g[f"sf2_ext_{dn}"]))
    assert torch.equal(codes, torch.from_numpy(g[f"codes_ext_{dn}"]))
    assert torch.equal(sf.view(torch.uint8), torch.from_numpy(g[f"sf_ext_{dn}"]))
    codes, sf, sf2 = opack.quantize(x[:, :184].contiguous(), 16)
    assert torch.equal(codes, torch.from_numpy(g[f"codes_rag_{dn}"]))
    assert torch.equal(sf.view(torch.uint8), torch.from_numpy(g[f"sf_rag_{dn}"]))
    assert torch.equal(sf2, torch.from_numpy(g[f"sf2_rag_{dn}"]))


# ---- reference's own known-answer vectors (tests/gpu/torch/quantization/test_tensor_quant_cuda.py:205-261)
E2M1 = [0, 0.5, 1, 1.5, 2, 3, 4, 6]
TIES_IN = [0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5, 6]


@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("block", [16, 32])
def test_e2m1_known_answers(dn, block):
    dt = DTYPES[dn]
    torch.manual_seed(0)
    sign = (torch.randint(0, 2, (1, 8)) * 2 - 1).float()
    rep = block // 8

    def run(tin, tout, dynamic=True):
        x = torch.cat((tin,) * rep, dim=-1).to(dt)
        want = torch.cat((tout,) * rep, dim=-1).to(dt)
        if dynamic:
            got = ofq.nvfp4_dynamic_qdq(x, x.float().abs().amax(), block)
            assert torch.allclose(got.float(), want.float())
        got = ofq.nvfp4_static_qdq(x.repeat(4, 1), torch.full((4,), 6.0), quantize_block_scales=False)
        assert torch.allclose(got.float(), want.repeat(4, 1).float(), atol=1e-6)

    grid = torch.tensor([E2M1]) * sign
    run(grid, grid)
    ties = torch.tensor([TIES_IN])
    run(ties * sign, torch.tensor([[0.0, 1, 1, 2, 2, 4, 4, 6]]) * sign)  # exact ties: IEEE division keeps them
    below = ties.clone(); below[:, :-1] -= 0.1
    run(below * sign, torch.tensor([E2M1]) * sign)
    above = ties.clone(); above[:, :-1] += 0.1
    run(above * sign, torch.tensor([[0.5, 1, 1.5, 2, 3, 4, 6, 6]]) * sign)


def test_block_scale_underflow_and_zero_block():
    # test_tensor_quant_cuda.py:183-192: scale below half the smallest E4M3 subnormal -> 0 (-> 1.0 guard)
    gs = torch.tensor(1.0)
    s = (torch.tensor(6.0 * 2**-11) / (6.0 * gs)).to(torch.float8_e4m3fn).float() * gs
    assert s.item() == 0.0
    # :194-203: zero static scale zeroes the block
    y = ofq.nvfp4_static_qdq(torch.ones(1, 16), torch.zeros(1), quantize_block_scales=False)
    assert torch.equal(y, torch.zeros(1, 16))
    # fp4 cast
    x = torch.tensor([0.25, -0.75, 1.25, -1.75, 2.5, -3.5, 5.0, 7.0, -0.0, float("nan")])
    y = ofq.fp4_cast(x)
    assert torch.equal(y[:8], torch.tensor([0.0, -1.0, 1.0, -2.0, 2.0, -4.0, 4.0, 6.0]))
    assert y[9].item() == -6.0  # NaN: all compares false -> 6, sign test false -> negative


def test_nvfp4_oracle_vs_reference_triton_on_b200():
    """The IEEE restatement against outputs of the reference's compiled Triton kernel captured on a B200
    (tests/golden/nvfp4_triton_b200.npz).  f32/f16: bit-equal.  bf16: the kernel's div.full.f32 moves a
    few elements per 10^4 across an E2M1 boundary; the product's "triton" numerics
    mode is what must be bit-equal there (tests/test_gpu_qdq.py)."""
    import numpy as np  # noqa: F401

    from conftest import DTYPES, from_np, load_golden, n_mismatch
    from oracle import fake_quant as ofq

    g = load_golden("nvfp4_triton_b200.npz")
    for dn, dt in DTYPES.items():
        for nm in ("rand", "adv"):
            x = from_np(g[f"x_{nm}_{dn}"], dt)
            for an in ("max", "clip"):
                ga = torch.from_numpy(g[f"ga_{nm}_{an}_{dn}"])
                bad = n_mismatch(ofq.nvfp4_dynamic_qdq(x, ga, 16), from_np(g[f"y_{nm}_{an}_{dn}"], dt))
                if dn == "bf16":
                    assert bad <= 1e-3 * x.numel(), (dn, nm, an, bad)
                else:
                    assert bad == 0, (dn, nm, an, bad)
