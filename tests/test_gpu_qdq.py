"""GPU parity: the CUDA path (through the C ABI) vs the CPU oracle and the reference golden vectors.

Bar: value-exact (NaN == NaN, -0 == +0) for every dequantized output — i.e. 0 ulp of the I/O dtype."""

import pytest
import torch

from conftest import DTYPES, bits_equal, from_np, load_golden, n_mismatch
from oracle import amax as oamax
from oracle import fake_quant as ofq

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from modelopt_b200 import ops as _ops

    return _ops


def gpu(t):
    return t.cuda()


def rand(shape, dt, seed=0, scale=3.0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(shape, generator=g) * scale
    x.view(-1)[:: max(1, x.numel() // 7)] *= 25.0  # outliers
    return x.to(dt)


# ---------------------------------------------------------------------------------------- self-tests
def test_device_is_sm100(ops):
    import ctypes as C

    from modelopt_b200 import _lib

    name = C.create_string_buffer(128)
    sm, cc = C.c_int(0), C.c_int(0)
    assert _lib.load().mo_device_info(name, 128, C.byref(sm), C.byref(cc)) == 0
    assert cc.value >= 100, f"{name.value} cc {cc.value}"


def test_selftest_e2m1_cvt_equals_reference_compare_chain(ops):
    assert ops.selftest_e2m1() == 0


def test_selftest_exact_div_equals_ieee_division(ops):
    assert ops.selftest_exact_div(65536, 8192, 0) == 0
    assert ops.selftest_exact_div(2048, 50000, 7) == 0


# ---------------------------------------------------------------------------------------- golden
@pytest.mark.parametrize("dn", list(DTYPES))
def test_golden_int_fp8(ops, dn):
    g = load_golden("qdq_cpu.npz")
    dt = DTYPES[dn]
    x = from_np(g[f"x_{dn}"], dt)
    for bits in (3, 4, 5, 7, 8, 11):
        for u in (False, True):
            for nr in (False, True):
                k = f"int_{dn}_b{bits}_u{int(u)}_n{int(nr)}"
                xi = x.abs() if u else x
                y = ops.qdq_int(gpu(xi), gpu(torch.from_numpy(g[k + "_amax"])), bits, u, nr).cpu()
                assert bits_equal(y, from_np(g[k + "_y"], dt)), k
    for case, xx, bits in (("axis0", x, 8), ("axis1", x, 8), ("blk128", x.reshape(-1, 128), 4)):
        y = ops.qdq_int(gpu(xx), gpu(torch.from_numpy(g[f"int_{dn}_{case}_amax"])), bits, False, False).cpu()
        assert bits_equal(y, from_np(g[f"int_{dn}_{case}_y"], dt)), case
    for nm, am in (("tiny", 1e-26), ("zero", 0.0), ("eps2", 1.5 / (1 << 24))):
        y = ops.qdq_int(gpu(x), torch.tensor(am).cuda(), 8, False, False).cpu()
        assert bits_equal(y, from_np(g[f"int_{dn}_{nm}_y"], dt)), nm
    for case in ("max", "clip", "zero", "small", "none", "axis0", "axis1", "axis2"):
        xx = x.reshape(6, 8, 256) if case.startswith("axis") else x
        am = None if case == "none" else gpu(torch.from_numpy(g[f"fp8_{dn}_{case}_amax"]))
        y = ops.qdq_fp8(gpu(xx), am, numerics="cpu").cpu()
        assert bits_equal(y, from_np(g[f"fp8_{dn}_{case}_y"], dt)), case


def test_golden_fp4_scales(ops):
    g = load_golden("fp4_scales_cpu.npz")
    amax = torch.from_numpy(g["amax"])
    for nm in ("auto", "given", "small"):
        ga = None if nm == "auto" else gpu(torch.from_numpy(g[f"ga_{nm}"]))
        for q in (True, False):
            for m in (448.0, 256.0):
                s = ops.fp4_scales(gpu(amax), ga, q, m, numerics="ieee").cpu()
                assert bits_equal(s, torch.from_numpy(g[f"scale_{nm}_q{int(q)}_m{int(m)}"])), (nm, q, m)


@pytest.mark.parametrize("dn", list(DTYPES))
def test_golden_amax(ops, dn):
    g = load_golden("amax_cpu.npz")
    dt = DTYPES[dn]
    xs = [from_np(g[f"x{i}_{dn}"], dt) for i in range(3)]
    assert bits_equal(ops.reduce_amax(gpu(xs[0])).cpu().float(), torch.from_numpy(g[f"amax_all_{dn}"]))
    for axis, an in ((0, "0"), (1, "1"), (2, "2"), (-1, "-1"), ((0, 2), "0_2")):
        run = None
        for x in xs:
            red = oamax.quant_axis_to_reduce_axis(x, axis)
            loc = ops.reduce_amax(gpu(x), axis=red)
            run = loc if run is None else torch.max(run, loc)
        assert bits_equal(run.cpu().float(), torch.from_numpy(g[f"calib_axis{an}_{dn}"])), axis
    xb = xs[1].reshape(-1, 6, 16)
    got = ops.reduce_amax(gpu(xb), axis=-1, keepdims=False, squeeze_scalar=False).cpu().float()
    assert bits_equal(got, torch.from_numpy(g[f"blk16_{dn}"]))


# ---------------------------------------------------------------------------------------- vs oracle
SHAPES = [(1, 16), (3, 48), (7, 4096), (129, 1024), (64, 14336), (5, 17, 64), (1000,), (33, 7)]


@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("shape", SHAPES)
def test_fp8_vs_oracle(ops, dn, shape):
    dt = DTYPES[dn]
    x = rand(shape, dt, seed=len(shape) + shape[-1])
    for am in (x.float().abs().max(), x.float().abs().max() * 0.3, torch.tensor(0.0)):
        assert bits_equal(ops.qdq_fp8(gpu(x), gpu(am), numerics="cpu").cpu(), ofq.fp8_qdq(x, am))
    assert bits_equal(ops.qdq_fp8(gpu(x), None).cpu(), ofq.fp8_qdq(x, None))
    if len(shape) >= 2:
        for axis in range(len(shape)):
            red = [i for i in range(len(shape)) if i != axis]
            am = x.float().abs().amax(dim=red, keepdim=True)
            assert bits_equal(ops.qdq_fp8(gpu(x), gpu(am), numerics="cpu").cpu(), ofq.fp8_qdq(x, am)), axis


@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("shape", SHAPES)
def test_int_vs_oracle(ops, dn, shape):
    dt = DTYPES[dn]
    x = rand(shape, dt, seed=11 + shape[-1])
    for bits, u, nr in ((8, False, False), (8, False, True), (4, False, False), (8, True, False), (16, False, False)):
        xi = x.abs() if u else x
        for am in (xi.float().abs().max(), xi.float().abs().max() * 0.4):
            got = ops.qdq_int(gpu(xi), gpu(am), bits, u, nr).cpu()
            assert bits_equal(got, ofq.int_qdq(xi, am, bits, u, nr)), (bits, u, nr)
    if len(shape) >= 2:
        for axis in range(len(shape)):
            red = [i for i in range(len(shape)) if i != axis]
            am = x.float().abs().amax(dim=red, keepdim=True) * 0.8
            assert bits_equal(ops.qdq_int(gpu(x), gpu(am), 8, False, False).cpu(), ofq.int_qdq(x, am, 8, False, False))
    # extreme scales exercise the ExactDiv fallback window
    for am in (1e-20, 1e-6, 1e20, 3e38):
        a = torch.tensor(am)
        xx = (x.float() * am).to(dt) if am < 1e30 else x
        assert bits_equal(ops.qdq_int(gpu(xx), gpu(a), 8, False, False).cpu(), ofq.int_qdq(xx, a, 8, False, False)), am


def test_int_special_values_and_status(ops):
    x = torch.tensor([float("nan"), float("inf"), -float("inf"), 0.0, -0.0, 0.49, -0.49, 0.5, 1.5, 2.5, -2.5, 1e30,
                      -1e30, 126.5, 127.5, -128.5] * 4)
    am = torch.tensor(127.0)
    got = ops.qdq_int(gpu(x), gpu(am), 8, False, False).cpu()
    assert bits_equal(got, ofq.int_qdq(x, am, 8, False, False))
    assert torch.equal(torch.signbit(got[:16][3:7]), torch.tensor([False, True, False, True]))
    st = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.qdq_int(gpu(x.abs()), gpu(am), 8, True, False, status=st)
    assert st.item() == 0
    ops.qdq_int(gpu(x), gpu(am), 8, True, False, status=st)
    assert st.item() & 1  # MO_FLAG_NEGATIVE_IN_UNSIGNED


def test_int_block128_awq_layout(ops):
    w = rand((256, 512), torch.bfloat16, seed=5)
    wb = w.reshape(-1, 128)
    am = wb.float().abs().amax(dim=1, keepdim=True)
    assert bits_equal(ops.qdq_int(gpu(wb), gpu(am), 4, False, False).cpu(), ofq.int_qdq(wb, am, 4, False, False))


def test_unaligned_views_and_inplace(ops):
    base = rand((4099,), torch.bfloat16, seed=9).cuda()
    x = base[3:]  # 6-byte offset: not 32-byte aligned
    am = x.float().abs().max()
    assert bits_equal(ops.qdq_fp8(x, am, numerics="cpu").cpu(), ofq.fp8_qdq(x.cpu(), am.cpu()))
    assert bits_equal(ops.qdq_int(x, am, 8, False, False).cpu(), ofq.int_qdq(x.cpu(), am.cpu(), 8, False, False))
    want = ofq.fp8_qdq(base.cpu(), am.cpu())
    ops.qdq_fp8(base, am, out=base, numerics="cpu")
    assert bits_equal(base.cpu(), want)


# ---------------------------------------------------------------------------------------- NVFP4
NV_SHAPES = [(1, 16), (4, 64), (16, 4096), (33, 1024), (8, 14336), (2, 3, 128), (5, 100), (7, 4100), (3, 24)]


@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("shape", NV_SHAPES)
def test_nvfp4_dynamic_ieee_vs_oracle(ops, dn, shape):
    dt = DTYPES[dn]
    for seed, sc in ((1, 1.0), (2, 10.0), (3, 1e-3)):
        x = rand(shape, dt, seed=seed, scale=sc)
        for ga in (x.float().abs().max(), x.float().abs().max() * 0.7, torch.tensor(0.0)):
            got = ops.qdq_nvfp4_dynamic(gpu(x), gpu(ga), 16, numerics="ieee").cpu()
            assert bits_equal(got, ofq.nvfp4_dynamic_qdq(x, ga, 16)), (seed, float(ga))


@pytest.mark.parametrize("block", [32, 64])
def test_nvfp4_dynamic_other_blocks(ops, block):
    x = rand((9, 512), torch.bfloat16, seed=4)
    ga = x.float().abs().max()
    got = ops.qdq_nvfp4_dynamic(gpu(x), gpu(ga), block, numerics="ieee").cpu()
    assert bits_equal(got, ofq.nvfp4_dynamic_qdq(x, ga, block))


def test_nvfp4_dynamic_adversarial(ops):
    grid = torch.tensor([0, 0.5, 1, 1.5, 2, 3, 4, 6.0])
    ties = torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5, 6.0])
    rows = [torch.cat([grid, -grid]), torch.cat([ties, -ties]), torch.cat([ties - 0.1, -(ties + 0.1)]),
            torch.zeros(16), torch.full((16,), 1e-12), torch.cat([torch.tensor([1000.0]), torch.full((15,), 1e-3)]),
            torch.tensor([float("nan")] + [1.0] * 15), torch.tensor([float("inf")] + [1.0] * 15),
            torch.tensor([-0.0] * 8 + [6.0 * 2**-11] * 8)]
    x = torch.stack(rows)
    for dt in (torch.float32, torch.bfloat16):
        xx = x.to(dt)
        for ga in (torch.tensor(6.0), torch.tensor(1000.0), torch.tensor(float("nan"))):
            got = ops.qdq_nvfp4_dynamic(gpu(xx), gpu(ga), 16, numerics="ieee").cpu()
            want = ofq.nvfp4_dynamic_qdq(xx, ga, 16)
            assert bits_equal(got, want), (dt, ga, n_mismatch(got, want))


def test_nvfp4_triton_mode_is_within_one_code_of_ieee(ops):
    """div.full.f32 vs IEEE division differ by <= 2 ulp of the quotient: an element may land on the
    neighbouring E2M1 code only when the quotient sits on a rounding boundary."""
    x = rand((256, 4096), torch.bfloat16, seed=21)
    ga = x.float().abs().max().cuda()
    a = ops.qdq_nvfp4_dynamic(gpu(x), ga, 16, numerics="triton").float()
    b = ops.qdq_nvfp4_dynamic(gpu(x), ga, 16, numerics="ieee").float()
    frac = (a != b).float().mean().item()
    assert frac < 3e-3, frac
    _, s = ofq.nvfp4_dynamic_qdq(x, ga.cpu(), 16, return_scales=True)
    step = (s.repeat_interleave(16, dim=1) * 2.0).cuda()  # largest E2M1 step (4 -> 6) times the scale
    assert bool(((a - b).abs() <= step * 1.01).all())


@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("block", [8, 16, 32, 64])
@pytest.mark.parametrize("qs", [True, False])
def test_nvfp4_static_vs_oracle(ops, dn, block, qs):
    dt = DTYPES[dn]
    x = rand((64, block * 6), dt, seed=block)
    xb = x.reshape(-1, block)
    amax = xb.float().abs().amax(dim=1)
    amax[3] = 0.0
    for ga in (None, amax.max() * 2):
        for od in (None, torch.float32):
            got = ops.qdq_nvfp4_static(gpu(xb), gpu(amax), None if ga is None else gpu(ga), qs, 448.0, od,
                                       numerics="ieee").cpu()
            want = ofq.nvfp4_static_qdq(xb, amax, ga, qs, 448.0, od)
            assert bits_equal(got, want), (ga, od)


def test_fp4_cast(ops):
    x = torch.cat([torch.linspace(-8, 8, 4001), torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5.0, -0.0])])
    for dt in DTYPES.values():
        xx = x.to(dt)
        assert bits_equal(ops.fp4_cast(gpu(xx)).cpu(), ofq.fp4_cast(xx))


# ---------------------------------------------------------------------------------------- amax
@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("shape", [(1,), (17,), (4096,), (33, 7), (128, 4096), (4, 14336), (5, 17, 64), (2, 3, 4, 40),
                                   (300000,), (2, 70000)])
def test_amax_vs_oracle(ops, dn, shape):
    dt = DTYPES[dn]
    x = rand(shape, dt, seed=sum(shape))
    assert bits_equal(ops.reduce_amax(gpu(x)).cpu(), oamax.reduce_amax(x))
    for d in range(len(shape)):
        red = [i for i in range(len(shape)) if i != d]
        if not red:
            continue
        got = ops.reduce_amax(gpu(x), axis=red).cpu()
        assert bits_equal(got, oamax.reduce_amax(x, axis=red)), d
        got = ops.reduce_amax(gpu(x), axis=d, keepdims=False).cpu()
        assert bits_equal(got, oamax.reduce_amax(x, axis=d, keepdims=False)), d


def test_amax_nan_inf_propagate(ops):
    x = torch.randn(1000, 64)
    x[17, 3] = float("nan")
    assert torch.isnan(ops.reduce_amax(gpu(x))).item()
    assert torch.isnan(ops.reduce_amax(gpu(x.bfloat16()))).item()
    assert torch.isnan(ops.reduce_amax(gpu(x.half()))).item()
    r = ops.reduce_amax(gpu(x.half()), axis=1).cpu().flatten()
    assert torch.isnan(r[17]) and not torch.isnan(r[16])
    x[17, 3] = -float("inf")
    assert torch.isinf(ops.reduce_amax(gpu(x))).item()


def test_amax_block_and_unaligned(ops):
    x = rand((512, 4096), torch.bfloat16, seed=8)
    for bs in (16, 128):
        got = ops.reduce_amax(gpu(x.reshape(-1, bs)), axis=1).cpu()
        assert bits_equal(got, oamax.reduce_amax(x.reshape(-1, bs), axis=1))
    base = rand((8195,), torch.float16, seed=2).cuda()
    assert bits_equal(ops.reduce_amax(base[5:]).cpu(), oamax.reduce_amax(base[5:].cpu()))


# ---------------------------------------------------------------------------------------- full-size properties
def test_full_size_properties(ops):
    """BASELINE.json sizes: [16384, 14336] bf16.  Size-independent properties instead of a CPU rerun:
    idempotence (QDQ of a QDQ output with the same amax is a fixed point), amax of the output <= amax,
    per-block agreement with the oracle on a random sample of rows."""
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn((16384, 14336), generator=g, device="cuda", dtype=torch.bfloat16)
    am = ops.reduce_amax(x)
    assert torch.equal(am.float(), x.float().abs().max())
    for fn in (lambda t: ops.qdq_fp8(t, am, numerics="cpu"), lambda t: ops.qdq_int(t, am, 8, False, False)):
        y = fn(x)
        assert torch.equal(fn(y), y)
        assert ops.reduce_amax(y).float() <= am.float()
    rows = torch.randint(0, 16384, (64,), generator=torch.Generator().manual_seed(1))
    y = ops.qdq_nvfp4_dynamic(x, am.float(), 16, numerics="ieee")
    want = ofq.nvfp4_dynamic_qdq(x[rows.cuda()].cpu(), am.cpu().float(), 16)
    assert bits_equal(y[rows.cuda()].cpu(), want)
    y8 = ops.qdq_fp8(x, am, numerics="cpu")
    assert bits_equal(y8[rows.cuda()].cpu(), ofq.fp8_qdq(x[rows.cuda()].cpu(), am.cpu()))


# ---------------------------------------------------------------------------------------- reference GPU kernels
# tests/golden/nvfp4_triton_b200.npz holds outputs of the reference's OWN compiled kernels captured on a B200
# (oracle/gen_golden_gpu.py): Triton fp4_fake_quant_block / static_blockwise_fp4_fake_quant /
# compute_fp4_scales and the JIT CUDA extensions fake_e4m3fy / fake_tensor_quant[_with_axis].
# This is the comparison that pins the div.full.f32 step ("triton" numerics).  Bar: bit-exact.
def _same_bits(a, b):
    iv = torch.int32 if a.dtype == torch.float32 else torch.int16
    return int((a.contiguous().view(iv) != b.contiguous().view(iv)).sum())


@pytest.mark.parametrize("dn", list(DTYPES))
def test_reference_triton_nvfp4_dynamic_bit_exact(ops, dn):
    g = load_golden("nvfp4_triton_b200.npz")
    dt = DTYPES[dn]
    for nm in ("rand", "adv"):
        x = from_np(g[f"x_{nm}_{dn}"], dt)
        for an in ("max", "clip"):
            ga = torch.from_numpy(g[f"ga_{nm}_{an}_{dn}"])
            got = ops.qdq_nvfp4_dynamic(gpu(x), gpu(ga), 16, numerics="triton").cpu()
            assert _same_bits(got, from_np(g[f"y_{nm}_{an}_{dn}"], dt)) == 0, (nm, an)
    x = from_np(g[f"x_rag_{dn}"], dt)
    got = ops.qdq_nvfp4_dynamic(gpu(x), gpu(torch.from_numpy(g[f"ga_rag_{dn}"])), 16, numerics="triton").cpu()
    assert _same_bits(got, from_np(g[f"y_rag_{dn}"], dt)) == 0, "ragged"


@pytest.mark.parametrize("dn", list(DTYPES))
def test_reference_triton_nvfp4_static_bit_exact(ops, dn):
    g = load_golden("nvfp4_triton_b200.npz")
    dt = DTYPES[dn]
    xb = from_np(g[f"x_rand_{dn}"], dt).reshape(-1, 16)
    bam, gam = torch.from_numpy(g[f"bam_{dn}"]), torch.from_numpy(g[f"gam_{dn}"])
    for q in (True, False):
        s = ops.fp4_scales(gpu(bam), gpu(gam), q, 448.0, numerics="triton").cpu()
        assert _same_bits(s, torch.from_numpy(g[f"sstatic_q{int(q)}_{dn}"])) == 0, ("scales", q)
        y = ops.qdq_nvfp4_static(gpu(xb), gpu(bam), gpu(gam), q, 448.0, None, numerics="triton").cpu()
        assert _same_bits(y, from_np(g[f"ystatic_q{int(q)}_{dn}"], dt)) == 0, ("static", q)


@pytest.mark.parametrize("dn", list(DTYPES))
def test_reference_cuda_ext_fp8_int8_bit_exact(ops, dn):
    g = load_golden("nvfp4_triton_b200.npz")
    if f"fp8_y_{dn}" not in g:
        pytest.skip("reference JIT extensions were not captured")
    dt = DTYPES[dn]
    x = from_np(g[f"x_rand_{dn}"], dt)
    am = torch.from_numpy(g[f"fp8_amax_{dn}"]).to(dt)
    assert _same_bits(ops.qdq_fp8(gpu(x), gpu(am)).cpu(), from_np(g[f"fp8_y_{dn}"], dt)) == 0
    assert _same_bits(ops.qdq_int(gpu(x), gpu(am), 8, False, False).cpu(), from_np(g[f"int8_y_{dn}"], dt)) == 0
    amc = torch.from_numpy(g[f"int8c_amax_{dn}"]).to(dt)
    assert _same_bits(ops.qdq_int(gpu(x), gpu(amc), 8, False, False).cpu(), from_np(g[f"int8c_y_{dn}"], dt)) == 0


@pytest.mark.parametrize("kind", ["nvfp4", "fp8"])
def test_calib_qdq_host_batch_equals_separate_calls(kind):
    """mo_calib_qdq_host_batch == mo_amax_host followed by mo_qdq_*_host per tensor, with a running amax carried in."""
    from modelopt_b200 import ops

    torch.manual_seed(4)
    shapes = [(300, 256), (1, 64), (1024, 4096), (77, 1024), (0, 64), (513, 128)]
    xs = [(torch.randn(s) * (i + 1)).to(torch.bfloat16).pin_memory() for i, s in enumerate(shapes)]
    running = [0.0, 100.0, 0.5, 0.0, 3.0, 1e-3]
    want_am, want = [], []
    for x, r in zip(xs, running):
        am = max(r, ops.amax_host(x)) if x.numel() else r
        want_am.append(am)
        if not x.numel():
            want.append(torch.empty_like(x))
        elif kind == "nvfp4":
            want.append(ops.qdq_nvfp4_dynamic_host(x, am, 16))
        else:
            want.append(ops.qdq_fp8_host(x, am))
    got_am = list(running)
    got = ops.calib_qdq_host_batch(xs, got_am, kind)
    assert got_am == pytest.approx(want_am, rel=0, abs=0)
    for a, b in zip(got, want):
        assert torch.equal(a.view(torch.int16), b.view(torch.int16))
    # second pass: the running amax only grows
    again = list(got_am)
    ops.calib_qdq_host_batch(xs, again, kind)
    assert again == got_am


@pytest.mark.parametrize("kind", ["nvfp4", "fp8"])
def test_calib_qdq_host_batch_device_inputs_and_calibration_only(kind):
    """Device-resident inputs are read in place, a device output is written in place, a NULL output (calibration-only)
    returns only the running amax: all three agree with the all-host form."""
    from modelopt_b200 import ops

    torch.manual_seed(5)
    shapes = [(300, 256), (1024, 4096), (77, 1024), (513, 128), (64, 512)]
    xs = [(torch.randn(s) * (i + 1)).to(torch.bfloat16).pin_memory() for i, s in enumerate(shapes)]
    running = [0.0, 100.0, 0.5, 0.0, 3.0]
    want_am = list(running)
    want = ops.calib_qdq_host_batch(xs, want_am, kind)
    # tensors 1 and 3 live on the device; output 1 is a device tensor, outputs 2 and 4 are not wanted back
    mixed = [x.cuda() if i in (1, 3) else x for i, x in enumerate(xs)]
    outs = [torch.empty_like(xs[0]), torch.empty_like(mixed[1]), None, torch.empty_like(xs[3]), None]
    got_am = list(running)
    got = ops.calib_qdq_host_batch(mixed, got_am, kind, outs=outs)
    assert got_am == want_am
    for i in (0, 1, 3):
        assert torch.equal(got[i].cpu().view(torch.int16), want[i].view(torch.int16)), i
    only_am = list(running)
    assert ops.calib_qdq_host_batch(mixed, only_am, kind, download=False) == [None] * 5
    assert only_am == want_am
