"""GPU parity: the AWQ-clip block search kernel (mo_awq_clip_loss through the C ABI) against the oracle, and
mtq.quantize with awq_clip / awq_full against what the REAL reference produced on CPU (tests/golden/awq_clip_cpu.npz)."""

import copy

import pytest
import torch
from torch import nn

from conftest import DTYPES, load_golden
from oracle import fake_quant as ofq
from oracle import smooth as osm

pytestmark = pytest.mark.gpu


def _exact_losses(w, tok, w_amax, ratios, block, bits):
    """The same quantity with the difference taken first and everything after it in fp64."""
    co, ci = w.shape
    pad = (-ci) % block
    wp = torch.nn.functional.pad(w, (0, pad)) if pad else w
    xp = torch.nn.functional.pad(tok, (0, pad)) if pad else tok
    xb = xp.double().reshape(xp.shape[0], -1, block)
    out = []
    for r in ratios:
        amax = (w_amax * r).reshape(co, -1, 1)
        wq = ofq.int_qdq(wp.reshape(co, -1, block), amax, bits, False, False)
        d = (wq.double() - wp.reshape(co, -1, block).double())
        o = torch.einsum("tnb,cnb->ctn", xb, d)
        out.append(o.pow(2).mean(dim=1))
    return torch.stack(out)


@pytest.mark.parametrize("dn", list(DTYPES))
@pytest.mark.parametrize("co,ci,block,ntok,stride", [(96, 512, 128, 64, 1), (70, 256, 64, 5, 1), (64, 320, 32, 100, 2),
                                                     (33, 512, 256, 130, 1), (40, 200, 48, 17, 3), (8, 128, 128, 1, 1)])
def test_clip_loss_kernel(dn, co, ci, block, ntok, stride):
    from modelopt_b200 import ops

    torch.manual_seed(co + ci + block)
    dt = DTYPES[dn]
    w = (torch.randn(co, ci) * 0.05).to(dt)
    w[::5, ::9] *= 8
    w[3] = 0
    x_all = (torch.randn(ntok * stride, ci) * 2).to(dt)
    tok = x_all[0::stride]
    assert tok.shape[0] == ntok
    nb = -(-ci // block)
    wpad = torch.nn.functional.pad(w, (0, nb * block - ci))
    w_amax = wpad.reshape(co, nb, block).float().abs().amax(-1)
    ratios = osm.awq_clip_ratios(0.5, 0.05)
    for bits in (4, 8):
        want = _exact_losses(w, tok, w_amax, ratios, block, bits)
        loss = torch.zeros(len(ratios), co, nb, device="cuda")
        xg = x_all.cuda()
        ops.awq_clip_loss_into(w.cuda(), xg[0::stride], w_amax.cuda().view(-1), ratios, loss, block, num_bits=bits)
        torch.testing.assert_close(loss.cpu().double(), want, rtol=2e-4, atol=1e-12)
        # accumulates over calls
        ops.awq_clip_loss_into(w.cuda(), xg[0::stride], w_amax.cuda().view(-1), ratios, loss, block, num_bits=bits)
        torch.testing.assert_close(loss.cpu().double(), 2 * want, rtol=2e-4, atol=1e-12)
        if dn == "f32":  # the reference's own order of operations (two dots, then the difference)
            ref = osm.awq_clip_block_losses(w, tok, w_amax, ratios, block, bits)
            torch.testing.assert_close(loss.cpu() / 2, ref, rtol=2e-3, atol=1e-9)


def test_clip_loss_many_ratios():
    """More than 12 ratios: the host splits the list over launches."""
    from modelopt_b200 import ops

    torch.manual_seed(0)
    w = torch.randn(32, 256) * 0.1
    tok = torch.randn(48, 256)
    w_amax = w.reshape(32, 2, 128).abs().amax(-1)
    ratios = osm.awq_clip_ratios(0.2, 0.03)
    assert len(ratios) > 24
    loss = torch.zeros(len(ratios), 32, 2, device="cuda")
    ops.awq_clip_loss_into(w.cuda(), tok.cuda(), w_amax.cuda().view(-1), ratios, loss, 128)
    torch.testing.assert_close(loss.cpu().double(), _exact_losses(w, tok, w_amax, ratios, 128, 4), rtol=2e-4, atol=1e-12)


def test_clip_loss_rejects():
    from modelopt_b200 import _lib, ops

    w = torch.randn(8, 1024, device="cuda")
    with pytest.raises(AssertionError):
        ops.awq_clip_loss_into(w, w, torch.ones(8, device="cuda"), [1.0], torch.zeros(1, 8, 1, device="cuda"), 1024)
    arr = (_lib.C.c_float * 1)(1.0)
    rc = _lib.load().mo_awq_clip_loss(w.data_ptr(), w.data_ptr(), 8, 1024, 1024, 8, 1024, 0, w.data_ptr(), 0, arr, 1, 4,
                                      0, 0, w.data_ptr(), None)
    assert rc == _lib.MO_ERR_SHAPE if hasattr(_lib, "MO_ERR_SHAPE") else rc < 0


def _tiny(g):
    m = nn.Sequential(nn.Linear(256, 512), nn.ReLU(), nn.Linear(512, 256, bias=False))
    m.load_state_dict({k[2:]: torch.from_numpy(g[k]) for k in ("w_0.weight", "w_0.bias", "w_2.weight")})
    return m.cuda()


CASES = {
    "int4_clip_t64": ("INT4_AWQ_CFG", {"method": "awq_clip", "debug": True}),
    "int4_clip_t100": ("INT4_AWQ_CFG", {"method": "awq_clip", "max_tokens_per_batch": 100, "debug": True}),
    "int4_clip_step": ("INT4_AWQ_CFG", {"method": "awq_clip", "min_clip_ratio": 0.3, "shrink_step": 0.05, "debug": True}),
    "int4_full": ("INT4_AWQ_CFG", {"method": "awq_full", "alpha_step": 0.1, "debug": True}),
    "w4a8_clip": ("W4A8_AWQ_BETA_CFG", {"method": "awq_clip", "debug": True}),
}


@pytest.mark.parametrize("case", list(CASES))
def test_quantize_awq_clip_vs_reference(case):
    import modelopt_b200 as mtq
    from modelopt_b200 import ops
    from modelopt_b200.nn import TensorQuantizer

    g = load_golden("awq_clip_cpu.npz")
    name, algo = CASES[case]
    cfg = copy.deepcopy(getattr(mtq, name))
    cfg["algorithm"] = algo
    batches = [torch.from_numpy(g[f"batch{i}"]).cuda() for i in range(2)]
    m = _tiny(g)
    n0 = ops.LAUNCHES[0]
    with torch.no_grad():
        m = mtq.quantize(m, cfg, lambda model: [model(b) for b in batches])
    assert ops.LAUNCHES[0] > n0
    for mod_name in ("0", "2"):
        h = getattr(m[int(mod_name)], "awq_clip")
        want = torch.from_numpy(g[f"{case}|{mod_name}|loss"])
        assert h.num_tokens == int(g[f"{case}|{mod_name}|num_tokens"])
        assert torch.equal(h.w_amax.float().cpu().view(-1), torch.from_numpy(g[f"{case}|{mod_name}|w_amax"]).view(-1)) \
            or case == "int4_full"  # after awq_lite the weight was folded by a searched scale: compared below
        got = h.loss_acc.float().cpu()
        if case != "int4_full":
            torch.testing.assert_close(got, want, rtol=2e-3, atol=1e-9)
    checked = 0
    for n, mod in m.named_modules():
        if not isinstance(mod, TensorQuantizer):
            continue
        assert bool(g[f"{case}|{n}|enabled"]) == mod.is_enabled, n
        for buf in ("_amax", "_pre_quant_scale"):
            key = f"{case}|{n}|{buf}"
            if key not in g:
                assert not hasattr(mod, buf), (n, buf)
                continue
            want = torch.from_numpy(g[key])
            got = getattr(mod, buf).float().cpu()
            assert got.shape == want.shape, (n, buf)
            if buf == "_amax" and "weight_quantizer" in n and want.numel() > 1:
                # clip choice per block: identical except where two ratios tie to within the fp32 summation noise
                # (after awq_lite the folded weight carries the 1e-7 differences of the searched scale, hence relative)
                same = ((got - want).abs() <= 1e-5 * want.abs()).float().mean().item()
                assert same >= (0.97 if case == "int4_full" else 0.99), (n, same)
            else:
                torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-7)
            checked += 1
    assert checked >= 2
    with torch.no_grad():
        out = m(torch.from_numpy(g["test"]).cuda()).cpu()
    want = torch.from_numpy(g[f"{case}|out"])
    assert (out - want).abs().max().item() <= 0.05 * want.abs().max().item()
    assert (out - want).abs().mean().item() <= 2e-3 * want.abs().max().item()


def test_clip_loss_bf16_amax_rounding():
    """A bf16 amax buffer: the candidate amax is bf16(w_amax * ratio), what the reference's tensor product gives."""
    from modelopt_b200 import ops

    torch.manual_seed(5)
    w = (torch.randn(64, 256) * 0.1).bfloat16()
    tok = torch.randn(64, 256).bfloat16()
    w_amax = w.reshape(64, 2, 128).abs().amax(-1)                      # bf16
    ratios = osm.awq_clip_ratios()
    cand = torch.stack([(w_amax * r).float() for r in ratios])          # rounded to bf16 by torch
    assert not torch.equal(cand[1], w_amax.float() * ratios[1])
    want = torch.stack([_exact_losses(w, tok, cand[k], [1.0], 128, 4)[0] for k in range(len(ratios))])
    loss = torch.zeros(len(ratios), 64, 2, device="cuda")
    ops.awq_clip_loss_into(w.cuda(), tok.cuda(), w_amax.float().cuda().view(-1), ratios, loss, 128,
                           amax_dtype=torch.bfloat16)
    torch.testing.assert_close(loss.cpu().double(), want, rtol=2e-4, atol=1e-12)
