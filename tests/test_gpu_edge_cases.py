"""Edge cases of the round's new entry points through the C ABI: empty inputs, one-element / one-block inputs, ragged and
unaligned views, non-finite values, and rejected arguments (the cases the reference's own tests poke)."""

import pytest
import torch

pytestmark = pytest.mark.gpu

from modelopt_b200 import ops  # noqa: E402
from modelopt_b200.qtensor import FP8QTensor, INT4QTensor, INT8QTensor, MXFP4QTensor, MXFP8QTensor, NF4QTensor  # noqa: E402
from oracle import mx as omx  # noqa: E402
from oracle import qtensor as oq  # noqa: E402

bf = torch.bfloat16


def test_empty_inputs():
    e = torch.empty(0, 64, device="cuda", dtype=bf)
    assert ops.qdq_mx(e, 32, (4, 3)).shape == (0, 64)
    q, s = ops.mxfp8_pack(e)
    assert q.shape == (0, 64) and s.shape == (0, 2)
    assert ops.mx_unpack(q, s, 32, bf).shape == (0, 64)
    q, s = ops.int4_pack(e.reshape(-1), 32)
    assert q.numel() == 0 and s.numel() == 0
    assert ops.int4_unpack(q, s, 32).numel() == 0
    assert ops.q8_pack(e, torch.ones((), device="cuda"), "int8").shape == (0, 64)
    h = torch.zeros(4, 16, 16, device="cuda")
    ops.block_hessian_into(e, h)
    assert not h.any()


@pytest.mark.parametrize("fmt", ["E4M3", "E2M1", "INT8"])
def test_mx_nonfinite_and_extremes(fmt):
    bits = {"E4M3": (4, 3), "E2M1": (2, 1), "INT8": 8}[fmt]
    x = torch.zeros(4, 32, dtype=torch.float32)
    x[0, :4] = torch.tensor([3.0e38, -3.0e38, 1e-38, -1e-45])  # near fp32 max, denormals
    x[1, 0] = float("inf")   # block amax inf -> scale 1 (tensor_quant_mx.cu:137-139)
    x[2, :3] = torch.tensor([1.0, -2.0, 0.5])
    x[3] = 1e-30
    got = ops.qdq_mx(x.cuda(), 32, bits, numerics="cpu").cpu()
    want = omx.mx_fake_quant(x, 32, fmt)
    fin = torch.isfinite(want) & torch.isfinite(got)
    assert torch.equal(got[fin], want[fin])
    assert torch.equal(torch.isfinite(got), torch.isfinite(want))


def test_single_block_and_tiny_shapes():
    x = torch.tensor([[0.3, -1.7, 2.2, 0.0, 5.0, -6.0, 0.26, 0.74] * 4], dtype=bf)
    for fmt, bits in (("E4M3", (4, 3)), ("E2M1", (2, 1)), ("E3M2", (3, 2))):
        assert torch.equal(ops.qdq_mx(x.cuda(), 32, bits, numerics="cpu").cpu(), omx.mx_fake_quant(x, 32, fmt))
    with ops.numerics("cpu"):
        qt, s = INT4QTensor.quantize(x.cuda(), 32)
        oq_q, oq_s = oq.int4_quantize(x, 32)
        assert torch.equal(qt._quantized_data.cpu(), oq_q)
        qt, s = MXFP4QTensor.quantize(x.cuda(), 32)
        assert torch.equal(qt._quantized_data.cpu(), oq.mxfp4_quantize(x, 32)[0])


def test_padding_paths():
    """INT4 / NF4 pad the FLATTENED tensor to whole blocks (int4_tensor.py:59-60); FP8 / MXFP8 pad the last dim."""
    torch.manual_seed(0)
    x = (torch.randn(1, 350) * 2).to(bf)  # 350 elements: not a multiple of 128 (a padded multi-row tensor fails the
    # final reshape in the reference too, int4_tensor.py:86-87)
    with ops.numerics("cpu"):
        qt, s = INT4QTensor.quantize(x.cuda(), 128)
        want_q, want_s = oq.int4_quantize(x, 128)
        assert torch.equal(qt._quantized_data.cpu(), want_q) and torch.equal(s.cpu(), want_s)
        assert torch.equal(qt.dequantize(scale=s, block_sizes={-1: 128}).cpu(), oq.int4_dequantize(want_q, want_s, 128, x.shape, bf))
        qt, s = NF4QTensor.quantize(x.cuda(), 64, 8)
        want_q, want_s = oq.nf4_quantize(x, 64, 8)
        assert torch.equal(qt._quantized_data.cpu(), want_q) and torch.equal(s.cpu(), want_s)
        qt, s = MXFP8QTensor.quantize(x.cuda())
        want_q, want_s = oq.mxfp8_quantize(x)
        assert torch.equal(qt._quantized_data.view(torch.uint8).cpu(), want_q.view(torch.uint8)) and torch.equal(s.cpu(), want_s)
        x2 = (torch.randn(7, 50) * 2).to(bf)
        qt, s = FP8QTensor.quantize(x2.cuda(), block_sizes={-1: 16, -2: 4})
        want_q, want_s = oq.scaled8_quantize(x2, "fp8", block_sizes={-1: 16, -2: 4})
        assert torch.equal(qt._quantized_data.view(torch.uint8).cpu(), want_q.view(torch.uint8))


def test_scales_given_and_3d():
    """FP8 / INT8 with caller-provided scales and a 3-D (MoE-style) weight with per-expert 2-D blocks."""
    torch.manual_seed(1)
    w = (torch.randn(3, 64, 128) * 0.1).to(bf)
    with ops.numerics("cpu"):
        for cls, kind in ((INT8QTensor, "int8"), (FP8QTensor, "fp8")):
            qt, s = cls.quantize(w.cuda(), block_sizes={-1: 32, -2: 16})
            want_q, want_s = oq.scaled8_quantize(w, kind, block_sizes={-1: 32, -2: 16})
            assert s.shape == (3, 4, 4)
            assert torch.equal(qt._quantized_data.view(torch.uint8).cpu(), want_q.view(torch.uint8))
            qt2, s2 = cls.quantize(w.cuda(), scales=s, block_sizes={-1: 32, -2: 16})
            assert torch.equal(qt2._quantized_data.view(torch.uint8), qt._quantized_data.view(torch.uint8))
            dq = qt.dequantize(scale=s, block_sizes={-1: 32, -2: 16})
            assert torch.equal(dq.cpu(), oq.scaled8_dequantize(want_q, want_s, bf, w.shape, {-1: 32, -2: 16}, kind))


def test_rejected_arguments():
    x = torch.randn(4, 64, device="cuda", dtype=bf)
    with pytest.raises(ValueError):
        ops.qdq_mx(x, 64, (4, 3))                       # block not in {8, 16, 32}
    with pytest.raises(NotImplementedError):
        ops.qdq_mx(x, 32, (7, 0))                       # unknown format
    with pytest.raises(RuntimeError):
        ops.int4_pack(x.cpu().reshape(-1), 32)          # CPU tensor: no fallback
    with pytest.raises(AssertionError):
        ops.int4_pack(x.reshape(-1), 48)                # numel % block
    with pytest.raises(Exception):
        ops.block_hessian_into(x, torch.zeros(2, 32, 32, device="cuda"), block=32)  # only block 16 has a kernel
    with pytest.raises(ValueError):
        ops.nvfp4_fp8_scale_sweep_hessian(x.reshape(-1, 16), x.float().abs().max(), torch.zeros(3, 16, 8, device="cuda"))
    with pytest.raises(Exception):
        ops.gptq_block_nvfp4(torch.zeros(4, 48, device="cuda"), torch.ones(4, 3, device="cuda"), 1.0,
                             torch.eye(48, device="cuda"), 16, 0)  # block must be 32 / 64 / 128 / 256
