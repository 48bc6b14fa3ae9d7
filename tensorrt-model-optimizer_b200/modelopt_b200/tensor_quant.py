"""Fake-quant functionals with the reference's names, positional signatures and autograd behaviour
(modelopt/torch/quantization/tensor_quant.py:319-707); the forward of every one is a single launch of a
hand-written sm_100a kernel through the C ABI (modelopt_b200.ops).  No Triton, no JIT, no eager fallback:
CPU tensors raise.

Backward is the reference's straight-through estimator with clipping (tensor_quant.py:290-316):
``grad_in = grad_out * (|x| <= amax)`` unless ``pass_through_bwd``.
"""

from __future__ import annotations

import torch
from torch.autograd import Function

from . import ops

__all__ = ["fake_tensor_quant", "scaled_e4m3", "dynamic_block_quant", "static_blockwise_fp4_fake_quant",
           "fp4_cast_ste", "FakeTensorQuantFunction", "ScaledE4M3Function", "DynamicBlockQuantizationFunction",
           "StaticBlockwiseFP4FakeQuantFunction", "FP4CastSTEFunction", "fp8_eager", "quantize_op",
           "dynamic_block_quantize_op", "OP_NAMESPACE", "int_cast_ste", "IntCastSTEFunction"]

E4M3_MAX = 448.0
mx_format_map = {(1, 2): "E1M2", (2, 1): "E2M1", (4, 3): "E4M3", (3, 2): "E3M2", (2, 3): "E2M3", (5, 2): "E5M2",
                 (8, 0): "E8M0", (0, 8): "INT8"}


# ------------------------------------------------------------------------------------------------ impls
def scaled_e4m3_impl(inputs: torch.Tensor, amax: torch.Tensor | None = None) -> torch.Tensor:
    """tensor_quant.py:67-92.  Scalar / single-axis amax -> mo_qdq_fp8 (the reference's CUDA-extension
    arithmetic); an N-D amax is what the reference sends to ``_fp8_eager`` and has no kernel here."""
    if amax is not None and amax.numel() > 1 and amax.squeeze().ndim > 1:
        raise NotImplementedError("modelopt_b200: FP8 fake-quant with an N-D amax has no CUDA path")
    return ops.qdq_fp8(inputs, amax)


fp8_eager = scaled_e4m3_impl


def fake_quant_impl(inputs, amax, num_bits=8, unsigned=False, narrow_range=True):
    """tensor_quant.py:95-112 -> mo_qdq_int (per-tensor, per-axis, and the [n/128, 128] block view)."""
    return ops.qdq_int(inputs, amax, num_bits, unsigned, narrow_range)


def _quantize_impl(inputs, amax, num_bits=8, exponent_bits=0, unsigned=False, narrow_range=True):
    """``tensorrt::quantize_op`` (tensor_quant.py:115-136, :222-229)."""
    if num_bits == 8 and exponent_bits == 4:
        return scaled_e4m3_impl(inputs, amax)
    if isinstance(num_bits, int):
        return fake_quant_impl(inputs, amax, num_bits, unsigned, narrow_range)
    raise ValueError(f"Invalid combination of (num_bits, exponent_bits): ({num_bits}, {exponent_bits}).")


def _dynamic_block_quantize_impl(inputs, block_size, amax, num_bits, exponent_bits, scale_num_bits,
                                 scale_exponent_bits):
    """``tensorrt::dynamic_block_quantize_op`` (tensor_quant.py:157-195, :231-237)."""
    scale_bits = (scale_exponent_bits, scale_num_bits - scale_exponent_bits - 1)
    fmt = (exponent_bits, num_bits - exponent_bits - 1) if exponent_bits != 0 else num_bits
    if fmt in ops.MX_FORMATS and fmt != (8, 0):
        assert scale_bits in ops.MX_FORMATS, f"Scale bits should be in {ops.MX_FORMATS.keys()}"
        if scale_bits != (8, 0):
            assert amax is not None and amax.is_cuda, "amax must be a CUDA tensor for dynamic block quantization."
        if fmt == (2, 1) and scale_bits == (4, 3) and amax is not None:
            return ops.qdq_nvfp4_dynamic(inputs, amax, block_size)  # the reference's Triton fp4_fake_quant_block
        return ops.qdq_mx(inputs, block_size, fmt, scale_bits, amax)  # cuda_ext_mx.fused_amax_convert
    raise NotImplementedError(f"Unsupported num_bits: {fmt}, scale_bits: {scale_bits} for dynamic block quantization.")


# ------------------------------------------------------------------------------------------------ op registration
def _register_library_ops():
    """tensor_quant.py:216-276: the two functionals as ``torch.library`` operators with fake (shape-only)
    implementations, so that ``torch.export`` / ``torch.compile`` can trace a quantized model without running a kernel.
    Same schemas and, when free, the same ``tensorrt::`` names as the reference (an exported graph then carries the
    node names its TensorRT consumers look for); if the reference itself is imported in this process and already owns
    them, the ops go to ``modelopt_b200::``.  CUDA implementations only - there is no CPU kernel to register."""
    schemas = {
        "quantize_op": "(Tensor input, Tensor amax, int num_bits, int exponent_bits, bool unsigned, "
                       "bool narrow_range) -> Tensor",
        "dynamic_block_quantize_op": "(Tensor input, int block_size, Tensor amax, int num_bits, int exponent_bits, "
                                     "int scale_num_bits, int scale_exponent_bits) -> Tensor",
        "dynamic_block_quantize_op.overload": "(Tensor input, int block_size, None amax, int num_bits, "
                                              "int exponent_bits, int scale_num_bits, int scale_exponent_bits) "
                                              "-> Tensor",
    }

    def fake(input, *args, **kwargs):
        return torch.empty_like(input)

    def none_amax(inputs, block_size, amax, num_bits, exponent_bits, scale_num_bits, scale_exponent_bits):
        return torch.empty_like(inputs)  # what the reference registers for this overload (:244-254)

    impls = {"quantize_op": _quantize_impl, "dynamic_block_quantize_op": _dynamic_block_quantize_impl,
             "dynamic_block_quantize_op.overload": none_amax}
    for ns in ("tensorrt", "modelopt_b200"):
        try:
            for name, schema in schemas.items():
                torch.library.define(f"{ns}::{name}", schema)
            for name, fn in impls.items():
                torch.library.impl(f"{ns}::{name}", ["cuda"])(fn)
                torch.library.register_fake(f"{ns}::{name}")(fake)
            return ns
        except (AttributeError, RuntimeError):
            continue
    return None


OP_NAMESPACE = _register_library_ops()


def quantize_op(inputs, amax, num_bits=8, exponent_bits=0, unsigned=False, narrow_range=True):
    """Eager calls go straight to the kernel wrapper (no dispatcher round trip per quantizer call); under
    ``torch.compile`` / ``torch.export`` tracing the registered operator is used so that the graph keeps one node."""
    if OP_NAMESPACE is not None and torch.compiler.is_compiling() and isinstance(amax, torch.Tensor):
        return getattr(torch.ops, OP_NAMESPACE).quantize_op(inputs, amax, num_bits, exponent_bits, unsigned,
                                                            narrow_range)
    return _quantize_impl(inputs, amax, num_bits, exponent_bits, unsigned, narrow_range)


def dynamic_block_quantize_op(inputs, block_size, amax, num_bits, exponent_bits, scale_num_bits, scale_exponent_bits):
    if OP_NAMESPACE is not None and torch.compiler.is_compiling() and isinstance(amax, torch.Tensor):
        return getattr(torch.ops, OP_NAMESPACE).dynamic_block_quantize_op(
            inputs, block_size, amax, num_bits, exponent_bits, scale_num_bits, scale_exponent_bits)
    return _dynamic_block_quantize_impl(inputs, block_size, amax, num_bits, exponent_bits, scale_num_bits,
                                        scale_exponent_bits)


# ------------------------------------------------------------------------------------------------ backward
def _save_for_backward_if_needed(ctx, pass_through_bwd, inputs, amax):
    if not pass_through_bwd and amax is not None:
        if not isinstance(amax, torch.Tensor):
            amax = torch.tensor(amax, device=inputs.device, dtype=inputs.dtype)
        ctx.save_for_backward(inputs, amax)


def _ste_backward(ctx, grad_outputs, num_args):
    saved = ctx.saved_tensors
    if len(saved) == 0:
        return (grad_outputs,) + (None,) * (num_args - 1)
    inputs, amax = saved
    grad = torch.where(inputs.abs() <= amax, grad_outputs, grad_outputs.new_zeros(1))
    return (grad,) + (None,) * (num_args - 1)


# ------------------------------------------------------------------------------------------------ Functions
class FakeTensorQuantFunction(Function):
    """tensor_quant.py:319-399."""

    @staticmethod
    def forward(ctx, inputs, amax, bias=None, num_bits=8, unsigned=False, narrow_range=True,
                trt_high_precision_dtype=None, pass_through_bwd=False, block_size=None, axis=None):
        if bias is not None:
            inputs = inputs - bias
        _save_for_backward_if_needed(ctx, pass_through_bwd, inputs, amax)
        if unsigned:
            status = torch.zeros(1, dtype=torch.int32, device=inputs.device)
            outputs = ops.qdq_int(inputs, amax, num_bits, unsigned, narrow_range, status=status)
            FakeTensorQuantFunction._pending_status = status  # checked lazily by check_unsigned_inputs()
        else:
            outputs = quantize_op(inputs, amax, num_bits=num_bits, exponent_bits=0, unsigned=unsigned,
                                  narrow_range=narrow_range)
        if bias is not None:
            outputs = outputs + bias
        return outputs

    _pending_status = None

    @staticmethod
    def backward(ctx, grad_outputs):
        return _ste_backward(ctx, grad_outputs, 10)


def check_unsigned_inputs():
    """The reference's CPU path raises TypeError on negative inputs to unsigned quantization
    (tensor_quant.py:610-611) and its CUDA kernel device-asserts; here the kernel sets a status word and this
    (one sync) turns it into the same TypeError."""
    st = FakeTensorQuantFunction._pending_status
    if st is not None and int(st.item()) & 1:
        FakeTensorQuantFunction._pending_status = None
        raise TypeError("Negative values encountered in unsigned quantization.")


class ScaledE4M3Function(Function):
    """tensor_quant.py:402-460."""

    @staticmethod
    def forward(ctx, inputs, amax, bias, E, M, trt_high_precision_dtype=None, pass_through_bwd=False):  # noqa: N803
        if E != 4 or M != 3:
            raise NotImplementedError("Only support E=4 & M=3 for now.")
        if bias is not None:
            inputs = inputs - bias
        _save_for_backward_if_needed(ctx, pass_through_bwd, inputs, amax)
        outputs = quantize_op(inputs, amax, num_bits=8, exponent_bits=4, unsigned=False, narrow_range=False)
        if bias is not None:
            outputs = outputs + bias
        return outputs

    @staticmethod
    def backward(ctx, grad_outputs):
        return _ste_backward(ctx, grad_outputs, 7)


class DynamicBlockQuantizationFunction(Function):
    """tensor_quant.py:497-568."""

    @staticmethod
    def forward(ctx, inputs, block_size, amax, bias, num_bits, scale_bits, trt_high_precision_dtype=None,
                onnx_quantizer_type="dynamic", pass_through_bwd=True):
        _save_for_backward_if_needed(ctx, pass_through_bwd, inputs, amax)
        if isinstance(num_bits, int):
            exponent_bits = 0
        else:
            assert isinstance(num_bits, tuple) and len(num_bits) == 2
            exponent_bits = num_bits[0]
            num_bits = num_bits[0] + num_bits[1] + 1
        assert isinstance(scale_bits, tuple) and len(scale_bits) == 2
        return dynamic_block_quantize_op(inputs, block_size, amax, num_bits, exponent_bits,
                                         scale_bits[0] + scale_bits[1] + 1, scale_bits[0])

    @staticmethod
    def backward(ctx, grad_outputs):
        return _ste_backward(ctx, grad_outputs, 9)


class StaticBlockwiseFP4FakeQuantFunction(Function):
    """tensor_quant.py:571-604."""

    @staticmethod
    def forward(ctx, x, amax, global_amax=None, quantize_block_scales=True, fp8_max_for_normalization=E4M3_MAX,
                out_dtype=None, pass_through_bwd=False):
        _save_for_backward_if_needed(ctx, pass_through_bwd, x, amax)
        return ops.qdq_nvfp4_static(x, amax, global_amax, quantize_block_scales, fp8_max_for_normalization, out_dtype)

    @staticmethod
    def backward(ctx, grad_outputs):
        return _ste_backward(ctx, grad_outputs, len(ctx.needs_input_grad))


class FP4CastSTEFunction(Function):
    """tensor_quant.py:648-669: round pre-scaled values to the E2M1 grid, identity backward."""

    @staticmethod
    def forward(ctx, x):
        return ops.fp4_cast(x)

    @staticmethod
    def backward(ctx, grad_outputs):
        return grad_outputs


class IntCastSTEFunction(Function):
    """tensor_quant.py:672-699: clamp(round(x)) to the INT-N range of already scaled values (the learned-scale branch of
    the static block quantizer, tensor_quantizer.py:1697); backward passes the gradient inside the range only.  A
    [min, max] clamp of a rounded tensor: two elementwise torch ops, used on weights during QAT, not on the PTQ path."""

    @staticmethod
    def forward(ctx, x, num_bits, unsigned=False, narrow_range=True):
        max_bound = (2.0 ** (num_bits - 1 + int(unsigned))) - 1.0
        min_bound = 0 if unsigned else (-max_bound if narrow_range else -max_bound - 1)
        ctx.save_for_backward(x)
        ctx.min_bound, ctx.max_bound = min_bound, max_bound
        return torch.clamp(x.round(), min_bound, max_bound)

    @staticmethod
    def backward(ctx, grad_outputs):
        (x,) = ctx.saved_tensors
        inside = (x >= ctx.min_bound) & (x <= ctx.max_bound)
        return torch.where(inside, grad_outputs, torch.zeros_like(grad_outputs)), None, None, None


class _NoGradCtx:
    """Stand-in for the autograd context when no gradient can flow: `Function.forward(_NOCTX, ...)` runs the same code
    as `.apply` without building a graph node (about 15 us of host time per quantizer call)."""

    needs_input_grad = ()

    @staticmethod
    def save_for_backward(*tensors):
        pass


_NOCTX = _NoGradCtx()


def call_functional(fn_cls, inputs, *args):
    """`fn_cls.apply(inputs, *args)`; bypasses autograd when grad mode is off or the input does not require grad."""
    if torch.is_grad_enabled() and isinstance(inputs, torch.Tensor) and inputs.requires_grad:
        return fn_cls.apply(inputs, *args)
    return fn_cls.forward(_NOCTX, inputs, *args)


fake_tensor_quant = FakeTensorQuantFunction.apply
int_cast_ste = IntCastSTEFunction.apply
scaled_e4m3 = ScaledE4M3Function.apply
dynamic_block_quant = DynamicBlockQuantizationFunction.apply
static_blockwise_fp4_fake_quant = StaticBlockwiseFP4FakeQuantFunction.apply
fp4_cast_ste = FP4CastSTEFunction.apply
