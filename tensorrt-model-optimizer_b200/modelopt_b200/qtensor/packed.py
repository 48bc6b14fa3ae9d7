"""INT4 / INT8 / FP8 / NF4 / MXFP4 / MXFP8 real-quantized tensors on the B200 path (interfaces of
modelopt/torch/quantization/qtensor/{int4,int8,fp8,nf4,mxfp4,mxfp8}_tensor.py).  Every ``quantize`` is ONE packing
kernel (plus one amax pass where the scale layout is not a last-dim block); every ``dequantize`` is one kernel.
The tiny second-level tensors (NF4 double quantization: numel / block_size scales) use torch ops on the device."""

from __future__ import annotations

import math

import numpy as np
import torch

from .. import ops
from ..utils import convert_quantization_axis_to_reduce_axis, reduce_amax, reduce_block_amax, reduce_block_padding
from .nvfp4_tensor import BaseQuantizedTensor

__all__ = ["INT4QTensor", "INT8QTensor", "FP8QTensor", "NF4QTensor", "MXFP4QTensor", "MXFP8QTensor"]


def _int_keys(block_sizes):
    return {k: v for k, v in (block_sizes or {}).items() if isinstance(k, int)}


class INT4QTensor(BaseQuantizedTensor):
    """int4_tensor.py:30-130: AWQ-style block INT4, even element in the high nibble, scales = 7 / block amax."""

    @staticmethod
    def _get_quant_maxbound(num_bits):
        return 2 ** (num_bits - 1) - 1

    @classmethod
    def quantize(cls, input: torch.Tensor, block_size: int) -> tuple:
        assert input.shape[-1] % 2 == 0, "Input tensor must have even number on last dimension."
        flat = reduce_block_padding(input.reshape(-1), {-1: block_size})
        q, scales = ops.int4_pack(flat, block_size)
        return cls(input.shape, flat.dtype, q.reshape(*input.shape[:-1], -1)), scales

    def dequantize(self, dtype: torch.dtype = None, **kwarg):
        dtype = dtype or self.metadata["dtype"]
        out = ops.int4_unpack(self._quantized_data, kwarg["scale"], kwarg["block_sizes"][-1])
        shape = self.metadata["shape"]
        return out[: int(np.prod(shape))].view(shape).to(dtype)


class _Scaled8(BaseQuantizedTensor):
    """Shared logic of INT8QTensor / FP8QTensor (int8_tensor.py:35-124, fp8_tensor.py:38-151)."""

    _kind = "int8"
    _maxbound = 127.0
    _pad_before_scales = False  # FP8 pads for given scales too (fp8_tensor.py:61-63); INT8 only when it computes them

    @classmethod
    def quantize(cls, input, scales=None, axis=None, block_sizes=None) -> tuple:
        original_input = input
        block_sizes = _int_keys(block_sizes)
        if block_sizes and (cls._pad_before_scales or scales is None):
            input = reduce_block_padding(input, block_sizes)
        if scales is None:
            if block_sizes:
                amax = reduce_block_amax(input, block_sizes)
            else:
                amax = reduce_amax(input, axis=convert_quantization_axis_to_reduce_axis(input, axis))
            scales = ops.div_scalar(amax, cls._maxbound)
        if block_sizes:
            expected = list(input.shape)
            for dim, bs in block_sizes.items():
                dim = dim if dim >= 0 else input.dim() + dim
                assert input.shape[dim] % bs == 0, f"Tensor dimension {dim}, {input.shape[dim]} is not divisible by {bs}."
                expected[dim] = input.shape[dim] // bs
            if scales.shape != tuple(expected) and scales.numel() == math.prod(expected) and cls._pad_before_scales:
                scales = scales.reshape(expected)
            assert scales.shape == tuple(expected), f"Mismatch in expected scale shape: {scales.shape} vs {tuple(expected)}"
        q = ops.q8_pack(input, scales, cls._kind)
        if q.shape != original_input.shape and cls._pad_before_scales:
            q = q[tuple(slice(0, d) for d in original_input.shape)]
        return cls(original_input.shape, original_input.dtype, q), scales

    def dequantize(self, dtype: torch.dtype = None, **kwarg):
        dtype = dtype or self.metadata["dtype"]
        assert "scale" in kwarg, "Require scale for dequantization."
        scales = kwarg["scale"]
        block_sizes = _int_keys(kwarg.get("block_sizes"))
        q = self._quantized_data
        if block_sizes:
            if self._pad_before_scales:
                q = reduce_block_padding(q.view(torch.uint8), block_sizes).view(q.dtype)
            expected = list(q.shape)
            for dim, bs in block_sizes.items():
                assert q.shape[dim] % bs == 0, f"Dimension {q.shape[dim]} is not divisible by {bs}."
                expected[dim] //= bs
            assert tuple(expected) == scales.shape, f"Scales shape {scales.shape} must match expected {tuple(expected)}."
        out = ops.q8_unpack(q, scales, dtype, self._kind)
        return out[tuple(slice(0, d) for d in self.metadata["shape"])]


class INT8QTensor(_Scaled8):
    _kind, _maxbound, _pad_before_scales = "int8", 127.0, False


class FP8QTensor(_Scaled8):
    _kind, _maxbound, _pad_before_scales = "fp8", 448.0, True


_NF4_TABLE = torch.tensor([-1.0000, -0.6962, -0.5251, -0.3949, -0.2844, -0.1848, -0.0911, 0.0000, 0.0796, 0.1609,
                           0.2461, 0.3379, 0.4407, 0.5626, 0.7230, 1.0000], dtype=torch.bfloat16)
nf4_table = _NF4_TABLE


class NF4QTensor(BaseQuantizedTensor):
    """nf4_tensor.py:68-199: NF4 indices (even element in the high nibble) + per-block amax, optional int8 double
    quantization of the scales."""

    @classmethod
    def quantize(cls, input: torch.Tensor, block_size: int, scale_block_size: int | None = None) -> tuple:
        flat = reduce_block_padding(input.reshape(-1), {-1: block_size})
        q, scales = ops.nf4_pack(flat, block_size)
        scales = reduce_block_padding(scales.view(-1), {-1: scale_block_size}) if scale_block_size else scales.view(-1)
        return cls(input.shape, input.dtype, q), scales

    @classmethod
    def double_quantization(cls, scales: torch.Tensor, scale_block_size: int, num_scale_bits: int):
        """nf4_tensor.py:126-151 on the (numel / block_size)-element scale vector."""
        assert scales.numel() % scale_block_size == 0, "Number of scales elements is not divisible by the scale block size."
        bound = 2 ** (num_scale_bits - 1) - 1
        block_scales = scales.view(-1, scale_block_size)
        zero_point = block_scales.mean()
        block_scales = block_scales - zero_point
        dq = bound / reduce_amax(block_scales, -1)
        q = (block_scales * dq.expand(block_scales.shape[0], scale_block_size)).round().clamp(-bound, bound).to(torch.int8)
        return q, dq.flatten(), zero_point

    def dequantize(self, dtype: torch.dtype = None, **kwarg):
        dtype = dtype or self.metadata["dtype"]
        block = kwarg["block_sizes"][-1]
        scales = kwarg["scale"].view(-1)[: (self._quantized_data.numel() * 2) // block]
        scales = ((scales / kwarg["double_scale"].unsqueeze(-1)).to(dtype) + kwarg["scale_zeros"]).flatten()
        # GPU numerics = the reference's NF4_dequantize kernel (bf16 table * bf16 scale -> bf16); CPU numerics = eager
        out = ops.nf4_unpack(self._quantized_data, scales, block, bf16_form=ops._mode(None) != ops.MO_DIV_IEEE)
        shape = self.metadata["shape"]
        return out[: int(np.prod(shape))].reshape(shape).to(dtype)


class MXFP4QTensor(BaseQuantizedTensor):
    """mxfp4_tensor.py:25-144: E2M1 codes (even element in the low nibble) under E8M0 block-32 scales."""

    E2M1_max = 6.0
    E2M1_values = [0, 0.5, 1, 1.5, 2, 3, 4, 6]
    E2M1_bounds = torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5])

    @classmethod
    def quantize(cls, input: torch.Tensor, block_size: int | None) -> tuple:
        block_size = 32 if block_size is None else block_size
        q, sf = ops.mxfp4_pack(input, block_size)
        return cls(input.shape, input.dtype, q), sf

    def dequantize(self, dtype: torch.dtype = None, **kwarg):
        dtype = dtype or self.metadata["dtype"]
        block = kwarg["block_sizes"][-1]
        q = self._quantized_data
        flat = q.reshape(-1, block // 2)  # scales follow the flattened blocks (mxfp4_tensor.py:131-139)
        out = ops.mx_unpack(flat, kwarg["scale"].reshape(-1, 1), block, dtype, fp4=True)
        return out.reshape(*q.shape[:-1], q.shape[-1] * 2)


class MXFP8QTensor(BaseQuantizedTensor):
    """mxfp8_tensor.py:26-262: E4M3 elements under E8M0 block-32 scales along the last dim."""

    E4M3_MAX = 448.0
    BLOCK_SIZE = 32
    SCALE_DTYPE = torch.uint8

    @classmethod
    def get_weights_scaling_factor(cls, weight: torch.Tensor) -> torch.Tensor:
        assert weight.dim() >= 2, f"Weight must be at least 2D, got {weight.dim()}D"
        assert weight.shape[-1] % cls.BLOCK_SIZE == 0, (
            f"Weight inner dimension ({weight.shape[-1]}) must be divisible by MXFP8 block size ({cls.BLOCK_SIZE})")
        return ops.mxfp8_pack(weight, scales_only=True)[1]

    @classmethod
    def get_weights_scaling_factor_from_quantizer(cls, weight, weight_quantizer) -> torch.Tensor:
        assert hasattr(weight_quantizer, "block_sizes"), "weight_quantizer must have 'block_sizes' attribute"
        assert weight_quantizer.block_sizes[-1] == cls.BLOCK_SIZE, (
            f"MXFP8 requires block size {cls.BLOCK_SIZE}, got {weight_quantizer.block_sizes[-1]}")
        assert weight.dim() >= 2, f"Weight must be at least 2D, got {weight.dim()}D"
        expected = (*weight.shape[:-1], weight.shape[-1] // cls.BLOCK_SIZE)
        scale = getattr(weight_quantizer, "_scale", None)
        if scale is not None:
            assert scale.dtype == cls.SCALE_DTYPE, f"MXFP8 scale must be {cls.SCALE_DTYPE} (E8M0 format), got {scale.dtype}"
            assert scale.shape == expected, f"Scale shape {scale.shape} does not match expected shape {expected}"
            return scale
        return cls.get_weights_scaling_factor(weight)

    @classmethod
    def quantize_with_scale(cls, weight: torch.Tensor, weights_scaling_factor: torch.Tensor) -> torch.Tensor:
        assert weights_scaling_factor.dtype == cls.SCALE_DTYPE, (
            f"weights_scaling_factor must be {cls.SCALE_DTYPE} (E8M0 format), got {weights_scaling_factor.dtype}")
        assert weight.shape[-1] % cls.BLOCK_SIZE == 0, (
            f"Weight inner dimension ({weight.shape[-1]}) must be divisible by MXFP8 block size ({cls.BLOCK_SIZE})")
        return ops.mxfp8_pack(weight, sf=weights_scaling_factor)[0]

    @classmethod
    def quantize(cls, input: torch.Tensor, weights_scaling_factor: torch.Tensor | None = None) -> tuple:
        q, sf = ops.mxfp8_pack(input, sf=weights_scaling_factor)  # a ragged last block is zero-padded in the kernel
        return cls(input.shape, input.dtype, q), sf

    def dequantize(self, dtype: torch.dtype = None, **kwargs) -> torch.Tensor:
        assert "scale" in kwargs, "dequantize requires 'scale' in kwargs"
        dtype = dtype or self.metadata["dtype"]
        return ops.mx_unpack(self._quantized_data, kwargs["scale"], self.BLOCK_SIZE, dtype, fp4=False)
