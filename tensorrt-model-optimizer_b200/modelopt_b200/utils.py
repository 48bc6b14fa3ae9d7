"""Quantization helpers with the reference's names (modelopt/torch/quantization/utils/core_utils.py)."""

from __future__ import annotations

import torch

from . import ops

__all__ = ["reduce_amax", "reduce_block_amax", "convert_quantization_axis_to_reduce_axis", "is_quantized",
           "is_quantized_linear"]


def convert_quantization_axis_to_reduce_axis(inputs: torch.Tensor, axis):
    """core_utils.py:128-143: quantization axis (dims that keep their own amax) -> dims to reduce."""
    if axis is None:
        return None
    axis = axis if isinstance(axis, (list, tuple)) else [axis]
    return [i for i in range(inputs.dim()) if i not in axis and (i - inputs.dim()) not in axis]


def reduce_amax(inputs: torch.Tensor, axis=None, keepdims=True, squeeze_scalar=True) -> torch.Tensor:
    """core_utils.py:146-183 — one read through mo_amax instead of max + min + 2 abs + maximum."""
    return ops.reduce_amax(inputs, axis, keepdims, squeeze_scalar)


def reduce_block_amax(input_tensor: torch.Tensor, block_sizes: dict) -> torch.Tensor:
    """core_utils.py:43-89: per-block amax; ``block_sizes`` maps dim -> block size."""
    amax = input_tensor
    for dim, bs in block_sizes.items():
        if not isinstance(dim, int):
            continue
        dim = dim if dim >= 0 else amax.dim() + dim
        if amax.shape[dim] % bs != 0:
            raise ValueError(f"Tensor dimension {amax.shape[dim]} at dim {dim} is not divisible by block size {bs}")
        new_shape = [*amax.shape[:dim], amax.shape[dim] // bs, bs, *amax.shape[dim + 1:]]
        amax = ops.reduce_amax(amax.reshape(new_shape), dim + 1, keepdims=False, squeeze_scalar=False)
    return amax


def is_quantized(module) -> bool:
    from .nn.tensor_quantizer import TensorQuantizer

    return any(isinstance(m, TensorQuantizer) for m in module.modules())


def is_quantized_linear(module) -> bool:
    from .nn.quant_module import QuantLinearConvBase

    return isinstance(module, QuantLinearConvBase) and hasattr(module, "weight") and module.weight.dim() == 2
