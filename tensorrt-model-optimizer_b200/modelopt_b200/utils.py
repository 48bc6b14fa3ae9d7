"""Quantization helpers with the reference's names (modelopt/torch/quantization/utils/core_utils.py)."""

from __future__ import annotations

import torch

from . import ops

__all__ = ["promote_static_block_weight_quantizers", "promote_nvfp4_static_quantizers", "find_shared_weight_groups",
           "SHARED_PATTERNS", "reduce_amax", "reduce_block_amax", "convert_quantization_axis_to_reduce_axis", "is_quantized",
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


def reduce_block_padding(input: torch.Tensor, block_sizes: dict, pad_value: float = 0) -> torch.Tensor:
    """core_utils.py:92-125: right-pad every blocked dim to a multiple of its block size."""
    import torch.nn.functional as F

    nd = input.dim()
    pad = [0] * (2 * nd)
    need = False
    for dim, bs in block_sizes.items():
        if not isinstance(dim, int):
            continue
        d = dim if dim >= 0 else nd + dim
        rem = input.size(d) % bs
        if rem:
            pad[(nd - 1 - d) * 2 + 1] = bs - rem
            need = True
    return F.pad(input, pad, value=pad_value) if need else input


def is_quantized(module) -> bool:
    from .nn.tensor_quantizer import TensorQuantizer

    return any(isinstance(m, TensorQuantizer) for m in module.modules())


def is_quantized_linear(module) -> bool:
    from .nn.quant_module import QuantLinearConvBase

    return isinstance(module, QuantLinearConvBase) and hasattr(module, "weight") and module.weight.dim() == 2


# ------------------------------------------------------------------------------------------------
# Static-block weight quantizers after max calibration (utils/core_utils.py:1077-1150, utils/shared_input.py)
# ------------------------------------------------------------------------------------------------
# Linears that export fuses into one GEMM (q/k/v -> qkv, gate/up, Mixtral w1/w3) must share ONE NVFP4 global scale.
# ``re.fullmatch`` on the module name; the captured parent keeps MoE experts apart (shared_input.py:58-62).
SHARED_PATTERNS = (
    r"(?:(.*)\.)?(?:q_proj|k_proj|v_proj)",
    r"(?:(.*)\.)?(?:gate_proj|up_proj)",
    r"(?:(.*)\.)?(?:w1|w3)",
)


def find_shared_weight_groups(model, patterns=SHARED_PATTERNS):
    """Lists of modules (with an enabled weight quantizer) whose names match the same pattern under the same parent,
    in module order (shared_input.py:420-470, name-based on purpose: "reads the same input" is broader than "gets
    fused")."""
    import re

    compiled = [re.compile(p) for p in (patterns or ())]
    buckets, order = {}, []
    for name, module in model.named_modules():
        q = getattr(module, "weight_quantizer", None)
        if q is None or not getattr(q, "is_enabled", False):
            continue
        for k, rx in enumerate(compiled):
            m = rx.fullmatch(name)
            if m is not None:
                key = (k, m.groups())
                if key not in buckets:
                    buckets[key] = []
                    order.append(key)
                buckets[key].append(module)
                break
    return [buckets[k] for k in order if len(buckets[k]) > 1]


def promote_static_block_weight_quantizers(model, shared_patterns=SHARED_PATTERNS) -> int:
    """After max calibration: every enabled static-block weight quantizer with an amax becomes a
    ``StaticBlockScaleQuantizer`` in place.  NVFP4 ones get their fp32 ``global_amax`` - the maximum of the per-block
    amax, taken over the whole fusible sibling group when there is one (the reference ties the siblings to one
    canonical buffer, `SharedWeightGlobalAmaxState`; here every member holds the same value) - and from then on
    fake-quantize with two-level scaling (E4M3 block scales under the global scale).  INT static-block quantizers are
    promoted for the learned-scale branch.  Returns the number of NVFP4 quantizers converted."""
    from .nn.quant_module import QuantModule
    from .nn.tensor_quantizer import SequentialQuantizer, StaticBlockScaleQuantizer, TensorQuantizer

    def first(q):
        if isinstance(q, SequentialQuantizer):
            return q[0] if len(q) else None
        return q if isinstance(q, TensorQuantizer) else None

    def eligible(q):
        return q is not None and q.is_enabled and q.is_static_block_quant and q.amax is not None

    shared = {}
    for group in find_shared_weight_groups(model, shared_patterns):
        qs = [first(m.weight_quantizer) for m in group]
        qs = [q for q in qs if eligible(q) and q.is_nvfp4_static and not q.amax.is_meta]
        if len(qs) > 1:
            g = torch.stack([reduce_amax(q.amax.detach().float(), axis=None).reshape(()) for q in qs]).max()
            for q in qs:
                shared[id(q)] = g
    converted = 0
    for module in list(model.modules()):
        if not isinstance(module, QuantModule):
            continue
        for _, q in module.iter_weights_for_calibration():
            q = first(q)
            if not eligible(q):
                continue
            if q.is_nvfp4_static:
                already = isinstance(q, StaticBlockScaleQuantizer)
                g = shared.get(id(q))
                if g is None:
                    g = reduce_amax(q.amax.clone().detach(), axis=None)
                StaticBlockScaleQuantizer.from_tensor_quantizer(q, global_amax=g)
                converted += 0 if already else 1
            elif isinstance(q._num_bits, int):
                StaticBlockScaleQuantizer.from_tensor_quantizer(q)
    return converted


promote_nvfp4_static_quantizers = promote_static_block_weight_quantizers
