"""``mtq.compress`` on the B200 path (interface of modelopt/torch/quantization/compress.py:52-227 and
qtensor/base_qtensor.py:196-238): weights of quantized linears whose weight quantizer supports real quantization are
replaced by their packed form (ONE packing kernel per weight: csrc/pack_nvfp4.cu, csrc/pack_int.cu, csrc/qdq_mx.cu) held
in a ``QTensorWrapper`` parameter; the forward dequantizes with one kernel and runs the tcgen05 GEMM."""

from __future__ import annotations

import fnmatch
import warnings

import torch
from torch import nn

from .nn.tensor_quantizer import SequentialQuantizer, TensorQuantizer
from .qtensor import QTensorWrapper
from .utils import is_quantized_linear

__all__ = ["compress", "pack_real_quantize_weight", "is_real_quantized"]


def pack_real_quantize_weight(module: nn.Module, force_quantize: bool = False) -> int:
    """base_qtensor.py:196-238: real-quantize every weight whose quantizer is enabled and not in fake-quant mode.
    Returns the number of weights packed."""
    n = 0
    with torch.no_grad():
        for name, m in module.named_modules():
            if name == "" or not hasattr(m, "weight_quantizer"):
                continue
            w = m._parameters.get("weight") if hasattr(m, "_parameters") else None
            q = m.weight_quantizer
            if isinstance(q, SequentialQuantizer):
                if len(q) != 1:
                    continue
                q = q[0]
            if w is None or w.is_meta or isinstance(w, QTensorWrapper) or not isinstance(q, TensorQuantizer):
                continue
            if q.is_enabled and not q._fake_quant and w.element_size() > 1:
                if force_quantize:
                    q._dequantize = False
                m._parameters["weight"] = QTensorWrapper(q(w.data))
                m.invalidate_weight_cache()
                n += 1
    return n


def compress(model: nn.Module, config: dict | None = None) -> nn.Module:
    """compress.py:179-222.  ``config`` maps module-name patterns to booleans (``"default"``: everything else);
    None compresses every supported layer.  Modifies the model in place."""
    cfg = dict(config.get("compress", config)) if config else {"default": True}
    quantizers = []
    for name, m in model.named_modules():
        if is_quantized_linear(m) and hasattr(m, "weight_quantizer"):
            q = m.weight_quantizer
            for sub in (q if isinstance(q, SequentialQuantizer) else [q]):
                quantizers.append((f"{name}.weight_quantizer", sub))
    for pattern, flag in [("default", cfg.get("default"))] + [(k, v) for k, v in cfg.items() if k != "default"]:
        if flag is None:
            continue
        if not isinstance(flag, bool):
            raise ValueError(f"Invalid compression configuration: {flag}, expected a boolean as value.")
        for qname, q in quantizers:
            if pattern == "default" or fnmatch.fnmatch(qname, pattern):
                if q.is_enabled and (not flag or q._is_real_quantize_support()):
                    q._fake_quant = not flag
    if pack_real_quantize_weight(model):
        warnings.warn("Real quantization has been applied to the model. Converting the model back to its original state "
                      "is not available.")
    return model


def is_real_quantized(model: nn.Module) -> bool:
    return any(isinstance(getattr(m, "_parameters", {}).get("weight"), QTensorWrapper) for m in model.modules())
