"""Real-quant GEMM registry (interface of modelopt/torch/quantization/backends/gemm_registry.py:31-156 with the two
default backends of backends/fp8_per_tensor_gemm.py and backends/nvfp4_gemm.py).  A quantized linear whose
``_use_real_quant_gemm`` flag is on (``enable_real_quant_gemm(model)`` / ``module.real_quant_gemm = True``) asks the
registry for a GEMM on the REAL codes; the defaults are the hand-written tcgen05 kernels: FP8 per-tensor
(``kind::f8f6f4``, csrc/gemm_fp8.cu) and NVFP4 block-scaled (``kind::mxf4nvf4``, csrc/gemm_nvfp4.cu)."""

from __future__ import annotations

from collections.abc import Callable

from torch import nn

__all__ = ["GEMMRegistry", "gemm_registry", "enable_real_quant_gemm", "disable_real_quant_gemm",
           "is_real_quant_gemm_enabled"]


class GEMMRegistry:
    """Ordered list of (gemm_func, availability_check); the first entry whose check accepts the call wins."""

    def __init__(self):
        self._registry: list[dict] = []

    def register(self, gemm_func: Callable, availability_check: Callable | None = None) -> None:
        self._registry.append({"gemm_func": gemm_func,
                               "availability_check": availability_check or (lambda module, input, args, kwargs: True)})

    def unregister(self, gemm_func: Callable) -> None:
        self._registry = [e for e in self._registry if e["gemm_func"] != gemm_func]

    def find_match(self, module: nn.Module, input, *args, **kwargs):
        for e in self._registry:
            if e["availability_check"](module, input, args, kwargs):
                return e["gemm_func"]
        return None

    def __contains__(self, gemm_func: Callable) -> bool:
        return any(e["gemm_func"] == gemm_func for e in self._registry)


gemm_registry = GEMMRegistry()


def enable_real_quant_gemm(model: nn.Module) -> None:
    """gemm_registry.py:104-124: mark every module whose input and weight quantizers are enabled."""
    for module in model.modules():
        iq, wq = getattr(module, "input_quantizer", None), getattr(module, "weight_quantizer", None)
        if iq is not None and wq is not None and iq.is_enabled and wq.is_enabled:
            module._use_real_quant_gemm = True


def disable_real_quant_gemm(model: nn.Module) -> None:
    for module in model.modules():
        if hasattr(module, "_use_real_quant_gemm"):
            module._use_real_quant_gemm = False


def is_real_quant_gemm_enabled(model: nn.Module) -> bool:
    return any(getattr(m, "_use_real_quant_gemm", False) for m in model.modules())


# ---- default backends ------------------------------------------------------------------------------------------------
def _plain_call(module, input, args, kwargs) -> bool:
    return not args and not kwargs and hasattr(module, "_tcgen05_ok") and module._tcgen05_ok(input, args, kwargs)


def _fp8_check(module, input, args, kwargs) -> bool:
    return _plain_call(module, input, args, kwargs) and module._real_fp8_ok(input)


def _nvfp4_check(module, input, args, kwargs) -> bool:
    return _plain_call(module, input, args, kwargs) and module._real_nvfp4_ok(input)


def fp8_per_tensor_gemm(module, input, weight=None, bias=None):
    """backends/fp8_per_tensor_gemm.py:53-91: `_to_fp8` of the input, cached E4M3 weight codes, scaled GEMM."""
    from . import ops

    scale_a = ops.div_scalar(module.input_quantizer._amax.float().reshape(1), 448.0)
    a = ops.q8_pack(input.reshape(-1, input.shape[-1]), scale_a, "fp8_sat")
    wq, scale_b = module.packed_fp8_weight()
    out = ops.linear_fp8(a, wq, scale_a, scale_b, module.bias, out_dtype=input.dtype)
    return out.reshape(*input.shape[:-1], out.shape[-1])


def nvfp4_gemm(module, input, weight=None, bias=None):
    """backends/nvfp4_gemm.py:28-125 (needs TensorRT-LLM there): pack the activation, cached packed weight, W4A4 GEMM."""
    from . import ops

    a = ops.nvfp4_quantize_operand(input.reshape(-1, input.shape[-1]), module.input_quantizer._amax.float().amax(), 128)
    out = ops.linear_nvfp4(*a, *module.packed_nvfp4_weight(), module.bias, out_dtype=input.dtype)
    return out.reshape(*input.shape[:-1], out.shape[-1])


gemm_registry.register(fp8_per_tensor_gemm, _fp8_check)
gemm_registry.register(nvfp4_gemm, _nvfp4_check)
