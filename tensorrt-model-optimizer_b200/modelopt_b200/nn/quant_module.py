"""Quantized wrappers of torch modules (interface of modelopt/torch/quantization/nn/modules/quant_module.py:41-300,
quant_linear.py:38-53, quant_conv.py, plugins/huggingface.py:559-575).

A module is converted IN PLACE: its class is swapped for a generated subclass ``Quant<Orig>`` that adds
``input_quantizer`` / ``output_quantizer`` (and ``weight_quantizer`` for linear / conv) and quantizes in ``forward``:

    out = output_quantizer(orig_forward(input_quantizer(x)))      with  self.weight -> weight_quantizer(weight)

Parameters, buffers, hooks and the state_dict layout of the original module are untouched, so checkpoints keep
loading.  One B200-specific behaviour: while a weight quantizer is only *collecting* with a ``MaxCalibrator`` the
result cannot change between forwards unless the weight does, so the re-collection the reference does on every
forward (a 2 x 14 GB read per micro-batch for Llama-3-8B) is skipped when the weight's
(data_ptr, version) is unchanged — max is idempotent, outputs are identical.
"""

from __future__ import annotations

import contextlib
import warnings
from typing import Any

import torch
from torch import nn

from ..calib import MaxCalibrator
from ..config import QuantizerAttributeConfig
from .tensor_quantizer import SequentialQuantizer, TensorQuantizer

__all__ = ["QuantModule", "QuantInputBase", "QuantLinearConvBase", "QuantModuleRegistry", "ParallelState",
           "RealQuantLinear"]

QUANT_DESC_8BIT_PER_TENSOR = QuantizerAttributeConfig(num_bits=8)
QUANT_DESC_8BIT_LINEAR_WEIGHT_PER_ROW = QuantizerAttributeConfig(num_bits=8, axis=(0,))


class ParallelState:
    """Process groups a quantized module synchronises calibration statistics over
    (modelopt/torch/utils/distributed.py:441-480).  ``None`` = the default (world) group; ``-1`` = no group."""

    def __init__(self, data_parallel_group=None, tensor_parallel_group=-1, expert_model_parallel_group=-1):
        self.data_parallel_group = data_parallel_group
        self.tensor_parallel_group = tensor_parallel_group
        self.expert_model_parallel_group = expert_model_parallel_group

    def __repr__(self):
        return (f"ParallelState(dp={self.data_parallel_group}, tp={self.tensor_parallel_group}, "
                f"ep={self.expert_model_parallel_group})")


class QuantModule(nn.Module):
    """Base of every generated quantized class."""

    _parallel_state: ParallelState
    _original_cls: type

    def _setup(self):
        raise NotImplementedError

    @property
    def parallel_state(self):
        return getattr(self, "_parallel_state", None)

    @parallel_state.setter
    def parallel_state(self, value):
        assert isinstance(value, ParallelState), "parallel_state must be a ParallelState object!"
        self._parallel_state = value

    def iter_weights_for_calibration(self):
        """Yield ``(weight, weight_quantizer)`` pairs for weight-only calibration."""
        if hasattr(self, "weight_quantizer") and getattr(self, "weight", None) is not None:
            yield self._parameters["weight"], self.weight_quantizer

    def fold_weight(self, keep_attrs: bool = False):
        """quant_module.py:160-186: replace the weight by its fake-quantized value (computed in fp32) and switch the
        weight quantizer off."""
        for weight, q in self.iter_weights_for_calibration():
            if isinstance(q, (TensorQuantizer, SequentialQuantizer)) and getattr(q, "fake_quant", True) \
                    and getattr(q, "is_enabled", True):
                with torch.no_grad():
                    weight.data.copy_(q(weight.float().contiguous()).to(weight.dtype))
                if hasattr(self, "invalidate_weight_cache"):
                    self.invalidate_weight_cache()  # `.data.copy_` does not move the weight's version counter
                q.disable()
                if not keep_attrs:
                    for name in ("_pre_quant_scale", "_amax"):
                        if hasattr(q, name):
                            delattr(q, name)

    def modelopt_post_restore(self, prefix: str = ""):
        ref = next((p for n, p in self.state_dict().items() if "_quantizer" not in n), None)
        if ref is None:
            warnings.warn(f"Could not identify the device for TensorQuantizer states of {prefix}.")
            return
        for m in self.modules():
            if isinstance(m, TensorQuantizer):
                m.to(ref.device)


class QuantInputBase(QuantModule):
    """Modules whose input (and optionally output) is quantized."""

    default_quant_desc_input = QUANT_DESC_8BIT_PER_TENSOR
    default_quant_desc_output = QUANT_DESC_8BIT_PER_TENSOR

    def _setup(self):
        self.input_quantizer = TensorQuantizer(self.default_quant_desc_input)
        self.output_quantizer = TensorQuantizer(self.default_quant_desc_output)
        self.output_quantizer.disable()

    def forward(self, input, *args, **kwargs):
        input = self.input_quantizer(input)
        output = self._original_cls.forward(self, input, *args, **kwargs)
        if isinstance(output, tuple):
            return (self.output_quantizer(output[0]), *output[1:])
        return self.output_quantizer(output)


_WEIGHT_CACHES = ("_wq_cache", "_w4_cache", "_w8_cache", "_weight_collect_key")
CHECK_CACHE = __import__("os").environ.get("MODELOPT_B200_CHECK_CACHE", "0") not in ("", "0")


def _no_cache() -> bool:
    from .. import ops

    return ops.NO_CACHE


def _fingerprint(w):
    """Debug aid (MODELOPT_B200_CHECK_CACHE=1): content checksum of a weight, to catch a write through `.data` / a raw
    pointer that left (data_ptr, _version) unchanged and would otherwise be served a stale fake-quantized weight."""
    f = w.detach().float()
    return torch.stack([f.sum(), f.abs().sum(), f.reshape(-1)[:: max(1, f.numel() // 997)].sum()]).tolist()


def _cache_lookup(module, slot, key, w):
    """Cached value of `slot` for `key`, or None.  With MODELOPT_B200_NO_CACHE=1 nothing is ever served from a cache (the
    reference's always-requantize behaviour); with MODELOPT_B200_CHECK_CACHE=1 a hit is verified against a checksum."""
    if _no_cache():
        return None
    hit = module.__dict__.get(slot)
    if hit is None or hit[0] != key:
        return None
    if CHECK_CACHE and len(hit) > 2 and hit[2] != _fingerprint(w):
        raise RuntimeError(
            f"modelopt_b200: the weight of {type(module).__name__} changed without its version counter moving (a write "
            "through `.data` or a raw pointer); call module.invalidate_weight_cache() after such writes or set "
            "MODELOPT_B200_NO_CACHE=1")
    return hit[1]


# Device memory the cached functions of the weights (fake-quantized copies, packed operands) may take in total, per
# device.  A cached fake-quantized weight doubles that weight's footprint: fine for an 8B model (14 GB), impossible for a
# 70B replica on one 180 GB GPU, where the weights past the budget are re-quantized on every forward like the reference
# does (quant_module.py:192-278).  MODELOPT_B200_WEIGHT_CACHE_GB overrides (0 = never cache).
_CACHE_BUDGET_GB = __import__("os").environ.get("MODELOPT_B200_WEIGHT_CACHE_GB")
_CACHE_USED: dict = {}


def _cache_budget(device) -> int:
    if _CACHE_BUDGET_GB is not None:
        return int(float(_CACHE_BUDGET_GB) * (1 << 30))
    total = torch.cuda.get_device_properties(device).total_memory
    return min(32 << 30, total // 5)


def _nbytes(value) -> int:
    if isinstance(value, torch.Tensor):
        return value.numel() * value.element_size()
    if isinstance(value, (tuple, list)):
        return sum(_nbytes(v) for v in value)
    return 0


def _first_tensor(value):
    if isinstance(value, torch.Tensor):
        return value
    if isinstance(value, (tuple, list)):
        for v in value:
            t = _first_tensor(v)
            if t is not None:
                return t
    return None


def _release(dev, n):
    _CACHE_USED[dev] = _CACHE_USED.get(dev, 0) - n


def _cache_drop(module, slot):
    module.__dict__.pop(slot, None)  # the budget is returned when the cached tensor dies (weakref.finalize below)


def _cache_store(module, slot, key, value, w):
    if _no_cache():
        return
    import weakref

    _cache_drop(module, slot)
    n, dev = _nbytes(value), w.device
    t = _first_tensor(value)
    if t is not None and t.data_ptr() == w.data_ptr():
        n = 0  # a disabled quantizer returns the weight itself: nothing extra is held
    if w.is_cuda and n:
        if _CACHE_USED.get(dev, 0) + n > _cache_budget(dev):
            return  # over the budget: this weight is re-quantized per forward
        if _CACHE_BUDGET_GB is None:
            # and never into the last 15 % of the device (a 70B replica leaves ~37 GB of a B200: activations need most of it)
            free, total = torch.cuda.mem_get_info(dev)
            free += torch.cuda.memory_reserved(dev) - torch.cuda.memory_allocated(dev)
            if free - n < 0.15 * total:
                return
    if t is not None and n:
        _CACHE_USED[dev] = _CACHE_USED.get(dev, 0) + n
        weakref.finalize(t, _release, dev, n)  # dropped, overwritten or the module collected: the bytes come back
    module.__dict__[slot] = (key, value, _fingerprint(w)) if CHECK_CACHE else (key, value)


def _weight_get(self):
    w = self._parameters["weight"] if "weight" in self._parameters else self._buffers.get("weight")
    if w is None or not self.__dict__.get("_enable_weight_quantization", False):
        return w
    q = self.weight_quantizer
    if (isinstance(q, TensorQuantizer) and q._if_calib and not q._if_quant and not q._disabled
            and isinstance(q._calibrator, MaxCalibrator) and q.pre_quant_scale is None):
        key = (w.data_ptr(), w._version, tuple(w.shape))
        if not _no_cache() and self.__dict__.get("_weight_collect_key") == key and q._calibrator._shape is not None:
            return w  # already folded into the running amax; max is idempotent
        self.__dict__["_weight_collect_key"] = key
    return q(w)


def _weight_set(self, value):
    # nn.Module.__setattr__ routes Parameters to register_parameter before a class property is consulted; this is
    # only reached for plain tensors / None
    if value is None or isinstance(value, nn.Parameter):
        self.register_parameter("weight", value)
    else:
        self._parameters["weight"] = value


class QuantLinearConvBase(QuantInputBase):
    """Linear / conv modules: input and weight are quantized."""

    default_quant_desc_weight = QUANT_DESC_8BIT_PER_TENSOR
    # generated subclasses get:  weight = property(_weight_get, _weight_set)

    def _setup(self):
        super()._setup()
        self.weight_quantizer = TensorQuantizer(self.default_quant_desc_weight)
        self.__dict__["_enable_weight_quantization"] = False

    def invalidate_weight_cache(self):
        """Drop every cached function of the weight (fake-quantized weight, packed NVFP4 / FP8 operands, the "already
        collected" mark) and the quantizers' fp32 amax copies.  The caches key on (data_ptr, _version) of the weight and
        of the quantizer buffers; call this after writing through `.data` or a raw pointer, which moves neither."""
        for k in _WEIGHT_CACHES:
            if k == "_weight_collect_key":
                self.__dict__.pop(k, None)
            else:
                _cache_drop(self, k)
        for q in self.modules():
            if isinstance(q, TensorQuantizer):
                q.__dict__.pop("_amax_f32", None)

    @contextlib.contextmanager
    def quantize_weight(self):
        """Context in which ``self.weight`` reads as the quantized weight."""
        self.__dict__["_enable_weight_quantization"] = True
        try:
            yield
        finally:
            self.__dict__["_enable_weight_quantization"] = False

    def forward(self, input, *args, **kwargs):
        with self.quantize_weight():
            return super().forward(input, *args, **kwargs)

    @staticmethod
    def quantized_linear_fn(package, func_name, self, input, weight, *args, **kwargs):
        """quant_linear.py:44-53."""
        out = getattr(package, func_name)(self.input_quantizer(input), self.weight_quantizer(weight), *args, **kwargs)
        return self.output_quantizer(out)


class _QuantLinear(QuantLinearConvBase):
    """quant_linear.py:38-53.  On a B200, when no gradient is needed and the operands are bf16 / f16, the forward is
    input fake-quant kernel -> tcgen05 GEMM (csrc/gemm.cu) on the CACHED fake-quantized weight -> output quantizer:
    the reference re-quantizes the weight on every forward; here it is re-done only when the
    weight or the quantizer state changed (``TensorQuantizer.cache_key``)."""

    default_quant_desc_weight = QUANT_DESC_8BIT_LINEAR_WEIGHT_PER_ROW

    def _tcgen05_ok(self, input, args, kwargs) -> bool:
        if args or kwargs or not isinstance(input, torch.Tensor) or not input.is_cuda:
            return False
        w = self._parameters.get("weight")
        if w is None or w.dim() != 2 or type(self)._original_cls.forward is not nn.Linear.forward:
            return False
        if torch.is_grad_enabled() and (input.requires_grad or w.requires_grad
                                        or (self.bias is not None and self.bias.requires_grad)):
            return False
        pqs = getattr(self.input_quantizer, "_pre_quant_scale", None)
        if pqs is not None and pqs.dtype != input.dtype:
            return False
        from .. import ops

        return ops.linear_supported(input, w, self.bias)

    @staticmethod
    def _fused_input_format(q, input):
        """num_bits of a per-tensor FP8 / signed INT input quantizer in plain quant mode, else None."""
        from .. import ops

        if (not isinstance(q, TensorQuantizer) or q._disabled or not q._if_quant or q._if_calib or not q._fake_quant
                or q._dynamic or q.backend is not None or q.block_sizes is not None or q._use_constant_amax
                or q._unsigned or getattr(q, "_bias", None) is not None or not hasattr(q, "_amax")
                or q._amax.numel() != 1 or input.numel() // input.shape[-1] > ops.FUSED_PROLOGUE["max_rows"]):
            return None
        if q._num_bits == (4, 3):
            return (4, 3) if ops.FUSED_PROLOGUE["fp8"] else None
        if isinstance(q._num_bits, int) and 2 <= q._num_bits <= 16:
            return q._num_bits if ops.FUSED_PROLOGUE["int"] else None
        return None

    def quantized_weight(self):
        """weight_quantizer(weight), cached while neither the weight nor the quantizer state changes."""
        w = self._parameters["weight"]
        q = self.weight_quantizer
        qk = q.cache_key() if isinstance(q, TensorQuantizer) else None
        if qk is None:
            _cache_drop(self, "_wq_cache")
            with self.quantize_weight():
                return self.weight
        key = (w.data_ptr(), w._version, tuple(w.shape), w.dtype, qk)
        hit = _cache_lookup(self, "_wq_cache", key, w)
        if hit is not None:
            return hit
        with torch.no_grad():
            wq = q(w)
        _cache_store(self, "_wq_cache", key, wq, w)
        return wq

    # ---- opt-in: real-quant GEMMs (FP8 per-tensor, NVFP4 W4A4) on the tensor-core paths, not the bf16 simulation ----
    # `module.real_quant_gemm = True` or `enable_real_quant_gemm(model)` (backends.py; quant_linear.py:196-244)
    @property
    def real_quant_gemm(self):
        return bool(self.__dict__.get("_use_real_quant_gemm", False))

    @real_quant_gemm.setter
    def real_quant_gemm(self, value):
        self.__dict__["_use_real_quant_gemm"] = bool(value)
        self.__dict__.pop("_real_quant_gemm_impl", None)

    @property
    def _should_run_real_quant_gemm(self):
        """quant_linear.py:203-210."""
        iq = self.input_quantizer
        return self.real_quant_gemm and not (iq.is_enabled and iq._if_calib)

    @staticmethod
    def _is_nvfp4_dynamic_calibrated(q):
        return (isinstance(q, TensorQuantizer) and not q._disabled and q._if_quant and not q._if_calib
                and q.is_nvfp4_dynamic and q.backend is None and hasattr(q, "_amax") and q.pre_quant_scale is None
                and (q.block_sizes.get(-1) == 16))

    def _real_nvfp4_ok(self, input):
        w = self._parameters["weight"]
        return (self.real_quant_gemm and self._is_nvfp4_dynamic_calibrated(self.input_quantizer)
                and self._is_nvfp4_dynamic_calibrated(self.weight_quantizer) and not self.output_quantizer.is_enabled
                and w.shape[1] % 256 == 0 and w.shape[0] % 8 == 0)

    def packed_nvfp4_weight(self):
        """(codes, swizzled scales, global scale) of the weight, cached like ``quantized_weight``."""
        from .. import ops

        w = self._parameters["weight"]
        q = self.weight_quantizer
        key = (w.data_ptr(), w._version, tuple(w.shape), w.dtype, q.cache_key())
        hit = _cache_lookup(self, "_w4_cache", key, w)
        if hit is not None:
            return hit
        with torch.no_grad():
            packed = ops.nvfp4_quantize_operand(w.detach(), q._amax.float().amax(), 256)
        _cache_store(self, "_w4_cache", key, packed, w)
        return packed

    def _forward_real_quant(self, input):
        """Weight held packed (``mtq.compress``): dequantize with one kernel, then the tcgen05 GEMM (cuBLAS via
        ``F.linear`` for operands it does not take).  Nothing is cached: the point of compress is the memory."""
        from .. import ops

        w = self.weight_quantizer(self._parameters["weight"])  # QTensorWrapper -> dequantize
        x = self.input_quantizer(input)
        if w.dtype != x.dtype:
            w = w.to(x.dtype)
        if (not torch.is_grad_enabled() or not (x.requires_grad or (self.bias is not None and self.bias.requires_grad))) \
                and x.is_cuda and ops.linear_supported(x, w, self.bias):
            out = ops.linear(x, w, self.bias)
        else:
            out = torch.nn.functional.linear(x, w, self.bias)
        return self.output_quantizer(out)

    # ---- opt-in: FP8 per-tensor W8A8 on tcgen05 kind::f8f6f4 (backends/fp8_per_tensor_gemm.py) ----
    @staticmethod
    def _is_fp8_per_tensor_calibrated(q):
        return (isinstance(q, TensorQuantizer) and not q._disabled and q._if_quant and not q._if_calib and q.is_fp8
                and q.backend is None and hasattr(q, "_amax") and q._amax.numel() == 1 and q.pre_quant_scale is None)

    def _real_fp8_ok(self, input):
        w = self._parameters["weight"]
        return (self.real_quant_gemm and self._is_fp8_per_tensor_calibrated(self.input_quantizer)
                and self._is_fp8_per_tensor_calibrated(self.weight_quantizer) and not self.output_quantizer.is_enabled
                and w.shape[1] % 16 == 0 and w.shape[0] % 8 == 0 and input.dtype in (torch.bfloat16, torch.float16))

    def packed_fp8_weight(self):
        """(E4M3 codes, scale_b) of the weight (`_to_fp8(weight, amax / 448)`), cached like ``quantized_weight``."""
        from .. import ops

        w = self._parameters["weight"]
        q = self.weight_quantizer
        key = (w.data_ptr(), w._version, tuple(w.shape), w.dtype, q.cache_key())
        hit = _cache_lookup(self, "_w8_cache", key, w)
        if hit is not None:
            return hit
        with torch.no_grad():
            scale_b = ops.div_scalar(q._amax.float().reshape(1), 448.0)
            packed = (ops.q8_pack(w.detach(), scale_b, "fp8_sat"), scale_b)
        _cache_store(self, "_w8_cache", key, packed, w)
        return packed

    def forward(self, input, *args, **kwargs):
        if hasattr(self._parameters.get("weight"), "get_qtensor") and not args and not kwargs \
                and type(self)._original_cls.forward is nn.Linear.forward:
            return self._forward_real_quant(input)
        if self._tcgen05_ok(input, args, kwargs):
            from .. import ops

            if self._should_run_real_quant_gemm and input.numel() > 1:
                from ..backends import gemm_registry

                impl = gemm_registry.find_match(self, input)  # quant_linear.py:212-221
                if impl is not None:
                    return impl(self, input, self._parameters["weight"], self.bias)
            q = self.input_quantizer
            fmt = self._fused_input_format(q, input)
            if fmt is not None:  # input fake-quant inside the GEMM's operand prologue: ONE kernel for the linear
                out = ops.linear_qdq(input, self.quantized_weight(), self.bias, q._amax, fmt, q._narrow_range,
                                     prescale=q.pre_quant_scale)
            else:
                out = ops.plain_linear(q(input), self.quantized_weight(), self.bias)
            return self.output_quantizer(out)
        return super().forward(input, *args, **kwargs)


class RealQuantLinear(_QuantLinear):
    """quant_linear.py:196-244 under its own name.  In this package one class serves both roles — `_QuantLinear` runs the
    fake-quant linear, and with `real_quant_gemm` on (or a packed `QTensorWrapper` weight from `mtq.compress`) the
    real-quant one — so this is that class plus the reference's attribute names."""

    list_of_scale_tensors = ["_scale", "_double_scale", "_scale_zeros"]
    allow_real_quant_gemm = True

    @property
    def _should_run_real_quant_gemm(self):
        return super()._should_run_real_quant_gemm and self.allow_real_quant_gemm

    def has_real_quant_gemm_impl(self, input, *args, **kwargs) -> bool:
        """quant_linear.py:211-221: the registry's match for this call, cached on the module."""
        if "_real_quant_gemm_impl" not in self.__dict__:
            from ..backends import gemm_registry

            self.__dict__["_real_quant_gemm_impl"] = gemm_registry.find_match(self, input, *args, **kwargs)
            if self.__dict__["_real_quant_gemm_impl"] is None:
                warnings.warn(f"RealQuantLinear: No real-quant GEMM found: {self}.")
        return self.__dict__["_real_quant_gemm_impl"] is not None


class _QuantConv(QuantLinearConvBase):
    default_quant_desc_weight = QUANT_DESC_8BIT_LINEAR_WEIGHT_PER_ROW


class _Registry:
    """original class -> quantized base (QuantModuleRegistry of the reference: ``in``, ``[name]``, ``convert``)."""

    def __init__(self):
        self._by_cls: dict[type, tuple[str, type, Any]] = {}
        self._generated: dict[type, type] = {}

    def register(self, mapping: dict[type, str], pre_convert=None):
        def deco(base):
            for cls, name in mapping.items():
                self._by_cls[cls] = (name, base, pre_convert)
            return base
        return deco

    def unregister(self, cls):
        self._by_cls.pop(cls, None)
        self._generated.pop(cls, None)

    def __contains__(self, cls):
        return cls in self._by_cls

    def __getitem__(self, name: str):
        """Generated quantized class registered under ``name`` (e.g. "nn.Linear")."""
        for cls, (n, _, _) in self._by_cls.items():
            if n == name:
                return self.get(cls)
        raise KeyError(name)

    def get(self, cls):
        if cls not in self._generated:
            name, base, _ = self._by_cls[cls]
            ns: dict[str, Any] = {"_original_cls": cls, "__module__": __name__}
            if issubclass(base, QuantLinearConvBase):
                ns["weight"] = property(_weight_get, _weight_set)
            self._generated[cls] = type(f"Quant{cls.__name__}", (base, cls), ns)
        return self._generated[cls]

    @torch.no_grad()
    def convert(self, module: nn.Module) -> nn.Module:
        cls = type(module)
        _, _, pre = self._by_cls[cls]
        if pre is not None:
            return pre(module, self)
        module.__class__ = self.get(cls)
        module._setup()
        if module.parallel_state is None:
            module.parallel_state = ParallelState(data_parallel_group=None)
        return module


QuantModuleRegistry = _Registry()
QuantModuleRegistry.register({nn.Linear: "nn.Linear"})(_QuantLinear)
QuantModuleRegistry.register({nn.Conv1d: "nn.Conv1d", nn.Conv2d: "nn.Conv2d", nn.Conv3d: "nn.Conv3d",
                              nn.ConvTranspose1d: "nn.ConvTranspose1d", nn.ConvTranspose2d: "nn.ConvTranspose2d",
                              nn.ConvTranspose3d: "nn.ConvTranspose3d"})(_QuantConv)
QuantModuleRegistry.register({
    nn.LeakyReLU: "nn.LeakyReLU", nn.BatchNorm1d: "nn.BatchNorm1d", nn.BatchNorm2d: "nn.BatchNorm2d",
    nn.BatchNorm3d: "nn.BatchNorm3d", nn.InstanceNorm1d: "nn.InstanceNorm1d", nn.InstanceNorm2d: "nn.InstanceNorm2d",
    nn.InstanceNorm3d: "nn.InstanceNorm3d", nn.LayerNorm: "nn.LayerNorm", nn.MaxPool1d: "nn.MaxPool1d",
    nn.MaxPool2d: "nn.MaxPool2d", nn.MaxPool3d: "nn.MaxPool3d", nn.AvgPool1d: "nn.AvgPool1d",
    nn.AvgPool2d: "nn.AvgPool2d", nn.AvgPool3d: "nn.AvgPool3d", nn.AdaptiveAvgPool1d: "nn.AdaptiveAvgPool1d",
    nn.AdaptiveAvgPool2d: "nn.AdaptiveAvgPool2d", nn.AdaptiveAvgPool3d: "nn.AdaptiveAvgPool3d",
})(QuantInputBase)


class _QuantEmbedding(QuantModule):
    """quant_embedding.py:31-: only the weight is quantized (disabled by the default configs)."""

    def _setup(self):
        self.weight_quantizer = TensorQuantizer(QUANT_DESC_8BIT_PER_TENSOR)

    def forward(self, input, *args, **kwargs):
        w = self._parameters["weight"]
        self._parameters["weight"] = self.weight_quantizer(w)
        try:
            return self._original_cls.forward(self, input, *args, **kwargs)
        finally:
            self._parameters["weight"] = w


QuantModuleRegistry.register({nn.Embedding: "nn.Embedding"})(_QuantEmbedding)

try:  # HF GPT-2's Conv1D is a Linear with a transposed weight (plugins/huggingface.py:559-575)
    from transformers.pytorch_utils import Conv1D as _HFConv1D

    def _conv1d_to_linear(module, registry):
        module.weight = nn.Parameter(module.weight.T.contiguous())
        module.out_features, module.in_features = module.weight.shape
        module.__class__ = registry.get(nn.Linear)  # nn.Linear.forward, not Conv1D.forward
        module._setup()
        module.parallel_state = ParallelState(data_parallel_group=None)
        return module

    QuantModuleRegistry.register({_HFConv1D: "Conv1D"}, pre_convert=_conv1d_to_linear)(_QuantLinear)
except Exception:  # noqa: BLE001  (transformers absent or broken: nothing to register)
    pass
