"""TensorQuantizer on the B200 path.

Same attributes, buffers (``_amax``, ``_pre_quant_scale``, ``_global_amax``), methods and state machine as
modelopt/torch/quantization/nn/modules/tensor_quantizer.py:136-1734, so that ``QuantModule``s, calibration
algorithms and checkpoints written for the reference work unchanged; every arithmetic step is a launch of a
hand-written sm_100a kernel (modelopt_b200.tensor_quant / .calib), never Triton, a JIT extension or eager torch.

Not carried over (out of scope per SURVEY.md section 8): affine bias calibration, Hadamard rotation, ONNX export
symbolic functions, MX (E8M0-scaled) formats and real-quantized formats other than NVFP4.  Asking for one raises.
"""

from __future__ import annotations

import contextlib
import math
import warnings
from collections.abc import Callable
from typing import Any

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch import nn

from .. import calib
from ..config import QuantizerAttributeConfig
from ..tensor_quant import (DynamicBlockQuantizationFunction, FakeTensorQuantFunction, ScaledE4M3Function,
                            call_functional, dynamic_block_quant, fake_tensor_quant, scaled_e4m3,
                            static_blockwise_fp4_fake_quant)
from .. import ops
from ..utils import convert_quantization_axis_to_reduce_axis, reduce_amax

try:
    from torch.distributed.tensor import DTensor
except ImportError:  # pragma: no cover
    DTensor = None

__all__ = ["TensorQuantizer", "SequentialQuantizer", "GroupedQuantizer", "StaticBlockScaleQuantizer", "NVFP4StaticQuantizer",
           "HardDisabledTensorQuantizer", "register_quant_backend", "unregister_quant_backend",
           "is_registered_quant_backend"]

QuantBackendEntrypoint = Callable[[torch.Tensor, "TensorQuantizer"], torch.Tensor]
_QUANT_FUNCTIONAL_BACKENDS: dict[str, QuantBackendEntrypoint] = {}


def register_quant_backend(name: str, entrypoint: QuantBackendEntrypoint) -> None:
    """tensor_quantizer.py:92-109: ``entrypoint(inputs, tensor_quantizer) -> tensor`` selected by cfg ``backend``."""
    if not isinstance(name, str) or not name:
        raise ValueError("Backend name must be a non-empty string.")
    if not callable(entrypoint):
        raise TypeError("Entrypoint must be callable.")
    if name in _QUANT_FUNCTIONAL_BACKENDS:
        warnings.warn(f"Overwriting existing backend: {name}")
    _QUANT_FUNCTIONAL_BACKENDS[name] = entrypoint


def unregister_quant_backend(name: str) -> None:
    if not isinstance(name, str) or not name:
        raise ValueError("Backend name must be a non-empty string.")
    _QUANT_FUNCTIONAL_BACKENDS.pop(name, None)


def is_registered_quant_backend(name: str) -> bool:
    return name in _QUANT_FUNCTIONAL_BACKENDS


def _standardize_constructor_args(val):
    """(cls) | (cls, args) | (cls, args, kwargs) -> (cls, args, kwargs)."""
    if isinstance(val, type):
        return val, (), {}
    val = tuple(val)
    cls = val[0]
    args = tuple(val[1]) if len(val) > 1 else ()
    kwargs = dict(val[2]) if len(val) > 2 else {}
    return cls, args, kwargs


def _set_calibrator(q, val):
    if val in ("max", "histogram"):
        cls = calib.MaxCalibrator if val == "max" else calib.HistogramCalibrator
        args, kwargs = (q._num_bits, q._axis, q._unsigned), {}
    else:
        cls, args, kwargs = _standardize_constructor_args(val)
    return cls(*args, **kwargs)


def _set_axis(q, val):
    if getattr(q, "_calibrator", None) is not None:
        q._calibrator._axis = val
    return val


def _set_block_sizes(q, val):
    if val is not None:
        q._axis = None
        if getattr(q, "_calibrator", None) is not None:
            q._calibrator._axis = None
    return val


def _set_constant_amax(q, val):
    if val is not None:
        q.amax = float(val)
    return val


def _unsupported(name):
    def f(q, val):
        if val:
            raise NotImplementedError(f"modelopt_b200: quantizer attribute '{name}' is not implemented "
                                      "(outside the hot path, SURVEY.md section 8)")
        return val
    return f


# config field -> (attribute name, setter(quantizer, value)); every other field `x` is stored as `_x`
_ATTRIBUTE_SETTERS: dict = {
    "enable": ("_disabled", lambda q, v: v is False),
    "type": ("_dynamic", lambda q, v: v == "dynamic"),
    "calibrator": ("_calibrator", _set_calibrator),
    "axis": ("_axis", _set_axis),
    "block_sizes": ("_block_sizes", _set_block_sizes),
    "backend": ("backend", lambda q, v: v),
    "backend_extra_args": ("backend_extra_args", lambda q, v: v or {}),
    "use_constant_amax": ("_use_constant_amax", lambda q, v: v),
    "constant_amax": ("_constant_amax", _set_constant_amax),
    "bias": ("_bias", _unsupported("bias")),
    "rotate": ("_rotate", _unsupported("rotate")),
}


class TensorQuantizer(nn.Module):
    """Quantizes (fake or real) or calibrates its input, depending on ``_if_quant`` / ``_if_calib``."""

    _skip_properties_for_save_restore = {"_calibrator", "_bias_calibrator", "_original_shape", "_block_reshape_size",
                                         "_padding", "_quantizer_cache", "_slices", "_amax_f32"}

    def __init__(self, quant_attribute_cfg=None, if_quant=True, if_calib=False, amax=None):
        super().__init__()
        cfg = quant_attribute_cfg if quant_attribute_cfg is not None else QuantizerAttributeConfig()
        if amax is not None:
            self.amax = amax
        put = object.__setattr__  # plain Python values: no need for nn.Module.__setattr__'s Parameter / Module checks
        put(self, "_use_constant_amax", False)
        put(self, "_constant_amax", None)
        self.set_from_attribute_config(cfg)
        put(self, "_if_quant", if_quant)
        put(self, "_if_calib", if_calib)
        put(self, "_enable_pre_quant_scale", True)
        put(self, "_dequantize", False)
        put(self, "_input_dtype", None)
        put(self, "_bias_calibrator", None)
        put(self, "_quantizer_cache", None)

    # ------------------------------------------------------------------ configuration
    def set_from_attribute_config(self, attribute_cfg):
        """tensor_quantizer.py:228-289."""
        fields = QuantizerAttributeConfig.model_fields
        setattr_ = object.__setattr__
        for attribute, val in attribute_cfg.items():
            assert attribute in fields, f"{attribute} is not a valid `TensorQuantizer` attribute"
            entry = _ATTRIBUTE_SETTERS.get(attribute)
            # plain Python values (never a Parameter / Module / buffer): skip nn.Module.__setattr__'s type checks, which
            # dominate quantizer construction and set_quantizer_by_cfg on a 676-quantizer model (properties still apply)
            if entry is None:
                setattr_(self, "_" + attribute, val)
            else:
                setattr_(self, entry[0], entry[1](self, val))
        if self.is_mx_format:
            self._pass_through_bwd = True  # tensor_quantizer.py:289-290

    # ------------------------------------------------------------------ simple properties
    @property
    def num_bits(self):
        return self._num_bits

    @num_bits.setter
    def num_bits(self, value):
        self._num_bits = value
        self._calibrator._num_bits = value

    @property
    def maxbound(self):
        if self._num_bits == (4, 3):
            return 448.0
        if self._num_bits == (2, 1) and self._block_sizes.get("scale_bits") == (4, 3):
            return 6.0
        return (1 << (self._num_bits - 1 + int(self._unsigned))) - 1

    @property
    def unsigned(self):
        return self._unsigned

    @unsigned.setter
    def unsigned(self, value):
        self._unsigned = value
        self._calibrator._unsigned = value

    @property
    def axis(self):
        return self._axis

    @axis.setter
    def axis(self, value):
        self._axis = value
        self._calibrator._axis = value

    @property
    def block_sizes(self):
        return self._block_sizes

    @block_sizes.setter
    def block_sizes(self, value):
        self._axis = None
        self._block_sizes = value

    @property
    def fake_quant(self):
        return self._fake_quant

    @property
    def narrow_range(self):
        return self._narrow_range

    @narrow_range.setter
    def narrow_range(self, value):
        self._narrow_range = value

    @property
    def trt_high_precision_dtype(self):
        return self._trt_high_precision_dtype

    @trt_high_precision_dtype.setter
    def trt_high_precision_dtype(self, value):
        self._trt_high_precision_dtype = value

    @property
    def bias(self):
        return None

    @property
    def bias_calibrator(self):
        return None

    @property
    def is_enabled(self):
        return not self._disabled

    def disable(self):
        """Bypass the module: no calibration, no quantization."""
        self._disabled = True

    def enable(self):
        self._disabled = False

    @property
    def is_mx_format(self):
        bs = self.block_sizes
        return bs is not None and bs.get("type") == "dynamic" and bs.get("scale_bits") == (8, 0)

    @property
    def is_fp8(self):
        return self._num_bits == (4, 3) and self._block_sizes is None and self._axis is None

    @property
    def is_nvfp4_dynamic(self):
        bs = self._block_sizes
        return (bs is not None and bs.get("type") == "dynamic" and self._num_bits == (2, 1)
                and bs.get("scale_bits") == (4, 3))

    @property
    def is_nvfp4_static(self):
        bs = self._block_sizes
        return self.is_static_block_quant and self._num_bits == (2, 1) and bs is not None and bs.get("scale_bits") == (4, 3)

    @property
    def is_static_block_quant(self):
        bs = self.block_sizes
        return bs is not None and bs.get("type") != "dynamic" and self._fake_quant

    @property
    def rotate_is_enabled(self):
        return False

    # ------------------------------------------------------------------ buffers
    @property
    def pre_quant_scale(self):
        if not hasattr(self, "_pre_quant_scale") or not self._enable_pre_quant_scale:
            return None
        return self._pre_quant_scale

    @pre_quant_scale.setter
    def pre_quant_scale(self, value):
        assert value is not None, "pre_quant_scale cannot be set to None."
        assert self._enable_pre_quant_scale, "pre_quant_scale cannot be set when forward_with_pre_quant_scale is False."
        if not isinstance(value, torch.Tensor):
            value = torch.tensor(value)
        if not hasattr(self, "_pre_quant_scale"):
            self.register_buffer("_pre_quant_scale", value.clone().detach())
        else:
            if self._pre_quant_scale.shape != value.shape:
                raise RuntimeError("Changing shape when setting pre_quant_scale is not allowed.")
            with torch.no_grad():  # copy_ on the buffer itself: bumps _version, which the weight cache keys on
                self._pre_quant_scale.copy_(value.detach().to(self._pre_quant_scale.device))

    @property
    def amax(self):
        if not hasattr(self, "_amax") or self.is_mx_format:
            return None
        assert not self._dynamic, "Dynamic quantization does not have fixed amax"
        return self._amax

    @amax.setter
    def amax(self, value):
        assert value is not None, "amax cannot be set to None."
        if not isinstance(value, torch.Tensor):
            value = torch.tensor(value)
        self._amax_setter_helper(value)

    def _amax_setter_helper(self, value):
        if not hasattr(self, "_amax"):
            self.register_buffer("_amax", value.clone().detach())
        else:
            if self._amax.shape != value.shape:
                raise RuntimeError("Changing shape when setting amax is not allowed.")
            with torch.no_grad():  # copy_ on the buffer itself: bumps _version, which the weight cache keys on
                self._amax.copy_(value.detach().to(self._amax.device))

    def reset_amax(self):
        if hasattr(self, "_amax"):
            delattr(self, "_amax")
        self._calibrator.reset()

    @property
    def step_size(self):
        if not hasattr(self, "_amax"):
            warnings.warn("step_size is undefined under dynamic amax mode!")
            return None
        assert isinstance(self._num_bits, int), "Step size is not defined for non-integer quantization."
        return self._amax / (2.0 ** (self._num_bits - 1 + int(self._unsigned)) - 1.0)

    def _set_buffer(self, key, value):
        if hasattr(self, key):
            setattr(self, key, value)
        else:
            self.register_buffer(key, value)

    # ------------------------------------------------------------------ calibration switches
    def disable_calib(self):
        self._if_calib = False

    def enable_calib(self):
        if self._calibrator is None:
            raise ValueError("Calibrator was not created, cannot enable calibration.")
        if self._dynamic:  # dynamic quantization does not need calibration
            return
        self._if_calib = True

    def disable_quant(self):
        self._if_quant = False

    def enable_quant(self):
        self._if_quant = True

    def load_calib_amax(self, *args, **kwargs):
        """tensor_quantizer.py:693-720."""
        assert not self._dynamic, "Dynamic quantization does not need calibration."
        strict = kwargs.pop("strict", True)
        if getattr(self, "_calibrator", None) is None:
            raise RuntimeError("Calibrator not created.")
        calib_amax = self._calibrator.compute_amax(*args, **kwargs)
        if calib_amax is None:
            msg = "Calibrator returned None. This usually happens when calibrator hasn't seen any tensor."
            if not strict:
                warnings.warn(msg)
                warnings.warn("Set amax to NaN!")
                calib_amax = torch.tensor(math.nan)
            else:
                raise RuntimeError(msg + " Passing 'strict=False' to `load_calib_amax()` will ignore the error.")
        self.amax = calib_amax

    def collect(self, inputs) -> None:
        if not self._if_calib or self._dynamic:
            return
        self._calibrator.collect(inputs)

    def validate_attr(self, attr_value=None, attr_name="amax", raise_error=False, warn_error=False, name=""):
        attr_value = attr_value if attr_value is not None else getattr(self, attr_name, None)
        if attr_value is None or (isinstance(attr_value, torch.Tensor) and attr_value.is_meta):
            return True
        ok = bool(torch.all(attr_value >= 0)) and not bool(torch.any(torch.isinf(attr_value))) \
            and not bool(torch.any(torch.isnan(attr_value)))
        if ok:
            return True
        name = f"{name}." if name else ""
        msg = f"{name}{attr_name} contains invalid values: {attr_value}"
        if warn_error:
            warnings.warn(msg)
        if raise_error:
            raise ValueError(msg)
        return False

    def sync_amax_across_distributed_group(self, parallel_group=None):
        """tensor_quantizer.py:1373-1385 (per-quantizer all-reduce MAX; the calibration driver prefers the ONE
        arena-wide exchange of ``modelopt_b200.arena.AmaxArena.sync``)."""
        group = getattr(parallel_group, "group", parallel_group)
        if dist.is_available() and dist.is_initialized() and getattr(self, "_amax", None) is not None:
            try:
                dist.all_reduce(self._amax, op=dist.ReduceOp.MAX, group=group)
            except RuntimeError as e:
                warnings.warn(f"Failed to synchronize amax: {e}")

    @contextlib.contextmanager
    def disable_pre_quant_scale(self):
        was = self._enable_pre_quant_scale
        self._enable_pre_quant_scale = False
        try:
            yield
        finally:
            self._enable_pre_quant_scale = was

    # ------------------------------------------------------------------ quantization
    def _get_amax(self, inputs):
        """Buffer, or the dynamic per-call amax (tensor_quantizer.py:736-751)."""
        if self._use_constant_amax:
            return torch.tensor(torch.finfo(torch.float8_e4m3fn).max, device=inputs.device)
        if hasattr(self, "_amax"):
            amax = self._amax
            if amax.dtype != torch.float32 and amax.is_cuda and not ops.NO_CACHE:
                # the kernels take fp32 amax (tensor_quant_gpu.cu:83 `amax.to(float)`): keep the converted copy while the
                # buffer's (address, version counter) is unchanged instead of launching a cast kernel per forward
                key = (amax.data_ptr(), amax._version, inputs.device)
                hit = self.__dict__.get("_amax_f32")
                if hit is not None and hit[0] == key:
                    return hit[1]
                f32 = amax.detach().to(device=inputs.device, dtype=torch.float32)
                self.__dict__["_amax_f32"] = (key, f32)
                return f32
            if amax.device != inputs.device:
                amax = amax.to(inputs.device)
        else:
            red = convert_quantization_axis_to_reduce_axis(inputs, self._axis)
            amax = reduce_amax(inputs, axis=red, keepdims=True).detach()
        return amax.data

    def _fake_quantize(self, inputs):
        """tensor_quantizer.py:890-949: pick the functional by format; each is one sm_100a kernel."""
        if self.backend is not None:
            if self.backend not in _QUANT_FUNCTIONAL_BACKENDS:
                raise KeyError(f"Quant backend '{self.backend}' is not registered.")
            return _QUANT_FUNCTIONAL_BACKENDS[self.backend](inputs, self)
        amax = None if self.is_mx_format else self._get_amax(inputs)  # tensor_quantizer.py:898-900
        bs = self.block_sizes
        if bs is not None and bs.get("type", "static") == "dynamic":
            block_size = bs.get(-1, None) or bs.get(inputs.dim() - 1, None)
            if block_size is None:
                raise ValueError("block size for dynamic quantization not found.")
            return call_functional(DynamicBlockQuantizationFunction, inputs, block_size, amax, None, self._num_bits,
                                   bs.get("scale_bits", None), getattr(self, "_trt_high_precision_dtype", None),
                                   getattr(self, "_onnx_quantizer_type", None), self._pass_through_bwd)
        if isinstance(self._num_bits, tuple):
            E, M = self._num_bits  # noqa: N806
            return call_functional(ScaledE4M3Function, inputs, amax, None, E, M, self._trt_high_precision_dtype,
                                   self._pass_through_bwd)
        return call_functional(FakeTensorQuantFunction, inputs, amax, None, self._num_bits, self._unsigned,
                               self._narrow_range, self._trt_high_precision_dtype, self._pass_through_bwd,
                               bs.get(-1) if bs else None,
                               self.axis[0] if isinstance(self.axis, tuple) else self.axis)

    def _is_real_quantize_support(self):
        """tensor_quantizer.py:785-794."""
        return bool((self._num_bits == 4 and self._block_sizes) or (self._num_bits == (2, 1) and self._block_sizes)
                    or self._num_bits in ((4, 3), 8))

    def _real_quantize(self, inputs):
        """tensor_quantizer.py:796-888: returns the packed tensor and keeps the scales as buffers.  Every branch is one
        packing kernel (csrc/pack_nvfp4.cu, csrc/pack_int.cu, csrc/qdq_mx.cu)."""
        from ..qtensor import (FP8QTensor, INT4QTensor, INT8QTensor, MXFP4QTensor, MXFP8QTensor, NF4QTensor,
                               NVFP4QTensor)

        assert self._is_real_quantize_support(), "Real quantization not supported for this format."
        bs = self._block_sizes
        buffers = {}
        if bs and bs.get("scale_bits") == (8, 0) and bs.get("type") == "dynamic":
            if self._num_bits == (2, 1):
                q, buffers["_scale"] = MXFP4QTensor.quantize(inputs, bs[-1])
            elif self._num_bits == (4, 3):
                assert bs[-1] == MXFP8QTensor.BLOCK_SIZE, f"MXFP8 requires block size {MXFP8QTensor.BLOCK_SIZE}, got {bs[-1]}"
                q, buffers["_scale"] = MXFP8QTensor.quantize(inputs)
            else:
                raise ValueError(f"Unsupported MX format: num_bits={self._num_bits}. "
                                 "Expected (2, 1) for MXFP4 or (4, 3) for MXFP8.")
        elif self._num_bits == (4, 3):
            use_amax = self.amax is not None and not (bs and self.amax.numel() == 1)
            q, buffers["_scale"] = FP8QTensor.quantize(inputs, axis=self._axis, block_sizes=bs,
                                                      scales=ops.div_scalar(self.amax, 448.0) if use_amax else None)
        elif self._num_bits == 8:
            q, buffers["_scale"] = INT8QTensor.quantize(inputs, axis=self._axis, block_sizes=bs)
        elif bs.get("scale_bits", 0) == 8 and bs.get("scale_block_sizes", None):
            q, scales = NF4QTensor.quantize(inputs, bs[-1], bs["scale_block_sizes"][-1])
            buffers["_scale"], buffers["_double_scale"], buffers["_scale_zeros"] = NF4QTensor.double_quantization(
                scales, bs["scale_block_sizes"][-1], bs["scale_bits"])
        elif bs.get("scale_bits") == (4, 3):
            if bs.get("four_over_six", False):
                raise NotImplementedError("NVFP4 Four-Over-Six (4/6) is not supported via mtq.compress")
            block = bs.get(-1, None) or bs.get(inputs.dim() - 1)
            sf2 = None
            if hasattr(self, "_amax"):
                sf2 = (self._amax.float().amax() / (6.0 * 448.0)).to(inputs.device)
            q, buffers["_scale"], buffers["_double_scale"] = NVFP4QTensor.quantize(inputs, block, weights_scaling_factor_2=sf2)
        else:
            q, buffers["_scale"] = INT4QTensor.quantize(inputs, bs[-1])
        for k, v in buffers.items():
            self._set_buffer(k, v)
        self._dequantize = True
        return q  # a BaseQuantizedTensor; pack_real_quantize_weight wraps it into a QTensorWrapper parameter

    def dequantize(self, inputs):
        q = inputs.get_qtensor() if hasattr(inputs, "get_qtensor") else inputs
        return q.dequantize(scale=self._scale, double_scale=getattr(self, "_double_scale", None),
                            scale_zeros=getattr(self, "_scale_zeros", None), block_sizes=self.block_sizes)

    # ------------------------------------------------------------------ static block quantization: reshape
    def _block_sizes_to_axis(self, x):
        """tensor_quantizer.py:1063-1085: block_sizes with all-None sizes is per-channel quantization."""
        bs = self._block_sizes
        if bs is None:
            return
        if all(v is None for k, v in bs.items() if isinstance(k, int)):
            assert self.axis is None, "Axis and block_sizes are both set."
            axis = tuple(k if k >= 0 else k + x.dim() for k in bs if isinstance(k, int))
            self.axis = tuple(i for i in range(x.dim()) if i not in axis) or None
            self._block_sizes = None

    def _setup_for_blockquant(self, inputs):
        """tensor_quantizer.py:975-1043: [.., N] -> [-1, block] (last-axis blocks) or the general nd split."""
        if hasattr(self, "_block_reshape_size"):
            return
        bs = self.block_sizes
        nd = inputs.dim()

        def axis_params(ax):
            ax = ax if ax in bs else ax - nd
            b = bs.get(ax, None)
            pad, sl = None, None
            if b is not None and inputs.shape[ax] % b != 0:
                pad = (b - (inputs.shape[ax] % b), 0)
                sl = slice(inputs.shape[ax])
            return b, pad, sl

        def commit(axis, reshape, padding, slices, amax_shape=None):
            self._axis = tuple(axis)
            if hasattr(self, "_calibrator"):
                self._calibrator._axis = self._axis
            self._original_shape = inputs.shape
            self._block_reshape_size = torch.Size(reshape)
            if padding is not None:
                self._padding = tuple(padding)
                self._original_shape = F.pad(inputs, self._padding, "constant", 0).shape
            if slices is not None:
                self._slices = slices
            if amax_shape:
                self._amax_shape_for_export = amax_shape

        block_axes = QuantizerAttributeConfig._get_block_quant_axes_and_sizes(bs)
        if (bs.get(nd - 1, None) or bs.get(-1, None)) and len(block_axes) == 1:
            b, pad, sl = axis_params(nd - 1)
            slices = None if sl is None else (*(slice(None),) * (nd - 1), sl)
            pad = pad if not pad else tuple(reversed(pad))
            commit((0,), (-1, b), pad, slices, (*inputs.shape[:-1], -1))
            return
        reshape, is_q, pads, slices = [], [], [], []
        for ax in range(nd):
            b, pad, sl = axis_params(ax)
            pads.append(pad)
            slices.append(sl)
            if b is not None:
                reshape += [math.ceil(inputs.shape[ax] / b), b]
                is_q += [True, False]
            else:
                reshape.append(inputs.shape[ax])
                is_q.append(True)
        slices = None if all(s is None for s in slices) else [s or slice(None) for s in slices]
        if all(p is None for p in pads):
            padding = None
        else:
            flat = []
            for p in pads:
                if not (flat or p):
                    continue
                flat.extend(p or (0, 0))
            padding = tuple(reversed(flat))
        commit([i for i, q in enumerate(is_q) if q], reshape, padding, slices)

    def _process_for_blockquant(self, inputs):
        if hasattr(self, "_padding"):
            inputs = F.pad(inputs, self._padding, "constant", 0)
        if inputs.shape != self._original_shape:
            raise ValueError(f"Input shape has changed from {self._original_shape} to {inputs.shape}."
                             " Block-quantization requires a fixed input shape.")
        return inputs.reshape(self._block_reshape_size)

    def _reset_to_original_shape(self, outputs):
        outputs = outputs.reshape(self._original_shape)
        if hasattr(self, "_slices"):
            outputs = outputs[self._slices]
        return outputs

    def export_amax(self):
        """tensor_quantizer.py:1087-1117."""
        if self.block_sizes is not None and self.block_sizes.get("type", None) == "dynamic":
            return self.amax
        if self.amax is None:
            return None
        amax = self.amax if not hasattr(self, "_amax_shape_for_export") else self.amax.reshape(self._amax_shape_for_export)
        amax[amax == 0] = self.maxbound
        amax = torch.nan_to_num(amax, nan=self.maxbound)
        amax = amax.clamp(min=torch.finfo(amax.dtype).tiny, max=torch.finfo(amax.dtype).max)
        self.validate_attr(attr_name="_amax", attr_value=amax)
        if self.block_sizes is None:
            if self.axis is None:
                amax = amax.unsqueeze(0)
            elif isinstance(self.axis, int) or (isinstance(self.axis, (list, tuple)) and len(self.axis) == 1):
                amax = amax.squeeze()
        return amax

    # ------------------------------------------------------------------ forward
    def forward(self, inputs):
        """tensor_quantizer.py:1119-1221."""
        if DTensor is not None and isinstance(inputs, DTensor):
            mesh, placements = inputs.device_mesh, inputs.placements
            return DTensor.from_local(self.forward(inputs.to_local()), mesh, placements)
        if hasattr(inputs, "get_qtensor") or type(inputs).__name__.endswith("QTensor"):
            assert self._dequantize, "No dequantization stats in the tensor quantizer."
            return self.dequantize(inputs)
        if inputs.numel() == 0:
            return inputs
        if self.pre_quant_scale is not None:
            fused = self._forward_fused_pre_quant_scale(inputs)
            if fused is not None:
                return fused
            inputs = self._apply_pre_quant_scale(inputs)
        if self._disabled:
            self._input_dtype = inputs.dtype if hasattr(inputs, "dtype") else None
            return inputs
        if self.block_sizes is not None and self._fake_quant:
            self._block_sizes_to_axis(inputs)
        if self.is_static_block_quant:
            self._setup_for_blockquant(inputs)
            inputs = self._process_for_blockquant(inputs)
        outputs = inputs
        if self._if_calib and not self._dynamic:
            if self._calibrator is None:
                raise RuntimeError("Calibrator was not created.")
            self.collect(inputs)
        if self._if_quant:
            if hasattr(inputs, "is_contiguous") and not inputs.is_contiguous():
                inputs = inputs.contiguous()
            if self.fake_quant:
                outputs = self._fake_quantize(inputs)
            elif not self._dequantize:
                outputs = self._real_quantize(inputs)
            else:
                raise ValueError("self._dequantize is True and self.fake_quant is False. "
                                 "This case should have been handled.")
        if self.is_static_block_quant:
            outputs = self._reset_to_original_shape(outputs)
        return outputs

    def cache_key(self):
        """Everything the fake-quantized value of a tensor depends on beside the tensor itself; None when the
        quantizer is collecting statistics or its result may not be reused."""
        if self._if_calib or self._dynamic or self.backend is not None:
            return None
        bufs = []
        for name in ("_amax", "_pre_quant_scale", "_global_amax", "_bias_value"):
            t = getattr(self, name, None)
            bufs.append(None if t is None else (t.data_ptr(), t._version, tuple(t.shape)))
        bs = self._block_sizes
        return (self._disabled, self._if_quant, self._fake_quant, self._num_bits, self._unsigned, self._narrow_range,
                self._axis if not isinstance(self._axis, list) else tuple(self._axis),
                None if bs is None else tuple(sorted(((str(k), str(v)) for k, v in bs.items()))),
                self._enable_pre_quant_scale, self._use_constant_amax, tuple(bufs))

    def _apply_pre_quant_scale(self, inputs):
        """inputs * pre_quant_scale (tensor_quantizer.py:1143); one mo_scale_cols_amax launch when no gradient is
        needed and the scale is a last-dim vector of the input dtype."""
        pqs = self._pre_quant_scale
        if (inputs.is_cuda and not (torch.is_grad_enabled() and (inputs.requires_grad or pqs.requires_grad))
                and pqs.dtype == inputs.dtype and pqs.numel() == inputs.shape[-1] and pqs.numel() > 1
                and inputs.dtype in (torch.float32, torch.float16, torch.bfloat16)):
            from .. import ops

            return ops.scale_cols(inputs.contiguous(), pqs)
        return inputs * pqs

    def _forward_fused_pre_quant_scale(self, inputs):
        """`inputs * pre_quant_scale` folded into the fake-quant kernel (one pass instead of two, row a5): taken when
        no gradient is needed and the format is INT-N / FP8 with a stored amax, or INT-N over last-dim blocks (stored
        or per-call amax).  Returns None when the general path must run."""
        from .. import ops

        pqs = self._pre_quant_scale
        if (self._disabled or not self._if_quant or self._if_calib or not self._fake_quant or self.backend is not None
                or not inputs.is_cuda or (torch.is_grad_enabled() and inputs.requires_grad)
                or pqs.dtype != inputs.dtype or pqs.numel() != inputs.shape[-1] or inputs.dtype == torch.float64
                or self._use_constant_amax or self._dynamic or getattr(self, "_bias", None) is not None):
            return None
        bs = self.block_sizes
        is_int = isinstance(self._num_bits, int)
        if bs is None:
            if not hasattr(self, "_amax"):
                return None
            amax = self._get_amax(inputs)
            try:
                if is_int:
                    return ops.qdq_int(inputs, amax, self._num_bits, self._unsigned, self._narrow_range, prescale=pqs)
                if self._num_bits == (4, 3):
                    return ops.qdq_fp8(inputs, amax, prescale=pqs)
            except ValueError:  # amax is not a scalar / single-axis vector
                return None
            return None
        keys = {k for k in bs if isinstance(k, int)}
        block = bs.get(-1) or bs.get(inputs.dim() - 1)
        if (not is_int or bs.get("type", "static") != "static" or len(keys) != 1 or not block
                or inputs.shape[-1] % block != 0):
            return None
        x = inputs.contiguous()
        if hasattr(self, "_amax"):
            amax = self._get_amax(inputs)
            if amax.numel() != x.numel() // block:
                return None
            y = ops.qdq_int(x.view(-1, block), amax.reshape(-1, 1), self._num_bits, self._unsigned,
                            self._narrow_range, prescale=pqs)
            return y.view(inputs.shape)
        return ops.qdq_int_block_dynamic(x, block, self._num_bits, self._unsigned, self._narrow_range, prescale=pqs)

    # ------------------------------------------------------------------ repr / state
    def _short_amax(self, fmt=".2e"):
        if not hasattr(self, "_amax"):
            return "dynamic"
        if self._amax is None:
            return "None"
        if self._amax.is_meta:
            return "meta"
        if self._amax.numel() == 1:
            return f"{self._amax.item():{fmt}}"
        return f"[{self._amax.min().item():{fmt}}, {self._amax.max().item():{fmt}}]({self._amax.numel()})"

    def extra_repr(self):
        if self._disabled:
            return "disabled"
        nb = self._num_bits if isinstance(self._num_bits, int) else f"({self._num_bits[0]}, {self._num_bits[1]})"
        s = f"{'unsigned ' if self._unsigned else ''}{nb} bit"
        s += " narrow" if self._narrow_range else ""
        s += " fake" if self._fake_quant else " real"
        if self.block_sizes is not None:
            s += f" block_sizes={self._block_sizes},"
        else:
            s += f" axis={self._axis}" if self._axis is not None else " per-tensor"
        s += f" amax={self._short_amax()}"
        s += " pre_quant_scale" if self.pre_quant_scale is not None else ""
        s += f" calibrator={self._calibrator.__class__.__name__}" if self._calibrator is not None else ""
        s += " quant" if self._if_quant else ""
        s += " calib" if self._if_calib else ""
        return s

    def _get_properties_for_modelopt_state(self):
        return self.__dict__.keys() - nn.Module().__dict__.keys() - self._skip_properties_for_save_restore

    def get_modelopt_state(self, properties_only: bool = False) -> dict[str, Any]:
        state = {k: getattr(self, k) for k in self._get_properties_for_modelopt_state()}
        if properties_only:
            return state
        meta: dict[str, Any] = {"params": {}, "buffers": {}}
        for k, v in self._parameters.items():
            meta["params"][k] = {"shape": v.shape, "dtype": v.dtype}
        for k, v in self._buffers.items():
            if k not in self._non_persistent_buffers_set:
                meta["buffers"][k] = {"shape": v.shape, "dtype": v.dtype}
        state["_pytorch_state_metadata"] = meta
        return state

    def set_from_modelopt_state(self, modelopt_state, properties_only: bool = False):
        for key in modelopt_state.keys() - self._skip_properties_for_save_restore - {"_pytorch_state_metadata"}:
            setattr(self, key, modelopt_state[key])
        for key in ("_num_bits", "_axis", "_unsigned"):
            setattr(self._calibrator, key, getattr(self, key))
        if not properties_only:
            meta = modelopt_state.get("_pytorch_state_metadata", {})
            self._parameters.clear()
            self._buffers.clear()
            self._non_persistent_buffers_set.clear()
            for k, v in meta.get("params", {}).items():
                self.register_parameter(k, nn.Parameter(torch.empty(v["shape"], dtype=v.get("dtype"))))
            for k, v in meta.get("buffers", {}).items():
                self.register_buffer(k, torch.empty(v["shape"], dtype=v.get("dtype")))


class HardDisabledTensorQuantizer(TensorQuantizer):
    """tensor_quantizer.py:1416-1457: a quantizer that can never be enabled."""

    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        self._disabled = True

    def enable(self):
        pass

    def enable_quant(self):
        pass

    def enable_calib(self):
        pass

    def set_from_attribute_config(self, attribute_cfg):
        super().set_from_attribute_config(attribute_cfg)
        self._disabled = True


class SequentialQuantizer(nn.Sequential):
    """Several quantizers applied in sequence (e.g. W4A8: INT4 block then FP8), tensor_quantizer.py:1737-."""

    _delegated = ("disable", "enable", "disable_quant", "enable_quant", "disable_calib", "enable_calib",
                  "load_calib_amax", "reset_amax")

    def __init__(self, *quantizers):
        assert quantizers and all(isinstance(q, TensorQuantizer) for q in quantizers), \
            "All quantizers must be a TensorQuantizer."
        super().__init__(*quantizers)

    def __getattr__(self, name):
        if name in SequentialQuantizer._delegated:
            def call(*a, **k):
                return [getattr(q, name)(*a, **k) for q in self]
            return call
        return super().__getattr__(name)

    def get_modelopt_state(self, properties_only=False):
        return {"num_quantizers": len(self), "is_sequential_quantizer": True}

    def set_from_attribute_config(self, attributes):
        if not isinstance(attributes, (list, tuple)):
            attributes = [attributes] * len(self)
        for a, q in zip(attributes, self):
            q.set_from_attribute_config(a)

    @staticmethod
    @contextlib.contextmanager
    def convert_to_single_quantizer(model, indx: int = 0):
        """Temporarily replace every SequentialQuantizer of ``model`` by its ``indx``-th member (AWQ search)."""
        swapped = []
        for _, module in list(model.named_modules()):
            for name, child in list(module.named_children()):
                if isinstance(child, SequentialQuantizer):
                    setattr(module, name, child[indx])
                    swapped.append((module, name, child))
        try:
            yield
        finally:
            for module, name, seq in swapped:
                setattr(module, name, seq)


class GroupedQuantizer(nn.ModuleList):
    """One quantizer per weight of a grouped linear (fused experts: ``grouped[i](weight_i)``), tensor_quantizer.py:
    1865-1892.  Unlike SequentialQuantizer the members act on DIFFERENT tensors; ``forward`` applies the first (the
    single-weight compatibility path), ``amax`` / ``is_enabled`` / ``fake_quant`` read the first member and writes go to
    all, lifecycle methods run on every member."""

    _props = ("fake_quant", "is_enabled", "amax")
    _methods = ("reset_amax", "disable", "enable", "load_calib_amax", "disable_quant", "enable_quant", "disable_calib",
                "enable_calib")

    def __init__(self, *quantizers):
        super().__init__(quantizers)
        assert len(self) and all(isinstance(q, (TensorQuantizer, SequentialQuantizer)) for q in self), \
            "All quantizers must be a TensorQuantizer or SequentialQuantizer."

    def forward(self, inputs):
        return self[0](inputs)

    def __getattr__(self, name):
        if name in GroupedQuantizer._props:
            return getattr(self[0], name)
        if name in GroupedQuantizer._methods:
            def call(*a, **k):
                return [getattr(q, name)(*a, **k) for q in self]
            return call
        return super().__getattr__(name)

    def __setattr__(self, name, value):
        if name in GroupedQuantizer._props:
            for q in self:
                setattr(q, name, value)
        else:
            super().__setattr__(name, value)

    def set_from_attribute_config(self, attributes):
        if not isinstance(attributes, (list, tuple)):
            attributes = [attributes] * len(self)
        for a, q in zip(attributes, self):
            q.set_from_attribute_config(a)


class StaticBlockScaleQuantizer(TensorQuantizer):
    """Static block quantization with two-level scaling (tensor_quantizer.py:1484-1734): ``_amax`` holds the
    per-block amax and ``_global_amax`` the tensor-wide one, both kept in fp32; NVFP4 goes through
    ``static_blockwise_fp4_fake_quant`` (one kernel for the E4M3 block scales, one for the E2M1 QDQ)."""

    _quant_max_bound: float = 6.0

    def _preserve_amax_in_fp32(self):
        for name in ("_amax", "_global_amax"):
            t = getattr(self, name, None)
            if t is not None and not isinstance(t, nn.Parameter) and t.dtype != torch.float32:
                setattr(self, name, t.to(dtype=torch.float32))

    def _amax_setter_helper(self, value):
        super()._amax_setter_helper(value)
        self._preserve_amax_in_fp32()

    @property
    def global_amax(self):
        return getattr(self, "_global_amax", None)

    @global_amax.setter
    def global_amax(self, value):
        if not isinstance(value, torch.Tensor):
            value = torch.tensor(value)
        self._set_buffer("_global_amax", value.clone().detach().to(torch.float32))

    @classmethod
    def from_tensor_quantizer(cls, tq: TensorQuantizer, global_amax=None):
        """In-place promotion of a calibrated static-block TensorQuantizer."""
        if not isinstance(tq, cls):
            nv = getattr(tq, "is_nvfp4_static", False)
            tq.__class__ = cls
            tq._is_static_block_scale_quantizer = True
            if nv:
                tq._is_nvfp4_static_quantizer = True
            tq._quant_max_bound = float(tq.maxbound)
        tq._preserve_amax_in_fp32()
        if global_amax is not None:
            tq.global_amax = global_amax
        return tq

    def _fake_quantize(self, inputs):
        if self.amax is not None and self.is_nvfp4_static:
            return static_blockwise_fp4_fake_quant(inputs, self.amax, self.global_amax, True, 448.0, inputs.dtype,
                                                   self._pass_through_bwd)
        return super()._fake_quantize(inputs)


NVFP4StaticQuantizer = StaticBlockScaleQuantizer
