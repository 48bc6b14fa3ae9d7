"""NVFP4QTensor on the B200 path (interface of modelopt/torch/quantization/qtensor/nvfp4_tensor.py:49-407 and
base_qtensor.py): packed E2M1 codes (two per byte, even element in the low nibble), E4M3 block scales, fp32 global
scale ``amax / (6 * 448)``.  ``quantize`` is one amax pass + ONE packing kernel (the reference: ~15 eager passes);
``dequantize`` is one kernel."""

from __future__ import annotations

import torch
import torch.nn.functional as F

from .. import ops

__all__ = ["NVFP4QTensor", "BaseQuantizedTensor", "QTensorWrapper"]


class BaseQuantizedTensor:
    """base_qtensor.py: original shape / dtype + the packed payload."""

    def __init__(self, original_shape, original_dtype, quantized_data):
        self.metadata = {"shape": original_shape, "dtype": original_dtype}
        self._quantized_data = quantized_data

    def get_qtensor(self):
        return self


class QTensorWrapper(torch.nn.Parameter):
    """A Parameter that carries a quantized tensor through ``module.weight`` (base_qtensor.py:87-145): the payload is
    the packed data, ``metadata`` keeps the original shape / dtype and the QTensor class."""

    def __new__(cls, qtensor, metadata=None):
        data = qtensor._quantized_data if isinstance(qtensor, BaseQuantizedTensor) else qtensor
        obj = torch.Tensor._make_subclass(cls, data, False)
        if metadata is None:
            obj.metadata = dict(qtensor.metadata)
            obj.metadata["qtensor_class"] = qtensor.__class__
        else:
            assert all(k in metadata for k in ("qtensor_class", "shape", "dtype")), f"metadata: {metadata}"
            obj.metadata = metadata
        return obj

    def dim(self):
        return len(self.metadata["shape"])

    def to(self, *args, **kwargs):
        device, dtype, *_ = torch._C._nn._parse_to(*args, **kwargs)
        data = self.data.to(device=device) if device else self.data
        return QTensorWrapper(self.metadata["qtensor_class"](self.metadata["shape"], dtype or self.metadata["dtype"], data))

    def get_qtensor(self):
        return self.metadata["qtensor_class"](self.metadata["shape"], self.metadata["dtype"], self.data)

    def get_state(self):
        return {"metadata": self.metadata, "quantized_data.shape": self.data.shape, "quantized_data.dtype": self.data.dtype}


class NVFP4QTensor(BaseQuantizedTensor):
    e2m1_bounds = torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5])
    e2m1_values = torch.tensor([0, 0.5, 1, 1.5, 2, 3, 4, 6, 0, -0.5, -1, -1.5, -2, -3, -4, -6])

    @classmethod
    def get_weights_scaling_factor_2(cls, input: torch.Tensor) -> torch.Tensor:
        """nvfp4_tensor.py:204-207: amax / (6 * 448) in fp32 (one read of the tensor)."""
        return ops.div_scalar(ops.reduce_amax(input).float(), 6.0 * 448.0)

    @classmethod
    def get_weights_scaling_factor(cls, input, block_size, weights_scaling_factor_2=None, keep_high_precision=False):
        """nvfp4_tensor.py:169-201: E4M3 per-block scales (and the global scale)."""
        if weights_scaling_factor_2 is None:
            weights_scaling_factor_2 = cls.get_weights_scaling_factor_2(input)
        if keep_high_precision:
            bam = ops.reduce_amax(input.reshape(*input.shape[:-1], -1, block_size), axis=-1, keepdims=False,
                                  squeeze_scalar=False).float()
            s = bam / (6.0 * weights_scaling_factor_2.to(bam.device))
            s[s == 0] = 1.0
            return s, weights_scaling_factor_2
        _, sf = ops.nvfp4_pack(input, block_size, weights_scaling_factor_2)
        return sf, weights_scaling_factor_2

    @classmethod
    def quantize(cls, input: torch.Tensor, block_size: int, weights_scaling_factor: torch.Tensor | None = None,
                 weights_scaling_factor_2: torch.Tensor | None = None, keep_high_precision: bool = False,
                 try_tensorrt: bool = False):
        """nvfp4_tensor.py:253-342.  Returns (NVFP4QTensor, e4m3 block scales, fp32 global scale)."""
        shape, dtype = input.shape, input.dtype
        pad = (-input.shape[-1]) % block_size
        if pad:
            input = F.pad(input, (0, pad))  # reduce_block_padding
        if weights_scaling_factor_2 is None:
            weights_scaling_factor_2 = cls.get_weights_scaling_factor_2(input)
        if keep_high_precision:
            raise NotImplementedError("modelopt_b200: keep_high_precision is not implemented")
        codes, sf = ops.nvfp4_pack(input, block_size, weights_scaling_factor_2, weights_scaling_factor)
        return cls(shape, dtype, codes), sf, weights_scaling_factor_2

    def dequantize(self, dtype: torch.dtype = None, fast=False, **kwarg):
        """nvfp4_tensor.py:344-407.  kwargs: scale (e4m3 block scales), double_scale (fp32), block_sizes {-1: n}."""
        dtype = dtype or self.metadata["dtype"]
        block = kwarg["block_sizes"][-1]
        scale = kwarg["scale"]
        if scale.dtype != torch.float8_e4m3fn:
            scale = scale.to(torch.float8_e4m3fn)
        out = ops.nvfp4_unpack(self._quantized_data, scale, kwarg["double_scale"], block, dtype, fast)
        shape = self.metadata["shape"]
        if out.shape[-1] != shape[-1]:
            out = out[..., : shape[-1]]
        return out.reshape(shape)
