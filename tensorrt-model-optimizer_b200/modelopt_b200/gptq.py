"""GPTQ on the B200 path (interface of modelopt/torch/quantization/utils/calib_utils.py:50-340 and
model_calib.py:2188-2272): Hessian collection, damped Cholesky inverse, blockwise column update.  With ``fused=True``
and a static NVFP4 weight quantizer the column loop of each block is ONE kernel (csrc/hessian.cu
``gptq_block_nvfp4_kernel``); the Hessian product, the Cholesky factorisations and the trailing update are plain library
GEMM / LAPACK calls through torch, as in the reference."""

from __future__ import annotations

import math
import time

import torch
from torch import nn

from . import ops
from .nn.tensor_quantizer import SequentialQuantizer, TensorQuantizer
from .utils import is_quantized_linear

__all__ = ["gptq", "GPTQHelper", "update_hessian", "compute_hessian_inverse", "gptq_blockwise_update",
           "gptq_blockwise_update_fused_scalar", "register_gptq_helper"]


def update_hessian(input, hessian, n_samples):
    """calib_utils.py:50-77: running mean of 2 X X^T over tokens."""
    x = input.reshape(-1, input.shape[-1]).t().float()
    bsz = x.shape[1]
    if bsz == 0:
        return hessian, n_samples
    hessian *= n_samples / (n_samples + bsz)
    n_samples += bsz
    x = math.sqrt(2 / n_samples) * x
    hessian.add_((x @ x.t()).to(hessian.device))
    return hessian, n_samples


def compute_hessian_inverse(hessian, weight, perc_damp):
    """calib_utils.py:80-113: upper Cholesky factor of the damped inverse Hessian (identity if not positive definite)."""
    h = hessian.clone()
    zero_cols = torch.nonzero(weight.eq(0).all(dim=0)).unsqueeze(-1)
    h[zero_cols, :] = 0
    h[:, zero_cols] = 0
    h[zero_cols, zero_cols] = 1
    damp = perc_damp * torch.mean(torch.diag(h))
    idx = torch.arange(h.shape[0], device=h.device)
    h[idx, idx] += damp
    try:
        h = torch.cholesky_inverse(torch.linalg.cholesky(h))
        return torch.linalg.cholesky(h, upper=True)
    except (RuntimeError, torch.linalg.LinAlgError):
        print("Warning: Hessian is not positive definite, using identity matrix")
        return torch.eye(h.shape[0], device=h.device, dtype=h.dtype)


def gptq_blockwise_update(weight, h_inv, block_size, quantize_fn):
    """calib_utils.py:252-288: generic column loop (one full-matrix fake-quant per column)."""
    num_cols = weight.shape[1]
    for bs in range(0, num_cols, block_size):
        be = min(bs + block_size, num_cols)
        hb = h_inv[bs:be, bs:be]
        wblk = weight.clone()
        errs = torch.zeros_like(weight[:, bs:be])
        for i in range(be - bs):
            w_ci = wblk[:, bs + i]
            qdq = quantize_fn(wblk)
            weight[:, bs + i] = qdq[:, bs + i]
            err = (w_ci - qdq[:, bs + i]) / hb[i, i]
            wblk[:, bs + i: be].addr_(err, hb[i, i:], alpha=-1)
            errs[:, i] = err
        weight[:, be:].addmm_(errs, h_inv[bs:be, be:], alpha=-1)


def gptq_blockwise_update_fused_scalar(weight, block_amax, global_scale, h_inv, block_size, quant_block_size):
    """calib_utils.py:291-327: one fused kernel per GPTQ block + one GEMM for the trailing columns."""
    num_cols = weight.shape[1]
    for bs in range(0, num_cols, block_size):
        be = min(bs + block_size, num_cols)
        qw, err = ops.gptq_block_nvfp4(weight[:, bs:be], block_amax, global_scale, h_inv[bs:be, bs:be], quant_block_size, bs)
        weight[:, bs:be] = qw
        if be < num_cols:
            weight[:, be:].addmm_(err, h_inv[bs:be, be:], alpha=-1)


_GPTQ_HELPER_REGISTRY: dict[str, type] = {}


def register_gptq_helper(backend: str, factory: type) -> None:
    _GPTQ_HELPER_REGISTRY[backend] = factory


class GPTQHelper:
    """calib_utils.py:116-249: per-module Hessian, forward patch during collection, blockwise weight update."""

    def __init__(self, module, name, offload_to_cpu=False, fused=False):
        self.module, self.name, self.fused = module, name, fused
        w = module._parameters["weight"]
        self.hessian = torch.zeros(w.shape[-1], w.shape[-1], dtype=torch.float32, device=w.device)
        self.n_samples = 0
        self.weight = None
        self.h_inv = None
        self._hook = None

    def setup(self):
        helper = self

        def pre_hook(mod, args):
            if not args:
                return
            x = args[0].to_local() if hasattr(args[0], "to_local") else args[0]
            iq = getattr(mod, "input_quantizer", None)
            hx = iq(x) if iq is not None and iq.is_enabled else x
            helper.hessian, helper.n_samples = update_hessian(hx, helper.hessian, helper.n_samples)

        self._hook = self.module.register_forward_pre_hook(pre_hook)

    def cleanup(self):
        if self._hook is not None:
            self._hook.remove()
            self._hook = None

    def free(self):
        self.hessian = self.weight = self.h_inv = None

    def update_weights(self, block_size, perc_damp):
        w = self.module._parameters["weight"]
        hessian = self.hessian.to(w.device)
        self.weight = w.data.float().clone()
        self.h_inv = compute_hessian_inverse(hessian, self.weight, perc_damp)
        self._blockwise_update(block_size)
        self.rel_mse = self._mse_error(hessian, w)
        w.data = self.weight.reshape(w.shape).to(w.dtype)
        self.module.invalidate_weight_cache()

    def _blockwise_update(self, block_size):
        q = self.module.weight_quantizer
        if self.fused and getattr(q, "is_nvfp4_static", False) and getattr(q, "global_amax", None) is not None:
            qb = q.block_sizes.get(-1) or q.block_sizes.get(1)
            if qb is not None and block_size % qb != 0:
                raise ValueError(f"GPTQ block_size ({block_size}) must be divisible by the quantizer group_size ({qb})")
            out_features, num_cols = self.weight.shape
            if block_size in (32, 64, 128, 256) and num_cols % block_size == 0:
                block_amax = q.amax.reshape(out_features, num_cols // qb).float()
                gs = q.global_amax.float().item() / (6.0 * 448.0)
                gptq_blockwise_update_fused_scalar(self.weight, block_amax, gs, self.h_inv, block_size, qb)
                return
        gptq_blockwise_update(self.weight, self.h_inv, block_size, q)

    def _mse_error(self, hessian, w):
        w_orig = w.float()
        delta = self.weight - w_orig
        return ((delta.mm(hessian).mul(delta).mean()) / (w_orig.mm(hessian).mul(w_orig).mean() + 1e-6)).item()


@torch.no_grad()
def gptq(model: nn.Module, forward_loop, perc_damp: float = 0.01, block_size: int = 128, fused: bool = False, **_):
    """model_calib.py:2188-2272: max-calibrate, collect per-linear Hessians from a forward with the weight quantizers
    disabled, then blockwise weight updates with the inverse Hessian."""
    from .model_calib import max_calibrate

    t0 = time.time()
    max_calibrate(model, forward_loop=forward_loop)
    layers = [(n, m) for n, m in model.named_modules()
              if is_quantized_linear(m) and isinstance(m.weight_quantizer, TensorQuantizer) and m.weight_quantizer.is_enabled]
    if not layers:
        print("No quantized linear layers found, skipping GPTQ")
        return
    handles = {}
    for name, m in layers:
        cls = _GPTQ_HELPER_REGISTRY.get(getattr(m.weight_quantizer, "backend", None), GPTQHelper)
        handles[name] = cls(m, name, offload_to_cpu=True, fused=fused)
        handles[name].setup()
    disabled = []
    for module in model.modules():
        q = getattr(module, "weight_quantizer", None)
        for sub in (q if isinstance(q, SequentialQuantizer) else [q]):
            if isinstance(sub, TensorQuantizer) and sub.is_enabled:
                sub.disable()
                disabled.append(sub)
    try:
        forward_loop(model)
    finally:
        for h in handles.values():
            h.cleanup()
        for sub in disabled:
            sub.enable()
    model._gptq_rel_mse = {}
    for name, h in handles.items():
        h.update_weights(block_size, perc_damp)
        model._gptq_rel_mse[name] = h.rel_mse
        h.free()
    print(f"GPTQ time: {time.time() - t0:.2f}s")
