"""CPU restatement of the reference's local-Hessian pieces.  TEST INFRASTRUCTURE ONLY.

* ``block_hessian``: `_LocalHessianAccumulator.accumulate` (modelopt/torch/quantization/model_calib.py:848-862).
* ``hessian_sweep``: Triton `_fp8_scale_sweep_hessian_kernel` + its host wrapper (modelopt/torch/kernels/quantization/gemm/
  nvfp4_fp8_sweep.py:173-290): candidates = the 126 positive finite E4M3 values / 448, candidate scale =
  compute_fp4_scales(c * global_amax) (gemm/fp4_kernel.py:217-251), dw = w - quant(w), loss = dw^T H dw, first minimum.

Parity: pinned against the reference's Triton kernel and accumulator run on a B200 (tests/golden/hessian_b200.npz,
oracle/gen_golden_gpu_mx.py).  The kernel's tl.dot summation order is not specified, so near-ties may pick another
candidate: the tests require identical choices in >= 99 % of blocks and, in fp64, an objective within 1e-5 of the chosen
minimum for the rest."""

from __future__ import annotations

import torch

from .fake_quant import compute_fp4_scales, e2m1_round_magnitude
from .mse_sweep import fp8_scale_candidates  # the 126 positive finite E4M3 values / 448


def block_hessian(x: torch.Tensor, block: int = 16, dtype=torch.float64) -> torch.Tensor:
    cin = x.shape[-1]
    xt = x.reshape(-1, cin).to(dtype).T.reshape(cin // block, block, -1)
    return xt @ xt.transpose(-1, -2)


def hessian_sweep(w: torch.Tensor, global_amax: torch.Tensor, hessian: torch.Tensor, block: int = 16, dtype=torch.float32):
    """IEEE-division restatement (= the product's MO_NUMERICS_CPU).  Returns (best_amax [n_blocks], losses [126, n_blocks])."""
    n_cin = hessian.shape[0]
    wb = w.detach().float().reshape(-1, n_cin, block)
    ga = torch.as_tensor(global_amax, dtype=torch.float32).reshape(())
    amax = fp8_scale_candidates() * ga                       # candidate_amaxes (:263)
    scales = compute_fp4_scales(amax, ga, True)              # candidate_scales (:264-266)
    h = hessian.to(dtype)
    wa = wb.abs()
    sign = torch.where(wb >= 0, 1.0, -1.0)                   # the sign does not cancel under the cross terms (:193-196)
    losses = []
    for k in range(scales.numel()):
        sc = scales[k] if scales[k] != 0 else torch.tensor(1.0)
        dw = (sign * (wa - e2m1_round_magnitude(wa / sc) * sc)).to(dtype)
        losses.append(torch.einsum("cnb,nbd,cnd->cn", dw, h, dw).reshape(-1))
    losses = torch.stack(losses)
    return amax[losses.argmin(dim=0)], losses                # torch.argmin: first minimum


def gptq_block_nvfp4(w_block, block_amax, global_scale, h_inv_blk, quant_block, block_start):
    """Torch restatement of the Triton _gptq_scalar_kernel (gemm/gptq_fused_kernel.py:43-94) with IEEE divisions and an
    unfused multiply-subtract (= the product's MO_NUMERICS_CPU).  Returns (qw, err), fp32 [rows, block]."""
    w = w_block.detach().float().clone()
    rows, bsz = w.shape
    gs = torch.tensor(float(global_scale), dtype=torch.float32)
    qw, err = torch.empty_like(w), torch.empty_like(w)
    for col in range(bsz):
        bam = block_amax[:, (block_start + col) // quant_block].float()
        scale = torch.minimum(bam / (6.0 * gs), torch.tensor(448.0)).to(torch.float8_e4m3fn).float() * gs  # fp8_quantize_scale
        zero = scale == 0
        safe = torch.where(zero | torch.isnan(scale) | torch.isinf(scale), torch.ones_like(scale), scale)
        ws = w[:, col]
        q = e2m1_round_magnitude(ws.abs() / safe) * safe
        q = torch.where(zero, torch.zeros_like(q), torch.where(ws >= 0, q, -q))
        e = (ws - q) / h_inv_blk[col, col]
        qw[:, col], err[:, col] = q, e
        if col + 1 < bsz:
            w[:, col + 1:] = w[:, col + 1:] - e[:, None] * h_inv_blk[col, col + 1:][None, :]
    return qw, err
