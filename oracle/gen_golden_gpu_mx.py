"""Capture outputs of the REAL reference's MX / INT4 / NF4 CUDA extensions and its qtensor classes on CUDA tensors, on a
B200.  TEST INFRASTRUCTURE ONLY.  Recipe (build container -> GPU box -> back):

    python /tmp/build_ref_ext.py ...        # torch.utils.cpp_extension.load of the reference's .cu files (arch 10.0, the MX one
                                            # with --use_fast_math as quantization/extensions.py:56-72 does) into
                                            # oracle/_ref/ext_mx, oracle/_ref/ext_main  (git-ignored, travel with gpurun)
    cp -r /root/reference/modelopt /root/reference/modelopt_recipes oracle/_ref/
    gpurun -- 'python oracle/gen_golden_gpu_mx.py'
    cp gpurun_out/mx_b200.npz gpurun_out/qtensor_b200.npz tests/golden/
"""

from __future__ import annotations

import importlib.util
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path[:0] = [os.path.join(HERE, "_ref"), os.path.join(HERE, "ref_shims"), ROOT,
                os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
sys.dont_write_bytecode = True
warnings.filterwarnings("ignore")

from oracle.gen_golden import DTYPES, to_np  # noqa: E402
from oracle.gen_golden_qtensor import base_tensor  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out")


def load_ext(sub, name):
    path = os.path.join(HERE, "_ref", sub, name + ".so")
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def mx_input():
    torch.manual_seed(5)
    x = torch.randn(64, 128) * torch.rand(64, 1) * 8
    x[0] = 0.0
    x[1] *= 1e-6
    x[2, 3] = 3000.0
    # rows whose block amax is exactly format_max * 2^k: the quotient is an exact power of two, where the fast-math division
    # decides the exponent
    for r, top in zip(range(3, 12), (448.0, 6.0, 28.0, 7.5, 127.0, 57344.0 / 64, 7.0, 3.5, 16.0)):
        for b in range(4):
            x[r, b * 32: (b + 1) * 32] = x[r, b * 32: (b + 1) * 32].clamp(-1, 1) * top * 2.0 ** (b - 2) * 0.99
            x[r, b * 32 + 5] = -top * 2.0 ** (b - 2)
    x[12, :8] = torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5.0, 6.0])
    x[12, 8:32] = x[12, 8:32].clamp(-5.9, 5.9)
    return x


ELEM = {"E4M3": (4, 3), "E5M2": (5, 2), "INT8": 8, "E0M3": (0, 3), "E1M2": (1, 2), "E3M0": (3, 0), "E2M1": (2, 1),
        "E3M2": (3, 2), "E2M3": (2, 3)}


def gen_mx(ext):
    g = {}
    base = mx_input()
    g["amax_sweep"] = to_np(torch.cat([torch.rand(4096) * 100, 2.0 ** torch.arange(-20, 20).float() * 448,
                                       2.0 ** torch.arange(-20, 20).float() * 6, 2.0 ** torch.arange(-20, 20).float() * 28]))
    for dn, dt in DTYPES.items():
        x = base.to(dt).cuda()
        g[f"x_{dn}"] = to_np(x)
        for fn in ELEM:
            for block in (32, 16, 8):
                if block != 32 and fn not in ("E4M3", "E2M1", "E3M0"):
                    continue
                y = ext.fused_amax_convert(x, block, getattr(ext.Types, fn), ext.Types.E8M0, None)
                g[f"y_{dn}_{fn}_{block}"] = to_np(y)
        # ragged last dim
        xr = x[:, :100].contiguous()
        g[f"yrag_{dn}"] = to_np(ext.fused_amax_convert(xr, 32, ext.Types.E4M3, ext.Types.E8M0, None))
        # low-precision scale formats, with and without a global amax (the non-Triton NVFP4 path and friends)
        ga = x.float().abs().amax().reshape(1)
        g[f"ga_{dn}"] = to_np(ga)
        for fn, sn in (("E2M1", "E4M3"), ("E4M3", "E5M2"), ("INT8", "E4M3")):
            for use_ga in (0, 1):
                y = ext.fused_amax_convert(x, 16, getattr(ext.Types, fn), getattr(ext.Types, sn), ga if use_ga else None)
                g[f"ys_{dn}_{fn}_{sn}_{use_ga}"] = to_np(y)
    # E8M0 scales alone: one-element-per-block trick (block of 8 with a single nonzero = amax) gives the unscale exponent
    a = torch.from_numpy(g["amax_sweep"]).cuda()
    probe = torch.zeros(a.numel(), 8, device="cuda")
    probe[:, 0] = a
    for fn in ("E4M3", "E2M1", "E3M2"):
        g[f"probe_{fn}"] = to_np(ext.fused_amax_convert(probe, 8, getattr(ext.Types, fn), ext.Types.E8M0, None)[:, 0])
    np.savez_compressed(os.path.join(OUT, "mx_b200.npz"), **g)
    print("mx_b200.npz:", len(g), "arrays")
    return g


def gen_qtensor(ext_main):
    import modelopt.torch.quantization.extensions as ex

    ex.get_cuda_ext.extension = ext_main  # the prebuilt extension instead of a JIT build on the box
    from modelopt.torch.quantization.qtensor import (FP8QTensor, INT4QTensor, INT8QTensor, MXFP4QTensor, MXFP8QTensor,
                                                     NF4QTensor)

    g = {}
    base = base_tensor()
    for dn, dt in DTYPES.items():
        x = base.to(dt).cuda()
        nz = x.clone()
        nz[1] = x[0]
        nz[2, :32] = x[0, :32]
        for block in (128, 32):
            for name, cls, extra in (("int4", INT4QTensor, ()), ("nf4", NF4QTensor, (nz.numel() // block,))):
                try:
                    qt, s = cls.quantize(nz, block, *extra)
                    g[f"{name}_q_{dn}_{block}"], g[f"{name}_s_{dn}_{block}"] = to_np(qt._quantized_data), to_np(s)
                    if name == "int4":
                        g[f"int4_dq_{dn}_{block}"] = to_np(qt.dequantize(scale=s, block_sizes={-1: block}))
                    elif dt != torch.float16:
                        sq, ds, zp = NF4QTensor.double_quantization(s.float(), nz.numel() // block, 8)
                        g[f"nf4_sq_{dn}_{block}"], g[f"nf4_ds_{dn}_{block}"], g[f"nf4_zp_{dn}_{block}"] = to_np(sq), to_np(ds), to_np(zp)
                        g[f"nf4_dq_{dn}_{block}"] = to_np(qt.dequantize(scale=sq, double_scale=ds, scale_zeros=zp,
                                                                        block_sizes={-1: block}))
                except Exception as e:  # AT_DISPATCH_FLOATING_TYPES has no half types
                    print(f"reference {name} {dn} block {block}: {type(e).__name__}: {str(e)[:100]}")
        for name, cls in (("int8", INT8QTensor), ("fp8", FP8QTensor)):
            for tag, kw in (("pt", {}), ("ax0", {"axis": 0}), ("b128", {"block_sizes": {-1: 128}}),
                            ("b2d", {"block_sizes": {-1: 64, -2: 8}})):
                qt, s = cls.quantize(nz, **kw)
                g[f"{name}_q_{dn}_{tag}"], g[f"{name}_s_{dn}_{tag}"] = to_np(qt._quantized_data.view(torch.uint8)), to_np(s)
                dkw = {"block_sizes": kw["block_sizes"]} if "block_sizes" in kw else {}
                g[f"{name}_dq_{dn}_{tag}"] = to_np(qt.dequantize(scale=s, **dkw))
        qt, s = MXFP4QTensor.quantize(x, 32)
        g[f"mxfp4_q_{dn}"], g[f"mxfp4_s_{dn}"] = to_np(qt._quantized_data), to_np(s)
        g[f"mxfp4_dq_{dn}"] = to_np(qt.dequantize(scale=s, block_sizes={-1: 32}))
        qt, s = MXFP8QTensor.quantize(x)
        g[f"mxfp8_q_{dn}"], g[f"mxfp8_s_{dn}"] = to_np(qt._quantized_data.view(torch.uint8)), to_np(s)
        g[f"mxfp8_dq_{dn}"] = to_np(qt.dequantize(scale=s))
    np.savez_compressed(os.path.join(OUT, "qtensor_b200.npz"), **g)
    print("qtensor_b200.npz:", len(g), "arrays")
    return g


def gen_hessian():
    """Reference local-Hessian pieces on the GPU: `_LocalHessianAccumulator.accumulate` (model_calib.py:848-862) and the
    Triton `nvfp4_fp8_scale_sweep_hessian` (gemm/nvfp4_fp8_sweep.py:236-290) + the 126-step Python sweep with the
    error_func (calib/mse.py:286-299)."""
    from modelopt.torch.kernels.quantization.gemm import nvfp4_fp8_scale_sweep_hessian
    from modelopt.torch.quantization.model_calib import _LocalHessianAccumulator

    g = {}
    torch.manual_seed(99)
    cout, cin = 192, 256
    for dn, dt in DTYPES.items():
        x1 = (torch.randn(300, cin) * (torch.rand(cin) * 2 + 0.1)).to(dt).cuda()
        x2 = (torch.randn(77, cin) * 0.5).to(dt).cuda()
        acc = _LocalHessianAccumulator(cout, cin, 16)
        acc.accumulate(x1)
        acc.accumulate(x2.reshape(7, 11, cin))
        h = acc.normalized_hessian()
        w = (torch.randn(cout, cin) * 0.05).to(dt).cuda()
        w[3] = 0
        w[5, :16] *= 30
        ga = w.float().abs().amax()
        best = nvfp4_fp8_scale_sweep_hessian(w.reshape(-1, 16), ga, h, block_size=16)
        g[f"x1_{dn}"], g[f"x2_{dn}"], g[f"w_{dn}"] = to_np(x1), to_np(x2), to_np(w)
        g[f"h_{dn}"], g[f"hraw_{dn}"], g[f"ga_{dn}"], g[f"best_{dn}"] = to_np(h), to_np(acc.hessian_per_block), to_np(ga), to_np(best)
    np.savez_compressed(os.path.join(OUT, "hessian_b200.npz"), **g)
    print("hessian_b200.npz:", len(g), "arrays")


def gen_gptq():
    """Reference fused GPTQ block kernel (Triton, gemm/gptq_fused_kernel.py) and the whole fused blockwise update
    (utils/calib_utils.py:291-327) on the GPU."""
    from modelopt.torch.kernels.quantization.gemm.gptq_fused_kernel import gptq_fused_block_scalar
    from modelopt.torch.quantization.utils.calib_utils import (compute_hessian_inverse,
                                                               gptq_blockwise_update_fused_scalar, update_hessian)

    g = {}
    torch.manual_seed(321)
    rows, cin, blk = 96, 256, 128
    w = (torch.randn(rows, cin) * 0.04).cuda()
    w[:, 7] = 0  # dead input column
    x = (torch.randn(700, cin) * (torch.rand(cin) * 2 + 0.2)).cuda()
    hess, n = update_hessian(x[:400], torch.zeros(cin, cin, device="cuda"), 0)
    hess, n = update_hessian(x[400:], hess, n)
    h_inv = compute_hessian_inverse(hess, w, 0.01)
    amax = w.view(rows, cin // 16, 16).abs().amax(-1)
    gs = float(amax.max()) / (6.0 * 448.0)
    qw, err = gptq_fused_block_scalar(w[:, :blk].clone().contiguous(), amax, gs, h_inv[:blk, :blk].contiguous(), 16, 0)
    full = w.clone()
    gptq_blockwise_update_fused_scalar(full, amax, gs, h_inv, blk, 16)
    for k, v in (("w", w), ("x", x), ("hess", hess), ("h_inv", h_inv), ("amax", amax), ("qw", qw), ("err", err), ("full", full)):
        g[k] = to_np(v)
    g["gs"] = np.asarray(gs, dtype=np.float64)
    g["n"] = np.asarray(n)
    np.savez_compressed(os.path.join(OUT, "gptq_b200.npz"), **g)
    print("gptq_b200.npz:", len(g), "arrays")


def compare_product(gm, gq):
    """For information: how the product kernels compare on the captured inputs (the tests assert this)."""
    from modelopt_b200 import ops

    bad = 0
    for dn, dt in DTYPES.items():
        x = torch.from_numpy(gm[f"x_{dn}"])
        x = (x.view(dt) if dt != torch.float32 else x).cuda()
        for fn, bits in ELEM.items():
            key = f"y_{dn}_{fn}_32"
            want = torch.from_numpy(gm[key])
            want = (want.view(dt) if dt != torch.float32 else want).cuda()
            got = ops.qdq_mx(x, 32, bits, (8, 0), numerics="gpu")
            n = int((got.float() != want.float()).sum())
            bad += n
            if n:
                print(f"  mx {key}: {n} mismatches")
    print("product vs reference MX ext:", "all equal" if bad == 0 else f"{bad} mismatching elements")


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if "--hessian-only" in sys.argv:
        gen_hessian()
        sys.exit(0)
    if "--gptq-only" in sys.argv:
        gen_gptq()
        sys.exit(0)
    gm = gen_mx(load_ext("ext_mx", "modelopt_cuda_ext_mx"))
    gq = gen_qtensor(load_ext("ext_main", "modelopt_cuda_ext"))
    gen_hessian()
    gen_gptq()
    try:
        compare_product(gm, gq)
    except Exception as e:  # noqa: BLE001
        print("product comparison failed:", type(e).__name__, e)
