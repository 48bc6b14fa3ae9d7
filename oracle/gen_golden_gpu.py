"""Capture outputs of the REAL reference's GPU kernels on a B200.  TEST INFRASTRUCTURE ONLY.

The reference's NVFP4 fake-quant has no CPU path and divides with div.full.f32 (approximate), so the
only way to pin bit parity with it is to run its compiled Triton kernels on the target GPU once and
commit the input/output vectors.  Recipe (build container -> GPU box -> back):

    mkdir -p oracle/_ref && cp -r /root/reference/modelopt /root/reference/modelopt_recipes oracle/_ref/
    gpurun -- 'python oracle/gen_golden_gpu.py'          # oracle/_ref is git-ignored, but travels
    cp gpurun_out/nvfp4_triton_b200.npz tests/golden/

Writes gpurun_out/nvfp4_triton_b200.npz (inputs + reference outputs) and prints, for information,
how the product kernels compare on the same inputs.
"""

from __future__ import annotations

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


def adversarial_rows():
    grid = torch.tensor([0, 0.5, 1, 1.5, 2, 3, 4, 6.0])
    ties = torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5, 6.0])
    rows = [torch.cat([grid, -grid]), torch.cat([ties, -ties]), torch.cat([ties - 0.1, -(ties + 0.1)]),
            torch.zeros(16), torch.full((16,), 1e-12), torch.cat([torch.tensor([1000.0]), torch.full((15,), 1e-3)]),
            torch.tensor([-0.0] * 8 + [6.0 * 2**-11] * 8), torch.tensor([-1e-4] * 16)]
    return torch.stack(rows)


def gen_sweep():
    """Reference NVFP4MSECalibrator on the GPU: Triton fast path and the 126-step Python sweep (calib/mse.py:175-311)."""
    from modelopt.torch.quantization.calib.mse import NVFP4MSECalibrator
    from modelopt.torch.quantization.tensor_quant import static_blockwise_fp4_fake_quant

    g = {}
    torch.manual_seed(2024)
    base = torch.randn(6000, 16) * torch.rand(6000, 1) * 3
    base[1] = 0
    base[2] *= 1e-4
    base[3, 0] = 30.0
    for dn, dt in DTYPES.items():
        x = base.to(dt).cuda()
        bam = x.float().abs().amax(dim=-1)
        ga = bam.max()
        g[f"x_{dn}"], g[f"ga_{dn}"] = to_np(x), to_np(ga)
        for key, flag in (("best", "1"), ("best_py", "0")):
            os.environ["MODELOPT_NVFP4_TRITON_SWEEP"] = flag
            cal = NVFP4MSECalibrator(amax=bam, axis=0, global_amax=ga,
                                     quant_func=lambda t, a: static_blockwise_fp4_fake_quant(t, a, ga))
            cal.collect(x)
            g[f"{key}_{dn}"] = to_np(cal.compute_amax())
        print(f"sweep {dn}: triton vs python differing blocks =",
              int((g[f'best_{dn}'] != g[f'best_py_{dn}']).sum()), "/", x.shape[0])
    os.environ.pop("MODELOPT_NVFP4_TRITON_SWEEP", None)
    out = os.path.join(ROOT, "gpurun_out", "nvfp4_sweep_b200.npz")
    np.savez_compressed(out, **g)
    print("wrote", out, len(g), "arrays")


def main():
    assert torch.cuda.is_available()
    import modelopt.torch.kernels.quantization.gemm as tk
    from modelopt.torch.quantization.tensor_quant import dynamic_block_quant, scaled_e4m3, fake_tensor_quant

    print("reference triton available:", tk.IS_AVAILABLE, "| device:", torch.cuda.get_device_name(0))
    g = {}
    torch.manual_seed(4242)
    base = torch.randn(64, 1024) * 2.0
    base[0, :16] = 0
    base[1] *= 1e-3
    base[2] *= 50
    adv = adversarial_rows()
    for dn, dt in DTYPES.items():
        for nm, src in (("rand", base), ("adv", adv)):
            x = src.to(dt).cuda()
            g[f"x_{nm}_{dn}"] = to_np(x)
            for an, fac in (("max", 1.0), ("clip", 0.7)):
                ga = (x.float().abs().max() * fac).float()
                y = tk.fp4_fake_quant_block(x, ga)
                g[f"ga_{nm}_{an}_{dn}"], g[f"y_{nm}_{an}_{dn}"] = to_np(ga), to_np(y)
                # the TensorQuantizer route: dynamic_block_quant(inputs, block, amax, bias, num_bits, scale_bits)
                y2 = dynamic_block_quant(x, 16, ga, None, (2, 1), (4, 3))
                assert torch.equal(y2.view(torch.int16 if dt != torch.float32 else torch.int32),
                                   y.view(torch.int16 if dt != torch.float32 else torch.int32))
        # ragged last dim
        xr = base[:, :1000].to(dt).cuda().contiguous()
        ga = xr.float().abs().max()
        g[f"x_rag_{dn}"], g[f"ga_rag_{dn}"] = to_np(xr), to_np(ga)
        g[f"y_rag_{dn}"] = to_np(tk.fp4_fake_quant_block(xr, ga))
        # static kernel
        xb = base.to(dt).cuda().reshape(-1, 16)
        bam = xb.float().abs().amax(dim=1)
        gam = bam.max()
        for q in (True, False):
            y = tk.static_blockwise_fp4_fake_quant(xb, bam, gam, q)
            g[f"ystatic_q{int(q)}_{dn}"] = to_np(y)
            g[f"sstatic_q{int(q)}_{dn}"] = to_np(tk.compute_fp4_scales(bam, gam, q))
        g[f"bam_{dn}"], g[f"gam_{dn}"] = to_np(bam), to_np(gam)
        # FP8 / INT8 through the reference's JIT CUDA extensions (if they compile on the box)
        try:
            x = base.to(dt).cuda()
            am = x.float().abs().max() * 0.8
            g[f"fp8_amax_{dn}"], g[f"fp8_y_{dn}"] = to_np(am), to_np(scaled_e4m3(x, am.to(dt), None, 4, 3))
            g[f"int8_y_{dn}"] = to_np(fake_tensor_quant(x, am.to(dt), None, 8, False, False))
            amc = x.float().abs().amax(dim=1, keepdim=True)
            g[f"int8c_amax_{dn}"], g[f"int8c_y_{dn}"] = to_np(amc), to_np(fake_tensor_quant(x, amc.to(dt), None, 8, False, False))
        except Exception as e:  # noqa: BLE001
            print("reference CUDA ext unavailable:", repr(e)[:200])
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    gen_sweep()
    out = os.path.join(ROOT, "gpurun_out", "nvfp4_triton_b200.npz")
    np.savez_compressed(out, **g)
    print("wrote", out, len(g), "arrays")

    # informational: compare the product kernels right away
    try:
        from modelopt_b200 import ops
        sys.path.insert(0, os.path.join(ROOT, "tests")); from conftest import from_np

        for k in sorted(g):
            if not k.startswith("y_"):
                continue
            _, nm, an, dn = k.split("_") if k.count("_") == 3 else (None, "rag", None, k.split("_")[-1])
            dt = DTYPES[dn]
            x = from_np(g[f"x_{nm}_{dn}"], dt).cuda()
            ga = torch.from_numpy(g[f"ga_{nm}_{an}_{dn}" if an else f"ga_rag_{dn}"]).cuda()
            want = from_np(g[k], dt)
            for mode in ("triton", "ieee"):
                got = ops.qdq_nvfp4_dynamic(x, ga, 16, numerics=mode).cpu()
                bad = int((got.view(torch.int16 if dt != torch.float32 else torch.int32)
                           != want.view(torch.int16 if dt != torch.float32 else torch.int32)).sum())
                print(f"{k:24s} mode={mode:6s} bit-mismatches={bad}/{want.numel()}")
    except Exception as e:  # noqa: BLE001
        print("product comparison skipped:", repr(e)[:300])


if __name__ == "__main__":
    main()
