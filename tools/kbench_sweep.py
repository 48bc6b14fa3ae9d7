"""Micro-benchmark of the scale searches on the reference's own benchmark shape
(8192 x 4096 weight, tests/gpu/torch/quantization/test_nvfp4_fp8_sweep_kernel.py): the fused FP8-scale sweep and the
fused MSE multiplier sweep, with the reference's Triton / Python paths beside them when oracle/_ref is present."""

from __future__ import annotations

import os
import sys
import warnings

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
warnings.filterwarnings("ignore")
from modelopt_b200 import ops  # noqa: E402
from modelopt_b200 import tensor_quant as tq  # noqa: E402
from modelopt_b200.calib import MseCalibrator  # noqa: E402
from tools.kbench import timeit  # noqa: E402


def line(name, fn, iters=10, warm=3):
    med, best = timeit(fn, iters, warm)
    print(f"{name:58s} {med:9.3f} ms   best {best:9.3f} ms", flush=True)
    return med


def main():
    torch.manual_seed(0)
    for dt in (torch.float32, torch.bfloat16):
        w = torch.randn(8192, 4096, device="cuda", dtype=dt)
        xb = w.view(-1, 16)
        ga = w.float().abs().max()
        print(f"--- weight 8192 x 4096 {dt}")
        line("nvfp4 fp8-scale sweep (mo_nvfp4_fp8_scale_sweep)", lambda: ops.nvfp4_fp8_scale_sweep(xb, ga, 16))
        am_row = ops.reduce_amax(w, axis=1).float()
        for nm, kw, qf in (("int8", {"num_bits": 8, "unsigned": False, "narrow_range": True},
                            lambda x, a: tq.fake_tensor_quant(x, a, None, 8, False, True)),
                           ("fp8", {"num_bits": (4, 3)}, lambda x, a: tq.scaled_e4m3(x, a, None, 4, 3))):
            def fused():
                c = MseCalibrator(amax=am_row, axis=0, quant_func=qf, fused=kw)
                c.collect(w)
                return c.compute_amax()

            def loop():
                c = MseCalibrator(amax=am_row, axis=0, quant_func=qf)
                c.collect(w)
                return c.compute_amax()

            line(f"MseCalibrator {nm} per-row, fused (mo_mse_sweep)", fused)
            line(f"MseCalibrator {nm} per-row, 39 x (qdq + mse + sum) loop", loop, 3, 1)
        cin = w.shape[1]
        act = torch.randn(16384, cin, device="cuda", dtype=w.dtype)
        hacc = torch.zeros(cin // 16, 16, 16, device="cuda")
        line("block Hessian accumulate [16384, cin] (mo_block_hessian)", lambda: ops.block_hessian_into(act, hacc))
        hn = hacc / 16384
        line("Hessian-weighted fp8-scale sweep (mo_nvfp4_fp8_scale_sweep_hessian)",
             lambda: ops.nvfp4_fp8_scale_sweep_hessian(xb, ga, hn))
        del act
        ref = os.path.join(ROOT, "oracle", "_ref")
        if os.path.isdir(os.path.join(ref, "modelopt")) and "--no-ref" not in sys.argv:
            sys.path[:0] = [ref, os.path.join(ROOT, "oracle", "ref_shims")]
            try:
                import modelopt.torch.kernels.quantization.gemm as tk
                from modelopt.torch.quantization.calib.mse import NVFP4MSECalibrator
                from modelopt.torch.quantization.tensor_quant import static_blockwise_fp4_fake_quant

                line("REF triton nvfp4_fp8_scale_sweep", lambda: tk.nvfp4_fp8_scale_sweep(xb, ga, 16))
                bam = xb.float().abs().amax(dim=-1)
                os.environ["MODELOPT_NVFP4_TRITON_SWEEP"] = "0"

                def ref_py():
                    c = NVFP4MSECalibrator(amax=bam, axis=0, global_amax=ga,
                                           quant_func=lambda t, a: static_blockwise_fp4_fake_quant(t, a, ga))
                    c.collect(xb)
                    return c.compute_amax()

                line("REF python 126-step sweep", ref_py, 2, 1)
                os.environ.pop("MODELOPT_NVFP4_TRITON_SWEEP", None)
                # local-Hessian pieces: H = sum X^T X of a [16384, cin] activation, Hessian-weighted sweep
                from modelopt.torch.quantization.model_calib import _LocalHessianAccumulator as RefAcc

                cin = w.shape[1]
                act = torch.randn(16384, cin, device="cuda", dtype=w.dtype)
                racc = RefAcc(w.shape[0], cin, 16)
                line("REF _LocalHessianAccumulator.accumulate [16384, cin]", lambda: racc.accumulate(act), 5, 2)
                hh = racc.normalized_hessian()
                line("REF triton nvfp4_fp8_scale_sweep_hessian", lambda: tk.nvfp4_fp8_scale_sweep_hessian(xb, ga, hh, 16))
            except Exception as e:  # noqa: BLE001
                print("REF failed:", repr(e)[:300])
        del w, xb
        torch.cuda.empty_cache()


def with_clocks(fn):
    """Run fn() with the SM clock and throttle reasons sampled (bench.py's sampler) and print them: a kernel figure without
    the clocks it was taken at is a claim, not a measurement."""
    sys.path[:0] = [ROOT]
    from bench import Clocks

    with Clocks(torch.cuda.current_device()) as clk:
        fn()
    print("# clocks during this run:", clk.summary(), flush=True)


if __name__ == "__main__":
    with_clocks(main)
