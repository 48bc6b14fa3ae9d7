"""Kernel micro-benchmark: GB/s of each product kernel on BASELINE shapes (CUDA events, L2-exceeding
inputs), and — when oracle/_ref holds the reference — the reference's own GPU kernels beside them."""

from __future__ import annotations

import json
import os
import sys
import warnings

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
warnings.filterwarnings("ignore")
from modelopt_b200 import ops  # noqa: E402

PEAK = 6544.3
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:  # noqa: BLE001
    pass


def timeit(fn, iters=20, warm=5):
    """(median, best) ms per call.  The calls are captured into a CUDA graph (10 per replay) so that the figure is the
    kernel's GPU time, not the ~15-30 us of Python + ctypes per launch that dominated the small-tensor rows of round 1
    (amax on a 134 MB tensor: 21 us of GPU time inside a 54 us launch-to-launch period); calls that cannot be captured
    (host syncs inside) fall back to one event pair per eager call."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    try:
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                for _ in range(10):
                    fn()
            g.replay()
            torch.cuda.synchronize()
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
            for a, b in ev:
                a.record(s); g.replay(); b.record(s)
            torch.cuda.synchronize()
        ts = sorted(a.elapsed_time(b) / 10 for a, b in ev)
        return ts[len(ts) // 2], ts[0]
    except Exception:  # noqa: BLE001  (not capturable)
        torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in ev)
    return ts[len(ts) // 2], ts[0]


def report(name, nbytes, fn):
    med, best = timeit(fn)
    print(f"{name:46s} {med:8.3f} ms  {nbytes / med / 1e6:8.1f} GB/s  ({nbytes / med / 1e6 / PEAK * 100:5.1f}% of {PEAK:.0f})"
          f"   best {nbytes / best / 1e6:8.1f}", flush=True)


def main():
    no_ref = "--no-ref" in sys.argv
    argv = [a for a in sys.argv[1:] if not a.startswith("--")]
    shapes = [(16384, 4096), (16384, 14336)] if not argv else [tuple(map(int, argv[0].split("x")))]
    for shape in shapes:
        x = torch.randn(shape, device="cuda", dtype=torch.bfloat16)
        y = torch.empty_like(x)
        n = x.numel()
        am = ops.reduce_amax(x).float()
        print(f"--- bf16 {shape}  ({n * 2 / 1e6:.0f} MB)")
        report("copy (torch y.copy_(x))", 4 * n, lambda: y.copy_(x))
        report("amax per-tensor", 2 * n, lambda: ops.reduce_amax(x))
        report("amax per-row (axis 0 kept)", 2 * n, lambda: ops.reduce_amax(x, axis=1))
        report("amax per-col (last axis kept)", 2 * n, lambda: ops.reduce_amax(x, axis=0))
        report("amax block-128", 2 * n, lambda: ops.reduce_amax(x.view(-1, 128), axis=1))
        report("qdq fp8 per-tensor", 4 * n, lambda: ops.qdq_fp8(x, am, out=y))
        report("qdq int8 per-tensor", 4 * n, lambda: ops.qdq_int(x, am, 8, False, False, out=y))
        amc = ops.reduce_amax(x, axis=1).float()
        report("qdq int8 per-channel", 4 * n, lambda: ops.qdq_int(x, amc, 8, False, False, out=y))
        xb = x.view(-1, 128)
        amb = ops.reduce_amax(xb, axis=1).float()
        report("qdq int4 block-128", 4 * n, lambda: ops.qdq_int(xb, amb, 4, False, False, out=y.view(-1, 128)))
        report("qdq nvfp4 dynamic (triton numerics)", 4 * n, lambda: ops.qdq_nvfp4_dynamic(x, am, 16, "triton", out=y))
        report("qdq nvfp4 dynamic (ieee numerics)", 4 * n, lambda: ops.qdq_nvfp4_dynamic(x, am, 16, "ieee", out=y))
        bam = ops.reduce_amax(x.view(-1, 16), axis=1).float()
        sc = ops.fp4_scales(bam, am, True)
        report("qdq nvfp4 static (scales given)", 4.25 * n, lambda: ops.qdq_nvfp4_static(x.view(-1, 16), bam, am))
        from modelopt_b200.calib import HistogramCalibrator
        hc = HistogramCalibrator(8, None, False)
        hc.collect(x)
        report("hist collect (steady state, fused amax)", 2 * n, lambda: hc.collect(x))
        rm = am.reshape(1).float().contiguous()
        cnt = torch.zeros(2048, dtype=torch.int64, device="cuda")
        report("hist kernel only (mo_hist, 2048 bins)", 2 * n, lambda: ops.hist_into(x, rm, cnt))
        cnt5 = torch.zeros(5000, dtype=torch.int64, device="cuda")
        report("hist kernel only (mo_hist, 5000 bins)", 2 * n, lambda: ops.hist_into(x, rm, cnt5))
        xf = x.float()
        report("REF-style torch: abs+float+max+histc", 2 * n, lambda: torch.histc(x.abs().float(), bins=2048, min=0, max=float(x.abs().float().max())))
        del xf
        import time as _t
        for m, kw in (("percentile", {}), ("mse", {}), ("entropy", {})):
            best = None  # the first call pays the lazy load of the kernel's module: report the best of three
            for _ in range(3):
                torch.cuda.synchronize(); t0 = _t.perf_counter(); v = hc.compute_amax(m, **kw); torch.cuda.synchronize()
                dt = (_t.perf_counter() - t0) * 1e3
                first = dt if best is None else first
                best = dt if best is None else min(best, dt)
            print(f"hist search {m:10s} ({hc._counts.numel()} bins): {best:8.2f} ms (first call {first:.2f}) -> {float(v):.4f}", flush=True)
        from modelopt_b200.qtensor import NVFP4QTensor
        sf2 = am / 2688.0
        report("nvfp4 pack (codes + e4m3 scales)", 2.5625 * n, lambda: ops.nvfp4_pack(x, 16, sf2))
        cq, sfq = ops.nvfp4_pack(x, 16, sf2)
        report("nvfp4 unpack", 2.5625 * n, lambda: ops.nvfp4_unpack(cq, sfq, sf2, 16, torch.bfloat16))
        del cq, sfq
        sc = (torch.rand(shape[1], device="cuda") + 0.5)
        scb = sc.to(torch.bfloat16)
        y = torch.empty_like(x)
        report("scale_cols (x * s[col])", 4 * n, lambda: ops.scale_cols(x, sc, out=y))
        bits = torch.zeros(shape[0], dtype=torch.int32, device="cuda")
        report("scale_cols + per-row amax (smooth fold)", 4 * n, lambda: ops.scale_cols(x, sc, out=y, amax_bits=bits, amax_mode="row"))
        report("int4 block-128 dynamic amax + QDQ (one pass)", 4 * n, lambda: ops.qdq_int_block_dynamic(x, 128, 4, out=y))
        report("  same with pre_quant_scale (AWQ search)", 4 * n, lambda: ops.qdq_int_block_dynamic(x, 128, 4, prescale=scb, out=y))
        report("int8 per-tensor QDQ with pre_quant_scale", 4 * n, lambda: ops.qdq_int(x, am, 8, False, False, out=y, prescale=scb))
        report("fp8 per-tensor QDQ with pre_quant_scale", 4 * n, lambda: ops.qdq_fp8(x, am, out=y, prescale=scb))
        report("awq weight_scale (block 128)", 2 * n, lambda: ops.awq_weight_scale(x, 128))
        report("awq act_scale (|x| column mean)", 2 * n, lambda: ops.awq_act_scale(x))
        acc = torch.zeros(1, dtype=torch.float64, device="cuda")
        report("sq_diff_mean (loss)", 4 * n, lambda: ops.sq_diff_mean_into(x, y, acc))
        report("REF-style 3-pass: x*s, block amax, qdq", 4 * n, lambda: ops.qdq_int((x * scb).view(-1, 128), ops.reduce_amax((x * scb).view(-1, 128), axis=1).float(), 4, False, False))
        del y
        ref = os.path.join(ROOT, "oracle", "_ref")
        if os.path.isdir(os.path.join(ref, "modelopt")) and not no_ref:
            sys.path[:0] = [ref, os.path.join(ROOT, "oracle", "ref_shims")]
            try:
                import modelopt.torch.kernels.quantization.gemm as tk
                from modelopt.torch.quantization.calib import MaxCalibrator
                from modelopt.torch.quantization.tensor_quant import fake_tensor_quant, scaled_e4m3
                from modelopt.torch.quantization.utils import reduce_amax as ref_amax

                report("REF triton fp4_fake_quant_block", 4 * n, lambda: tk.fp4_fake_quant_block(x, am))
                report("REF reduce_amax", 2 * n, lambda: ref_amax(x))
                report("REF MaxCalibrator.collect", 2 * n, lambda: MaxCalibrator(8, None, False).collect(x))
                try:
                    report("REF scaled_e4m3 (JIT ext)", 4 * n, lambda: scaled_e4m3(x, am, None, 4, 3))
                    report("REF fake_tensor_quant int8 (JIT ext)", 4 * n, lambda: fake_tensor_quant(x, am, None, 8, False, False))
                    report("REF fake_tensor_quant int8 per-channel", 4 * n,
                           lambda: fake_tensor_quant(x, amc.unsqueeze(1), None, 8, False, False))
                except Exception as e:  # noqa: BLE001
                    print("REF JIT ext failed:", repr(e)[:200])
            except Exception as e:  # noqa: BLE001
                print("REF import failed:", repr(e)[:300])
        del x
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
