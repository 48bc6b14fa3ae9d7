"""GPU time of the three streaming kernels of the bench step (amax collect, FP8 QDQ, NVFP4 QDQ) on the Llama-3-8B activation
shapes, measured as CUDA-graph replays of 4 x 8 back-to-back launches over 4 rotating buffers (no host gaps, L2 defeated)
with the SM clock sampled during the run.  Developer tool:  [MODELOPT_B200_LIB=<variant .so>] python tools/stream_bench.py"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from bench import Clocks  # noqa: E402
from modelopt_b200 import ops  # noqa: E402

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]


def graph_time(fns, reps=10):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for f in fns:
            f()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(8):
                for f in fns:
                    f()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(reps):
            g.replay()
        e1.record(s)
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * 8 * len(fns))


def main():
    print("library:", os.environ.get("MODELOPT_B200_LIB", "default"))
    with Clocks(0) as clk:
        for shape in [(16384, 1024), (16384, 4096), (16384, 14336), (4096, 4096), (14336, 4096)]:
            xs = [torch.randn(shape, device="cuda", dtype=torch.bfloat16) for _ in range(4)]
            ys = [torch.empty_like(x) for x in xs]
            n = xs[0].numel()
            am = ops.reduce_amax(xs[0]).float()
            bits = torch.zeros(8, dtype=torch.int32, device="cuda")
            rows = {
                "amax collect": (2 * n, [lambda x=x: ops.amax_into(x, 1, 1, n, bits[:1]) for x in xs]),
                "fp8 qdq": (4 * n, [lambda x=x, y=y: ops.qdq_fp8(x, am, out=y) for x, y in zip(xs, ys)]),
                "int8 qdq": (4 * n, [lambda x=x, y=y: ops.qdq_int(x, am, 8, False, False, out=y) for x, y in zip(xs, ys)]),
                "nvfp4 qdq": (4 * n, [lambda x=x, y=y: ops.qdq_nvfp4_dynamic(x, am, 16, None, out=y) for x, y in zip(xs, ys)]),
                "copy": (4 * n, [lambda x=x, y=y: y.copy_(x) for x, y in zip(xs, ys)]),
            }
            line = f"{str(shape):16s}"
            for name, (nbytes, fns) in rows.items():
                ms = graph_time(fns)
                line += f"  {name} {ms * 1e3:6.1f} us {nbytes / ms / 1e6 / PEAK * 100:5.1f}%"
            print(line, flush=True)
    print("clocks:", clk.summary())


if __name__ == "__main__":
    main()
