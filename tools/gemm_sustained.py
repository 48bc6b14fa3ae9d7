"""bf16 linear GEMMs of a Llama forward under SUSTAINED load (each candidate runs back to back for ~0.6 s, so the GPU sits
at its power-capped clock like inside a model forward): cuBLAS (F.linear) vs mo_linear_pair vs mo_linear, TFLOP/s and the
SM clock sampled during the run.  Developer tool:  python tools/gemm_sustained.py [--model 8b|70b] [--secs 0.6]"""
import argparse
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
from bench import SHAPES, Clocks  # noqa: E402
from modelopt_b200 import ops  # noqa: E402


def sustained(fn, secs):
    fn()
    torch.cuda.synchronize()
    n, t_end = 0, time.perf_counter() + 0.15 * secs
    while time.perf_counter() < t_end:  # ramp: reach the sustained clock before timing
        for _ in range(20):
            fn()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with Clocks(0) as clk:
        e0.record()
        t_end = time.perf_counter() + secs
        while time.perf_counter() < t_end:
            for _ in range(20):
                fn()
            n += 20
            torch.cuda.synchronize()
        e1.record()
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, clk.summary()["sm_mhz"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="8b")
    ap.add_argument("--secs", type=float, default=0.6)
    ap.add_argument("--tokens", type=int, default=16384)
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    M = a.tokens
    shapes = sorted({(co, ci) for _, co, ci in SHAPES[a.model]["linears"]}) + [(128256, SHAPES[a.model]["linears"][0][2])]
    print(f"M = {M}; ms per GEMM and TFLOP/s under sustained load (SM MHz during the run)")
    tot = {"cublas": 0.0, "pair": 0.0}
    for N, K in shapes:
        x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
        w = (torch.randn(N, K, device="cuda") * 0.02).to(torch.bfloat16)
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        flops = 2.0 * M * N * K
        cands = {"cublas": lambda: torch.nn.functional.linear(x, w), "pair": lambda: ops.linear(x, w, out=out, pair=True),
                 "single": lambda: ops.linear(x, w, out=out, pair=False)}
        line = f"N={N:6d} K={K:6d}:"
        for name, fn in cands.items():
            if a.only and name not in a.only.split(","):
                continue
            ms, mhz = sustained(fn, a.secs)
            if name in tot:
                tot[name] += ms
            line += f"  {name} {ms:7.3f} ms {flops / ms / 1e9:7.1f} TF ({mhz} MHz)"
        print(line, flush=True)
    print(f"sum over shapes: cuBLAS {tot['cublas']:.3f} ms, pair {tot['pair']:.3f} ms")


if __name__ == "__main__":
    main()
