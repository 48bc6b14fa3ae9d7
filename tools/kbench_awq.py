"""BASELINE.json configs[3] in miniature: INT4_AWQ_CFG (block-128, awq_lite) on ONE Llama-3-8B-shaped linear,
mtq.quantize end to end on the GPU, ours vs the reference (oracle/_ref) on identical weights and batches.
Prints wall time (synchronised), the chosen pre_quant_scale agreement and the number of kernels we launched."""

from __future__ import annotations

import copy
import os
import sys
import time
import warnings

import torch
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200")]
warnings.filterwarnings("ignore")


def run(mtq, model, batches, cfg):
    def loop(m):
        for b in batches:
            m(b)

    torch.cuda.synchronize()
    t0 = time.perf_counter()
    with torch.no_grad():
        mtq.quantize(model, cfg, loop)
    torch.cuda.synchronize()
    return time.perf_counter() - t0


def main():
    cout, cin = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 and "x" in sys.argv[1] else "4096x4096").split("x"))
    torch.manual_seed(0)
    base = nn.Sequential(nn.Linear(cin, cout, bias=False)).to("cuda", torch.bfloat16)
    with torch.no_grad():
        base[0].weight.mul_(0.5)
        base[0].weight[:, ::97] *= 6.0        # a few salient input channels, what AWQ is for
    batches = [torch.randn(8, 2048, cin, device="cuda", dtype=torch.bfloat16) * (1 + torch.rand(cin, device="cuda") * 3).to(torch.bfloat16)
               for _ in range(2)]
    import modelopt_b200 as ours
    from modelopt_b200 import ops

    algo = next((a.split("=", 1)[1] for a in sys.argv if a.startswith("--algo=")), "awq_lite")  # awq_clip, awq_full

    def cfg_of(mod):
        c = copy.deepcopy(mod.INT4_AWQ_CFG)
        c["algorithm"] = {"method": algo} if algo == "awq_clip" else {"method": algo, "alpha_step": 0.1}
        return c

    def pqs(mod):
        q = mod[0].input_quantizer
        return q._pre_quant_scale.float().clone() if hasattr(q, "_pre_quant_scale") else torch.ones(1, device="cuda")

    print(f"algorithm {algo}, linear {cout}x{cin}, 2 batches of 16384 tokens", flush=True)
    res = {}
    for rep in range(2):  # second repetition is the warm one
        m = copy.deepcopy(base)
        n0 = ops.LAUNCHES[0]
        t = run(ours, m, batches, cfg_of(ours))
        res["ours"] = (t, pqs(m), m[0].weight_quantizer._amax.float().clone(),
                       ops.LAUNCHES[0] - n0, m[0].weight_quantizer._amax.dtype)
    print(f"ours      : {res['ours'][0] * 1e3:9.1f} ms   ({res['ours'][3]} kernel launches of this library)", flush=True)
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if os.path.isdir(os.path.join(ref_dir, "modelopt")) and "--no-ref" not in sys.argv:
        sys.path[:0] = [ref_dir, os.path.join(ROOT, "oracle", "ref_shims")]
        import modelopt.torch.quantization as mtq

        for rep in range(2):
            m = copy.deepcopy(base)
            t = run(mtq, m, batches, cfg_of(mtq))
            res["ref"] = (t, pqs(m), m[0].weight_quantizer._amax.float().clone(), 0, m[0].weight_quantizer._amax.dtype)
        print(f"reference : {res['ref'][0] * 1e3:9.1f} ms   -> speed-up {res['ref'][0] / res['ours'][0]:.1f}x", flush=True)
        a, b = res["ours"][1], res["ref"][1]
        print(f"pre_quant_scale: max rel diff {((a - b).abs() / b.abs()).max().item():.3e}, identical {(a == b).float().mean().item() * 100:.1f}%")
        if algo == "awq_clip":
            # which choice is better?  exact (fp64) clip-search objective of both amax sets on the sampled tokens
            w = base[0].weight
            tok = torch.cat([b.reshape(-1, cin)[0::max(1, b.numel() // cin // 64)] for b in batches]).double()
            xb = tok.reshape(tok.shape[0], -1, 128)
            for who in ("ours", "ref"):
                am = res[who][2].reshape(-1, 1).contiguous()
                wq = ops.qdq_int(w.reshape(-1, 128).contiguous(), am, num_bits=4, narrow_range=False).reshape(cout, -1, 128)
                d = (wq.double() - w.reshape(cout, -1, 128).double())
                tot = 0.0
                for c0 in range(0, cout, 512):
                    tot += torch.einsum("tnb,cnb->ctn", xb, d[c0:c0 + 512]).pow(2).sum().item()
                print(f"exact clip objective with {who:4s} amax: {tot / tok.shape[0]:.6e}")
            # lower bound: the per-block optimum of the exact objective over the 11 ratios (amax products in fp32)
            w_amax = w.reshape(-1, 128).float().abs().amax(1, keepdim=True)
            ratios = [round(0.5 + 0.05 * i, 2) for i in range(10)] + [1.0]
            per = []
            for r in ratios:
                am = (w_amax * r).contiguous()
                wq = ops.qdq_int(w.reshape(-1, 128).contiguous(), am, num_bits=4, narrow_range=False).reshape(cout, -1, 128)
                d = (wq.double() - w.reshape(cout, -1, 128).double())
                per.append(torch.cat([torch.einsum("tnb,cnb->ctn", xb, d[c0:c0 + 512]).pow(2).sum(1)
                                      for c0 in range(0, cout, 512)]))
            per = torch.stack(per)                                   # [11, co, nb]
            best, k_best = per.min(0)
            print(f"exact clip objective, per-block optimum : {best.sum().item() / tok.shape[0]:.6e}")
            for who in ("ours", "ref"):
                am = res[who][2].reshape(cout, -1)
                k = ((am / w_amax.reshape(cout, -1)).unsqueeze(0) - torch.tensor(ratios, device="cuda").view(-1, 1, 1)).abs().argmin(0)
                print(f"  {who:4s}: amax dtype {res[who][4]}, ratio index == exact optimum in {(k == k_best).float().mean().item() * 100:.1f}% of blocks,"
                      f" histogram {torch.bincount(k.reshape(-1), minlength=11).tolist()}")
        a, b = res["ours"][2], res["ref"][2]
        print(f"weight block amax: max rel diff {((a - b).abs() / b.abs().clamp_min(1e-30)).max().item():.3e}, identical {(a == b).float().mean().item() * 100:.1f}%")


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
