"""Kernel-time breakdown (torch.profiler / CUPTI) of one Llama micro-batch forward: plain, calibrating, fake-quant.
Developer tool: python tools/model_profile.py [--layers 8] [--preset nvfp4]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200"), os.path.join(ROOT, "tools")]
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import model_level as ml  # noqa: E402


def prof(tag, fn, rows=22):
    fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as p:
        fn()
        torch.cuda.synchronize()
    ev = p.key_averages()
    total = sum(e.device_time_total for e in ev)
    print(f"=== {tag}: total device time {total / 1e3:.2f} ms")
    for e in sorted(ev, key=lambda e: -e.device_time_total)[:rows]:
        print(f"{e.device_time_total / 1e3:9.3f} ms {100 * e.device_time_total / total:5.1f}%  x{e.count:<5d} {e.key[:110]}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--preset", default="nvfp4")
    ap.add_argument("--model", default="8b")
    a = ap.parse_args()
    import modelopt_b200 as mtq
    from modelopt_b200 import config as qcfg
    from modelopt_b200 import model_calib

    dev = torch.device("cuda", 0)
    model = ml.build_model(a.model, dev, layers=a.layers)
    ids = torch.randint(0, 1000, (8, 2048), device=dev)

    def fwd():
        with torch.no_grad():
            model(input_ids=ids, use_cache=False)

    prof("plain forward", fwd)
    mtq.quantize(model, getattr(qcfg, ml.PRESETS[a.preset]), lambda m: fwd())
    model_calib.enable_stats_collection(model)
    prof("calibrating forward", fwd)
    model_calib.finish_stats_collection(model)
    prof("fake-quant forward", fwd)


if __name__ == "__main__":
    main()
