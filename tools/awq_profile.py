"""Kernel-time breakdown (torch.profiler) of the awq_lite search on one Llama-3-8B decoder layer.  Developer tool."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200"), os.path.join(ROOT, "tools")]
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import model_level as ml  # noqa: E402
import modelopt_b200 as mtq  # noqa: E402
from modelopt_b200 import config as qcfg  # noqa: E402

dev = torch.device("cuda", 0)
model = ml.build_model("8b", dev, layers=1).model
ids = [torch.randint(0, 1000, (8, 2048), device=dev) for _ in range(2)]


def loop(m):
    with torch.no_grad():
        for b in ids:
            m(input_ids=b, use_cache=False)


loop(model)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as p:
    mtq.quantize(model, qcfg.INT4_AWQ_CFG, loop)
    torch.cuda.synchronize()
ev = p.key_averages()
total = sum(e.device_time_total for e in ev)
print(f"awq_lite on one layer, 2 micro-batches: total device time {total / 1e3:.2f} ms")
for e in sorted(ev, key=lambda e: -e.device_time_total)[:16]:
    print(f"{e.device_time_total / 1e3:9.3f} ms {100 * e.device_time_total / total:5.1f}%  x{e.count:<5d} {e.key[:120]}")
