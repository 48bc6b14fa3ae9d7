"""Where the fixed cost of histogram calibration goes on a full-depth Llama (tiny batch, GPU work negligible): wall time of
replace / configure / enable / one forward / finish_stats_collection(method='entropy') and a cProfile of the finish phase.
Developer tool."""
import cProfile
import copy
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200"), os.path.join(ROOT, "tools")]
import torch  # noqa: E402

import model_level as ml  # noqa: E402
import modelopt_b200 as mtq  # noqa: E402
from modelopt_b200 import config as qcfg  # noqa: E402
from modelopt_b200 import model_calib  # noqa: E402

dev = torch.device("cuda", 0)
model = ml.build_model("8b", dev)
ids = torch.randint(0, 1000, (2, 512), device=dev)
hist_cfg = copy.deepcopy(qcfg.INT8_DEFAULT_CFG["quant_cfg"])
for e in hist_cfg:
    if e["quantizer_name"] == "*weight_quantizer":
        e.pop("cfg", None)
        e["enable"] = False
    elif e["quantizer_name"] == "*input_quantizer":
        e["cfg"] = {"num_bits": 8, "axis": None, "calibrator": "histogram"}


def fwd():
    with torch.no_grad():
        model(input_ids=ids, use_cache=False)


fwd()
torch.cuda.synchronize()


def timed(name, fn):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    fn()
    torch.cuda.synchronize()
    print(f"{name:44s} {1e3 * (time.perf_counter() - t0):8.1f} ms", flush=True)


timed("replace_quant_module", lambda: mtq.replace_quant_module(model))
timed("set_quantizer_by_cfg", lambda: mtq.set_quantizer_by_cfg(model, hist_cfg))
timed("enable_stats_collection", lambda: model_calib.enable_stats_collection(model))
timed("forward 1 (first batch: amax + init + bin)", fwd)
timed("forward 2", fwd)
timed("forward 3", fwd)
pr = cProfile.Profile()
pr.enable()
timed("finish_stats_collection(entropy)", lambda: model_calib.finish_stats_collection(model, method="entropy"))
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(35)
