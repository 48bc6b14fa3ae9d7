"""cProfile of the host side of quantize() and of a quantized forward on a full-depth Llama with a tiny batch (GPU work
negligible): where the Python time per call goes.  Developer tool."""
import cProfile
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200"), os.path.join(ROOT, "tools")]
import torch  # noqa: E402

import model_level as ml  # noqa: E402
import modelopt_b200 as mtq  # noqa: E402
from modelopt_b200 import config as qcfg  # noqa: E402

preset = sys.argv[1] if len(sys.argv) > 1 else "nvfp4"
dev = torch.device("cuda", 0)
model = ml.build_model("8b", dev)
ids = torch.randint(0, 1000, (1, 300), device=dev)


def fwd(m=None):
    with torch.no_grad():
        model(input_ids=ids, use_cache=False)


fwd()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
mtq.quantize(model, getattr(qcfg, ml.PRESETS[preset]), fwd)
torch.cuda.synchronize()
pr.disable()
print("=== quantize() with one tiny batch")
pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
for _ in range(3):
    fwd()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(10):
    fwd()
torch.cuda.synchronize()
pr.disable()
print("=== 10 quantized forwards")
pstats.Stats(pr).sort_stats("tottime").print_stats(40)
