"""Scratch: locate mismatches of per-axis FP8 QDQ against the oracle."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tensorrt-model-optimizer_b200"), os.path.join(ROOT, "tests")]
import torch
from modelopt_b200 import ops
from oracle import fake_quant as ofq

torch.manual_seed(0)
for dt in (torch.float32, torch.bfloat16):
    x = (torch.randn(6, 8, 256) * 3).to(dt)
    for axis in (0, 1, 2):
        red = [i for i in range(3) if i != axis]
        am = x.float().abs().amax(dim=red, keepdim=True)
        got = ops.qdq_fp8(x.cuda(), am.cuda()).cpu()
        want = ofq.fp8_qdq(x, am)
        bad = (got.float() != want.float()).nonzero()
        print(dt, "axis", axis, "mismatch", len(bad), "of", x.numel())
        for idx in bad[:6]:
            i = tuple(idx.tolist())
            a = am.flatten()[i[axis]].item()
            print("   idx", i, "x", x[i].item(), "amax", a, "got", got[i].item(), "want", want[i].item(),
                  "scale", 448.0 / a)
