"""Golden vectors for `disable_pre_quant_scale_and_resmooth` (reference model_calib.py:1169-1204), from the REAL reference on CPU.
TEST INFRASTRUCTURE ONLY.  Build container:  PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden_resmooth.py
Reuses the model and batches of tests/golden/quantize_cpu.npz; writes tests/golden/resmooth_cpu.npz."""
import copy
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.dirname(HERE)]
from oracle.gen_golden import OUT, to_np  # noqa: E402  (also puts the reference and its shims on sys.path)


def main():
    import modelopt.torch.quantization as mtq
    from modelopt.torch.quantization.model_calib import disable_pre_quant_scale_and_resmooth
    from torch import nn

    q = np.load(os.path.join(OUT, "quantize_cpu.npz"))
    g = {}
    for name in ("INT8_SMOOTHQUANT_CFG", "INT4_AWQ_CFG"):
        m = nn.Sequential(nn.Linear(256, 512), nn.ReLU(), nn.Linear(512, 256, bias=False))
        m.load_state_dict({k[2:]: torch.from_numpy(q[k]) for k in q.files if k.startswith("w_")})
        batches = [torch.from_numpy(q[f"batch{i}"]) for i in range(3)]
        m = mtq.quantize(m, copy.deepcopy(getattr(mtq, name)), lambda model: [model(b) for b in batches])
        for idx, delete in (("0", False), ("2", True)):
            disable_pre_quant_scale_and_resmooth(m[int(idx)], delete_pre_quant_scale=delete)
        for n, mod in m.named_modules():
            if isinstance(mod, mtq.nn.TensorQuantizer):
                g[f"{name}|{n}|has_pqs"] = np.asarray(hasattr(mod, "_pre_quant_scale"))
                g[f"{name}|{n}|pqs_enabled"] = np.asarray(bool(mod._enable_pre_quant_scale))
                if hasattr(mod, "_amax"):
                    g[f"{name}|{n}|_amax"] = to_np(mod._amax.float())
        for k, v in m.state_dict().items():
            if "quantizer" not in k:
                g[f"{name}|weight|{k}"] = to_np(v)
        with torch.no_grad():
            g[f"{name}|out"] = to_np(m(torch.from_numpy(q["test"])))
    np.savez_compressed(os.path.join(OUT, "resmooth_cpu.npz"), **g)
    print("resmooth_cpu.npz", len(g), "arrays")


if __name__ == "__main__":
    main()
