"""GPU: INTEGRATION.md section 3 executed against the UNMODIFIED reference — `register_quant_backend("b200", fn)` with `fn`
written only against include/modelopt_b200.h through ctypes, the reference's own `mtq.quantize` with `{"backend": "b200"}`,
amax and outputs compared bit for bit with the reference's own GPU kernels (mirrors the reference's
tests/unit/torch/quantization/test_custom_backend.py:33-72).  Needs the reference package: /root/reference (build container)
or a copy under oracle/_ref (see tools/reference_backend_check.py); skipped otherwise.  Runs in a fresh process because the
reference and modelopt_b200 both register `tensorrt::quantize_op`."""

import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def test_reference_quantize_with_b200_backend_is_bit_identical():
    if not any(os.path.isdir(os.path.join(p, "modelopt")) for p in ("/root/reference", os.path.join(ROOT, "oracle", "_ref"))):
        pytest.skip("the reference package is not present on this machine")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "reference_backend_check.py")], capture_output=True,
                       text=True, timeout=1500)
    print(r.stdout[-4000:], r.stderr[-2000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "REFERENCE BACKEND CHECK OK" in r.stdout
