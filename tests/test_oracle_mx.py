"""CPU: oracle/mx.py's element rounding equals the reference's own header (oracle/_ref/libref_mx.so, g++-compiled
from /root/reference in the build container) and a few hand-checked E8M0 cases."""

import numpy as np
import pytest
import torch

from oracle import mx

ELEM = ["E4M3", "E5M2", "INT8", "E0M3", "E1M2", "E3M0", "E2M1", "E3M2", "E2M3"]


@pytest.mark.skipif(mx.ref_lib() is None, reason="oracle/_ref/libref_mx.so not built (needs /root/reference)")
@pytest.mark.parametrize("fmt", ELEM)
def test_convert_matches_reference_header(fmt):
    rng = np.random.default_rng(0)
    top = mx.FMT_MAX[fmt]
    vals = torch.tensor(mx._TABLES[fmt]) if fmt in mx._TABLES else torch.linspace(0, top, 513)
    mids = (vals[:-1] + vals[1:]) / 2
    a = np.concatenate([rng.uniform(0, top * 1.2, 20000), rng.uniform(0, 1, 5000) ** 4 * top, vals.numpy(), mids.numpy(),
                        np.nextafter(mids.numpy().astype(np.float32), np.float32(0)),
                        np.nextafter(mids.numpy().astype(np.float32), np.float32(1e9)),
                        [0.0, top, top * 2, 1e30, np.inf]]).astype(np.float32)
    if fmt == "INT8":
        a = a[a < 2e9]  # (int)rint(x) overflows on the host (UB); the device conversion saturates
    want = mx.ref_convert(a, fmt)
    got = mx.convert(torch.from_numpy(a), fmt).numpy()
    assert np.array_equal(got, want), np.flatnonzero(got != want)[:10]
    assert mx.ref_lib().ref_mx_format_max(mx.FMT[fmt]) == top


def test_e8m0_exponent_cases():
    # unscale = 2^ceil(log2(amax / dmax)): exact powers of two stay, anything above goes up
    amax = torch.tensor([448.0, 449.0, 224.0, 225.0, 6.0, 1e-30])
    e = mx.e8m0_exponent(amax, 448.0)
    assert e.tolist()[:4] == [0, 1, -1, 0]
    assert mx.e8m0_exponent(torch.tensor([6.0, 6.5, 3.0]), 6.0).tolist() == [0, 1, -1]


@pytest.mark.parametrize("fmt,block", [("E4M3", 32), ("E2M1", 32), ("E3M2", 32), ("INT8", 32), ("E2M3", 16), ("E3M0", 8)])
def test_mx_fake_quant_properties(fmt, block):
    torch.manual_seed(1)
    x = (torch.randn(8, 96) * 3).to(torch.bfloat16)
    y = mx.mx_fake_quant(x, block, fmt)
    assert y.shape == x.shape and y.dtype == x.dtype
    # idempotent: representable values under a power-of-two scale stay put
    assert torch.equal(mx.mx_fake_quant(y, block, fmt), y)
    # error bounded by half the largest grid step of the format times the block scale
    assert (y.float() - x.float()).abs().max() <= x.float().abs().max()
    z = torch.zeros(2, 64, dtype=torch.float16)
    assert torch.equal(mx.mx_fake_quant(z, block, fmt), z)
