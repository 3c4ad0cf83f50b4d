"""CPU: the sampling oracle (oracle/gstex_oracle_sample.c) against the golden vectors made by the reference's own
Python (tests/golden/make_golden_sample.py)."""
import os

import numpy as np
import pytest

import oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden", "sample_golden.npz")


@pytest.mark.parametrize("name", ["ref_test", "border", "wide"])
def test_sample_oracle_matches_reference_python(name):
    d = np.load(GOLD)
    tex = d[f"{name}_texture"]
    out = oracle.texture_sample((0, 1, tex.shape[1]), d[f"{name}_texture_dims"], tex, d[f"{name}_uvs"])
    np.testing.assert_allclose(out, d[f"{name}_out"], rtol=1.3e-6, atol=1e-5)  # tests/test_sample.py's assert_close


def test_sample_oracle_properties():
    g = np.random.default_rng(5)
    dims = np.array([[4, 5, 0], [1, 1, 20], [3, 2, 21]], np.int32)
    tex = g.random((27, 3)).astype(np.float32)
    # uv on the texel grid returns the texel itself; a 1x1 chart is constant; u = 1 replicates the last row
    qd = np.stack([dims[0], dims[0], dims[1], dims[1], dims[2]])
    uv = np.array([[0.25, 0.4], [0.0, 0.0], [0.3, 0.9], [1.0, 1.0], [1.0, 0.5]], np.float32)
    out = oracle.texture_sample((3, 1, 3), qd, tex, uv)
    np.testing.assert_allclose(out[0], tex[0 + 1 * 5 + 2], rtol=1e-6)
    np.testing.assert_allclose(out[1], tex[0], rtol=1e-6)
    np.testing.assert_allclose(out[2], tex[20], rtol=1e-6)
    np.testing.assert_allclose(out[3], tex[20], rtol=1e-6)
    np.testing.assert_allclose(out[4], tex[21 + 2 * 2 + 1], rtol=1e-6)
