"""CPU: the SH oracle (oracle/gstex_oracle_sh.c) against the golden vectors made by the reference's own Python
(tests/golden/make_golden_sh.py), plus the degrees_to_use < degree rule of the CUDA kernels (sh.cuh:47-119)."""
import os

import numpy as np
import pytest

import oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden", "sh_golden.npz")


@pytest.mark.parametrize("degree", [0, 1, 2, 3, 4])
def test_sh_oracle_matches_reference_python(degree):
    d = np.load(GOLD)
    v, c, w = d[f"viewdirs_{degree}"], d[f"coeffs_{degree}"], d[f"w_{degree}"]
    col = oracle.spherical_harmonics(degree, v, c)
    np.testing.assert_allclose(col, d[f"colors_{degree}"], rtol=1.3e-6, atol=1e-5)  # test_sh.py's assert_close
    vc = oracle.spherical_harmonics_vjp(degree, c.shape[1], v, w)
    np.testing.assert_allclose(vc, d[f"v_coeffs_{degree}"], rtol=1.3e-6, atol=1e-5)


def test_sh_oracle_partial_degree_and_scaling():
    g = np.random.default_rng(3)
    v = g.standard_normal((64, 3)).astype(np.float32) * 7.0  # not normalised: the kernels normalise (sh.cuh:62-67)
    c = g.standard_normal((64, 25, 3)).astype(np.float32)
    for use in range(5):
        nb = oracle.num_sh_bases(use)
        full = oracle.spherical_harmonics(use, v, c)
        cut = oracle.spherical_harmonics(use, v / np.linalg.norm(v, axis=-1, keepdims=True), np.ascontiguousarray(c[:, :nb]))
        np.testing.assert_allclose(full, cut, rtol=1e-5, atol=1e-6)
        vc = oracle.spherical_harmonics_vjp(use, 25, v, np.ones((64, 3), np.float32))
        assert np.all(vc[:, nb:] == 0) and np.all(vc[:, 0] == np.float32(0.28209479177387814))
    # linearity in the coefficients (size-independent property used at full size on the GPU)
    a = oracle.spherical_harmonics(4, v, c) + oracle.spherical_harmonics(4, v, 2 * c)
    np.testing.assert_allclose(a, oracle.spherical_harmonics(4, v, 3 * c), rtol=1e-5, atol=1e-5)
