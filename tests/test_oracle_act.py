"""CPU: the activation oracle (oracle/activations.py) against the golden vectors made from the reference's own source
text (tests/golden/make_golden_act.py executes gstex.py:1062-1066 and get_uv_mapping unchanged)."""
import os

import numpy as np

from common import GOLDEN_DIR


def _close(a, b, rtol, name):
    scale = float(np.abs(b).max())
    err = float(np.abs(a.astype(np.float64) - b).max())
    assert err <= rtol * scale + 1e-30, f"{name}: {err:.3e} at scale {scale:.3e}"


def test_activation_oracle_matches_reference_golden():
    import oracle.activations as A

    d = dict(np.load(os.path.join(GOLDEN_DIR, "act_golden.npz")))
    out = A.forward(d["quats_raw"], d["scales_raw"], d["opacities_raw"], d["mappings"])
    for nm, a in zip(["quats", "scales", "opacities", "uv0", "umap", "vmap"], out):
        assert a.shape == d[nm].shape and a.dtype == np.float32, nm
        _close(a, d[nm], 5e-7, nm)          # last-ulp differences between numpy's and torch's CPU kernels only
    assert np.array_equal(out[3], d["uv0"]) and float(out[1][0, 0]) == np.float32(1e-9)   # the clamp fired
    g = A.backward(d["quats_raw"], d["scales_raw"], d["quats"], d["scales"], d["opacities"], d["v_quats"], d["v_scales"],
                   d["v_opacities"])
    for nm, a in zip(["g_quats_raw", "g_scales_raw", "g_opacities_raw"], g):
        _close(a, d[nm], 2e-6, nm)
    assert np.all(g[1][:, 2] == 0) and np.all(g[1][:3, 0] == 0) and np.all(d["g_scales_raw"][:3, 0] == 0)
    no_maps = A.forward(d["quats_raw"], d["scales_raw"], d["opacities_raw"])
    assert no_maps[3] is None and np.array_equal(no_maps[0], out[0])
