"""CPU: the re-charting oracle (oracle/charts.py) against goldens made by the reference's own code
(tests/golden/make_golden_charts.py: build_charts source text executed unchanged, the real JaggedTexture class)."""
import os

import numpy as np
import pytest

from common import GOLDEN_DIR


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(os.path.join(GOLDEN_DIR, "charts_golden.npz")))


@pytest.mark.parametrize("name", ["small", "wide", "tiny"])
def test_build_charts_oracle_matches_reference(gold, name):
    import oracle.charts as C

    g = {k[len(f"charts_{name}_"):]: v for k, v in gold.items() if k.startswith(f"charts_{name}_")}
    dims, mappings, pixel_scale = C.build_charts(g["scales"], float(g["pixel_num"]), 3.0, area_sum=float(g["area_sum"]))
    # numpy's and torch's CPU exp differ in the last ulp for some inputs (on the GPU both sides call the same expf): the
    # lengths, and with them mappings and -- in principle -- a chart edge, inherit that; the bisection path is the same
    assert abs(pixel_scale - float(g["pixel_scale"][0])) <= 2e-7 * pixel_scale   # the model keeps it in an fp32 buffer
    assert dims.dtype == np.int32 and float((dims != g["texture_dims"]).mean()) <= 1e-3
    np.testing.assert_allclose(mappings, g["mappings"], rtol=3e-7, atol=0)
    total = int(dims[-1, 2] + dims[-1, 0] * dims[-1, 1])
    assert abs(total - float(g["pixel_num"])) <= 1e-3 * float(g["pixel_num"]) + 1
    # with numpy's own area sum (last bits may differ) the bisection still lands within the reference's tolerance
    dims2, _, ps2 = C.build_charts(g["scales"], float(g["pixel_num"]), 3.0)
    assert abs(ps2 - pixel_scale) <= 1e-5 * pixel_scale and float((dims2 != dims).mean()) < 1e-3


@pytest.mark.parametrize("name", ["rgb", "six"])
def test_resample_oracle_matches_reference_jagged_texture(gold, name):
    import oracle.charts as C

    g = {k[len(f"resample_{name}_"):]: v for k, v in gold.items() if k.startswith(f"resample_{name}_")}
    out = C.resample_texture(g["old_dims"], g["old_texture"], g["new_dims"])
    assert out.shape == g["new_texture"].shape
    np.testing.assert_allclose(out, g["new_texture"], rtol=1.3e-6, atol=1e-5)      # the reference test's own tolerance
