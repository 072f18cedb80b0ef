"""CPU: pins the C oracle (oracle/fmpnp_oracle.c) against
  (1) the reference notebook's stored known answer (toyexample_6: 27497.41105769231 -> 276.125), and
  (2) every fixture produced by running the unmodified reference (oracle/gen_golden.py).
Pixel indices and masks must be bit-exact; costs / poses within float64 re-association noise."""
import numpy as np
import pytest

import oracle
from util import (ORACLE_COST_RTOL, ORACLE_POSE_ATOL, assert_costs_close, chw64, golden_names, load_golden)


def run_oracle(g, track=True):
    args = (g["pts3D"], g["ref"].astype(np.float64), chw64(g["F"]), chw64(g["Gx"]), chw64(g["Gy"]), g["K"],
            g["im_width"], g["im_height"], g["R0"], g["t0"], g["n_iters"])
    kw = dict(loss=g["loss"], lambda_=g["lambda_"], ratio_threshold=g["ratio_threshold"],
              barron_alpha=g["barron_alpha"], track=track)
    if g["levels"] is None:
        return oracle.forward(*args, **kw)
    return oracle.multilevel(g["levels"], *args, **kw)


def test_notebook_known_answer():
    g = load_golden("toy6_squared_lm")
    # the numbers printed in FeatureBA_ToyExample.ipynb cell 28
    assert g["costs"][0] == 27497.41105769231 and g["costs"][-1] == 276.125 and len(g["costs"]) == 51
    o = run_oracle(g)
    c = o["track"]["costs"]
    assert len(c) == 51
    assert c[0] == 27497.41105769231
    assert c[-1] == 276.125
    # 17 "cost increased" lines in cell 24
    assert int(np.sum(np.diff(np.minimum.accumulate(c)) == 0) >= 17)


@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_reference_fixture(name):
    g = load_golden(name)
    o = run_oracle(g)
    tr = o["track"]
    assert_costs_close(tr["costs"], g["costs"], ORACLE_COST_RTOL, name)
    assert np.array_equal(tr["points2d"], g["points2d"]), name + ": projected pixel indices differ"
    assert np.array_equal(tr["mask"], g["mask"].astype(bool)), name + ": support masks differ"
    if g["ratio_threshold"] is not None:
        assert np.array_equal(tr["threshold_mask"], g["kept"].astype(bool)), name + ": ratio-test masks differ"
    np.testing.assert_allclose(tr["Rs"], g["Rs"], rtol=0, atol=ORACLE_POSE_ATOL)
    np.testing.assert_allclose(tr["ts"], g["ts"], rtol=0, atol=ORACLE_POSE_ATOL)
    np.testing.assert_allclose(o["R"], g["out_R"], rtol=0, atol=ORACLE_POSE_ATOL)
    np.testing.assert_allclose(o["t"], g["out_t"], rtol=0, atol=ORACLE_POSE_ATOL)
    assert o["best_num_inliers"] == g["best_num_inliers"]
    assert abs(o["best_cost"] - g["best_cost"]) <= ORACLE_COST_RTOL * abs(g["best_cost"])
    ic = o["initial_cost"] if g["levels"] is not None else o["first_cost"]
    assert abs(ic - g["initial_cost"]) <= ORACLE_COST_RTOL * abs(g["initial_cost"])


def test_find_inliers_and_compute_cost():
    g = load_golden("syn_inliers_cost")
    F = chw64(g["F"])
    ref = g["ref"].astype(np.float64)
    for name in ("squared_loss", "geman_mcclure_loss", "cauchy_loss"):
        m = oracle.find_inliers(g["pts3D"], g["R0"], g["t0"], F, ref, g["K"], g["im_width"], g["im_height"], 0.8,
                                loss=name)
        assert np.array_equal(m, g["inliers_" + name].astype(bool)), name
    c = oracle.compute_cost(g["pts3D"], ref, F, g["K"], g["im_width"], g["im_height"], g["R0"], g["t0"])
    assert abs(c - float(g["compute_cost"])) <= ORACLE_COST_RTOL * abs(float(g["compute_cost"]))
    c2 = oracle.compute_cost(g["pts3D"], ref, F, g["K"], g["im_width"], g["im_height"], g["R0"], g["t0"],
                             ratio_threshold=0.7)
    assert abs(c2 - float(g["compute_cost_ratio07"])) <= ORACLE_COST_RTOL * abs(float(g["compute_cost_ratio07"]))


def test_fixture_generator_imports_the_reference_own_helpers():
    """Where the reference is present (the build container), oracle/pyref.py must hand gen_golden*.py the reference's
    model.py AND the reference's helpers/utils.py -- not the product's drop-in `helpers` package, which is a regular
    package and would otherwise shadow the reference's namespace package on sys.path."""
    import os
    import sys
    import pyref
    if not pyref.available():
        pytest.skip("reference not present (GPU box)")
    import helpers.utils as product_utils  # the product's package is importable first, as in the generators
    ref_model, ref_utils = pyref.load()
    root = os.path.realpath(pyref.REFERENCE_ROOT)
    assert os.path.realpath(ref_model.__file__).startswith(root) and os.path.realpath(ref_utils.__file__).startswith(root)
    assert ref_utils is not product_utils and hasattr(ref_utils, "kornia")
    assert ref_model.squared_loss is ref_utils.squared_loss and ref_model.so3exp_map is ref_utils.so3exp_map
    assert sys.modules.get("helpers.utils") is product_utils  # and the product's modules are back in place
