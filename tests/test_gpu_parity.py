"""GPU: the CUDA path (through the C ABI) against the reference-generated golden fixtures and
against the oracle.  Bars (BASELINE.json north_star): projected pixel indices and inlier masks
bit-exact, per-iteration cost within 1e-4 relative, final pose within 1e-3 deg / 1e-4 m.  The
implementation is float64 throughout, so the tests also hold it to a much tighter bar (TIGHT)."""
import numpy as np
import pytest
import torch

from util import (COST_RTOL, ROT_TOL_DEG, TRANS_TOL_M, assert_costs_close, chw64, golden_names, load_golden,
                  rot_angle_deg)

pytestmark = pytest.mark.gpu
TIGHT = 1e-9


def _fm():
    import fmpnp_b200 as fm
    return fm


def problem_from_golden(g, dev="cuda"):
    fm = _fm()
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(device=dev, dtype=dt).contiguous()  # noqa: E731
    return fm.Problem(t(g["F"], torch.float32), t(g["Gx"], torch.float32), t(g["Gy"], torch.float32),
                      t(g["ref"], torch.float32), t(g["pts3D"], torch.float64), g["K"], g["R0"], g["t0"],
                      g["im_width"], g["im_height"])


def check_against(name, res, hist, exp, use_ratio, b=0):
    """exp: dict with costs, points2d, mask, kept, Rs, ts, out_R, out_t, best_cost, best_num_inliers."""
    n = int(hist["len"][b])
    N = exp["points2d"].shape[1]
    assert n == len(exp["costs"]), "%s: %d history entries, expected %d" % (name, n, len(exp["costs"]))
    assert np.array_equal(hist["points2d"][b, :n, :N], exp["points2d"]), name + ": projected pixel indices differ"
    assert np.array_equal(hist["mask"][b, :n, :N].astype(bool), exp["mask"].astype(bool)), name + ": support masks differ"
    if use_ratio:
        assert np.array_equal(hist["kept"][b, :n, :N].astype(bool), exp["kept"].astype(bool)), name + ": ratio masks differ"
    assert_costs_close(hist["costs"][b, :n], exp["costs"], COST_RTOL, name)
    assert_costs_close(hist["costs"][b, :n], exp["costs"], TIGHT, name + " (tight)")
    R = np.array(res["R"]).reshape(3, 3)
    t = np.array(res["t"])
    assert rot_angle_deg(R, exp["out_R"]) <= ROT_TOL_DEG, name
    assert np.abs(t - exp["out_t"]).max() <= TRANS_TOL_M, name
    assert np.abs(R - exp["out_R"]).max() <= 1e-8 and np.abs(t - exp["out_t"]).max() <= 1e-8, name + " (tight pose)"
    np.testing.assert_allclose(hist["Rs"][b, :n], exp["Rs"], rtol=0, atol=1e-8)
    np.testing.assert_allclose(hist["ts"][b, :n], exp["ts"], rtol=0, atol=1e-8)
    assert int(res["best_num_inliers"]) == int(exp["best_num_inliers"]), name
    assert abs(float(res["best_cost"]) - exp["best_cost"]) <= TIGHT * abs(exp["best_cost"]), name


@pytest.mark.parametrize("solver,route", [("cholesky", "auto"), ("lu", "auto"), ("cholesky", "tasks2"),
                                          ("cholesky", "tasks128")], indirect=["route"])
@pytest.mark.parametrize("name", golden_names())
def test_c_abi_matches_reference_fixture(name, solver, route):
    fm = _fm()
    g = load_golden(name)
    prob = problem_from_golden(g)
    out = fm.refine_batch([prob], g["n_iters"], loss=g["loss"], lambda_=g["lambda_"],
                          ratio_threshold=g["ratio_threshold"], levels=g["levels"], track=True, solver=solver,
                          barron_alpha=g["barron_alpha"] if g["loss"] == "barron_loss" else None)
    res, hist = out["results"][0], out["history"]
    check_against(name, res, hist, g, g["ratio_threshold"] is not None)
    assert abs(float(res["initial_cost"]) - g["initial_cost"]) <= TIGHT * abs(g["initial_cost"])
    assert int(res["status"]) == 0


def test_notebook_known_answer_on_gpu():
    """FeatureBA_ToyExample.ipynb cell 28: 27497.41105769231 -> 276.125, through the drop-in model.py API."""
    import model
    g = load_golden("toy6_squared_lm")
    m = model.sparseFeaturePnP(n_iters=50, lambda_=0.01)
    R, t = m(torch.from_numpy(g["pts3D"]), torch.from_numpy(g["ref"]).double(), torch.from_numpy(chw64(g["F"])),
             torch.from_numpy(chw64(g["Gx"])), torch.from_numpy(chw64(g["Gy"])), torch.from_numpy(g["K"]),
             g["im_width"], g["im_height"], R_init=torch.from_numpy(g["R0"]), t_init=torch.from_numpy(g["t0"]),
             track=True)
    c = m.track_["costs"]
    assert len(c) == 51 and c[0] == 27497.41105769231 and c[-1] == 276.125
    assert R.dtype == torch.float64 and R.device.type == "cpu" and tuple(R.shape) == (3, 3) and tuple(t.shape) == (3,)
    assert m.best_num_inliers_ == 14 and m.best_cost_.item() == 276.125
    assert m.initial_cost_.item() == 27497.41105769231


@pytest.mark.parametrize("name", ["syn_geman_ratio05", "syn_pyramid_ratio06", "syn_oddC_pyramid", "syn_dyn_squared"])
def test_dropin_model_api(name):
    """Same class / call signature / attributes / track_ schema as the reference's model.py."""
    import helpers.utils as hu
    import model
    model.clear_pack_cache()
    g = load_golden(name)
    loss_fn = getattr(hu, g["loss"])
    m = model.sparseFeaturePnP(n_iters=g["n_iters"], loss_fn=loss_fn, lambda_=g["lambda_"],
                               ratio_threshold=g["ratio_threshold"])
    args = (torch.from_numpy(g["pts3D"]), torch.from_numpy(g["ref"]).double(), torch.from_numpy(chw64(g["F"])),
            torch.from_numpy(chw64(g["Gx"])), torch.from_numpy(chw64(g["Gy"])), torch.from_numpy(g["K"]),
            g["im_width"], g["im_height"])
    kw = dict(R_init=torch.from_numpy(g["R0"]), t_init=torch.from_numpy(g["t0"]), track=True)
    if g["levels"] is None:
        R, t = m(*args, **kw)
    else:
        R, t = m.multilevel_optimization([(s, e, None, None) for s, e in g["levels"]], *args, **kw)
    tr = m.track_
    assert set(tr) == {"Rs", "ts", "costs", "points2d", "mask", "threshold_mask"}
    n = len(g["costs"])
    assert len(tr["costs"]) == n and all(isinstance(c, float) for c in tr["costs"])
    assert_costs_close(tr["costs"], g["costs"], TIGHT, name)
    for e in range(n):
        assert tr["points2d"][e].dtype == torch.int32 and np.array_equal(tr["points2d"][e].numpy(), g["points2d"][e])
        assert tr["mask"][e].dtype == torch.bool and np.array_equal(tr["mask"][e].numpy(), g["mask"][e].astype(bool))
        if g["ratio_threshold"] is None:
            assert tr["threshold_mask"][e] is None
        else:
            exp = g["kept"][e].astype(bool)[g["mask"][e].astype(bool)]
            assert np.array_equal(tr["threshold_mask"][e].numpy(), exp)
    assert rot_angle_deg(R.numpy(), g["out_R"]) <= ROT_TOL_DEG and np.abs(t.numpy() - g["out_t"]).max() <= TRANS_TOL_M
    assert m.best_num_inliers_ == g["best_num_inliers"]
    assert abs(m.best_cost_.item() - g["best_cost"]) <= TIGHT * abs(g["best_cost"])
    assert abs(m.initial_cost_.item() - g["initial_cost"]) <= TIGHT * abs(g["initial_cost"])


def test_find_inliers_and_compute_cost():
    import helpers.utils as hu
    import model
    fm = _fm()
    g = load_golden("syn_inliers_cost")
    prob = problem_from_golden(g)
    for name in ("squared_loss", "geman_mcclure_loss", "cauchy_loss"):
        m = fm.find_inliers_batch([prob], 0.8, loss=name)[0].cpu().numpy().astype(bool)
        assert np.array_equal(m, g["inliers_" + name].astype(bool)), name
        m2 = model.find_inliers(torch.from_numpy(g["pts3D"]), torch.from_numpy(g["R0"]), torch.from_numpy(g["t0"]),
                                torch.from_numpy(chw64(g["F"])), torch.from_numpy(g["ref"]).double(),
                                torch.from_numpy(g["K"]), g["im_width"], g["im_height"], threshold=0.8,
                                loss_fn=getattr(hu, name))
        assert m2.dtype == torch.bool and np.array_equal(m2.numpy(), g["inliers_" + name].astype(bool))
    c, ns = fm.compute_cost_batch([prob])
    assert abs(c[0].item() - float(g["compute_cost"])) <= TIGHT * float(g["compute_cost"])
    c2, _ = fm.compute_cost_batch([prob], ratio_threshold=0.7)
    assert abs(c2[0].item() - float(g["compute_cost_ratio07"])) <= TIGHT * float(g["compute_cost_ratio07"])
    mm = model.sparseFeaturePnP(n_iters=1, ratio_threshold=0.7)
    c3 = mm.compute_cost(torch.from_numpy(g["pts3D"]), torch.from_numpy(g["R0"]), torch.from_numpy(g["t0"]),
                         torch.from_numpy(chw64(g["F"])), torch.from_numpy(g["ref"]).double(), torch.from_numpy(g["K"]),
                         g["im_width"], g["im_height"])
    assert abs(c3.item() - float(g["compute_cost_ratio07"])) <= TIGHT * float(g["compute_cost_ratio07"])


def test_forward_many_equals_forward_packed_per_query():
    """model.sparseFeaturePnP.forward_many: a list of device-resident problems in one launch gives, query by query, what
    forward_packed / multilevel_packed return one at a time (same records: the reduction order is fixed)."""
    import helpers.utils as hu
    import model
    import fmpnp_b200 as fm
    from fmpnp_b200 import synth
    Ks = ((400.0, 0.0, 255.5), (0.0, 400.0, 254.3), (0.0, 0.0, 1.0))
    scs = [synth.make_scene(7300 + i, N=n, C=64, Hf=32, Wf=32, im_width=512, im_height=512, K=Ks, grad_scale=1.0 / 64,
                            outlier_frac=0.1) for i, n in enumerate((700, 1024, 333, 2000))]
    probs = [fm.Problem(sc.F.cuda(), sc.Gx.cuda(), sc.Gy.cuda(), sc.ref.cuda(), sc.pts3D.cuda(), sc.K, sc.R0, sc.t0,
                        sc.im_width, sc.im_height) for sc in scs]
    m = model.sparseFeaturePnP(n_iters=10, loss_fn=hu.geman_mcclure_loss, lambda_=0.01, ratio_threshold=None)
    for levels in (None, [(32, 64), (0, 32)]):
        many = m.forward_many(probs, levels=levels)
        assert len(many) == len(probs) and m.last_results_.shape == (len(probs),)
        for p, (R, t) in zip(probs, many):
            R1, t1 = m.forward_packed(p) if levels is None else m.multilevel_packed(levels, p)
            assert R.dtype == torch.float64 and torch.equal(R, R1) and torch.equal(t, t1)
