"""GPU: shapes and exit paths beyond the golden fixtures, checked against the oracle, plus
size-independent properties at the BASELINE.json sizes (N = 4096, C = 1024, 128 x 128 maps)."""
import os

import numpy as np
import pytest
import torch

import oracle
from util import COST_RTOL, ROT_TOL_DEG, TRANS_TOL_M, assert_costs_close, rot_angle_deg

pytestmark = pytest.mark.gpu
TIGHT = 1e-9
KS = ((100.0, 0.0, 127.055733), (0.0, 100.0, 124.546844), (0.0, 0.0, 1.0))
SMALL = dict(Hf=32, Wf=32, im_width=256, im_height=256, K=KS)


def fm():
    import fmpnp_b200
    return fmpnp_b200


def same_records(a, b, skip=()):
    """Field-wise bit equality of fmpnp_result records (the struct has 4 padding bytes)."""
    a, b = np.atleast_1d(a), np.atleast_1d(b)
    return a.shape == b.shape and all(
        np.ascontiguousarray(a[f]).tobytes() == np.ascontiguousarray(b[f]).tobytes() for f in a.dtype.names
        if f not in skip)


def assert_reuse_is_invisible(out, full, B):
    """`out` (records of points that stay in their cell are reused) against `full` (gather_every_row=True): the
    same bits in every result field and history entry; only the row counters tell the two apart."""
    assert same_records(out["results"], full["results"], skip=("gathered_rows", "reused_rows"))
    g, r = out["results"]["gathered_rows"], out["results"]["reused_rows"]
    assert np.array_equal(g + r, full["results"]["gathered_rows"]) and not full["results"]["reused_rows"].any()
    assert (g >= 0).all() and (r >= 0).all()
    if "history" in out:
        for b in range(B):
            n = int(out["history"]["len"][b])
            assert n == int(full["history"]["len"][b])
            for k in ("Rs", "ts", "costs", "points2d", "mask", "kept"):
                assert np.array_equal(out["history"][k][b, :n], full["history"][k][b, :n], equal_nan=k in ("Rs", "ts", "costs")), k


def synth():
    from fmpnp_b200 import synth as s
    return s


def to_problem(sc, active=None, R0=None, t0=None):
    dev = "cuda"
    return fm().Problem(sc.F.to(dev), sc.Gx.to(dev), sc.Gy.to(dev), sc.ref.to(dev), sc.pts3D.to(dev), sc.K,
                        sc.R0 if R0 is None else R0, sc.t0 if t0 is None else t0, sc.im_width, sc.im_height,
                        active=None if active is None else torch.as_tensor(active, dtype=torch.uint8, device=dev))


def run_oracle(sc, n_iters, loss, lambda_, ratio, levels=None, sel=None, R0=None, t0=None):
    inp = sc.reference_inputs()
    pts, ref = inp["pts3D"].numpy(), inp["feature_ref"].numpy()
    if sel is not None:
        pts, ref = np.ascontiguousarray(pts[sel]), np.ascontiguousarray(ref[sel])
    args = (pts, ref, inp["feature_map_query"].numpy(), inp["feature_grad_x"].numpy(), inp["feature_grad_y"].numpy(),
            inp["K"].numpy(), sc.im_width, sc.im_height, inp["R_init"].numpy() if R0 is None else np.asarray(R0),
            inp["t_init"].numpy() if t0 is None else np.asarray(t0), n_iters)
    kw = dict(loss=loss, lambda_=lambda_, ratio_threshold=ratio, track=True)
    return oracle.forward(*args, **kw) if levels is None else oracle.multilevel(levels, *args, **kw)


def compare(name, res, hist, o, b=0, use_ratio=False, sel=None):
    tr = o["track"]
    n = int(hist["len"][b])
    assert n == len(tr["costs"]), "%s: %d entries vs oracle %d" % (name, n, len(tr["costs"]))
    N = tr["points2d"].shape[1] if n else 0
    px, mk, kp = hist["points2d"][b, :n], hist["mask"][b, :n].astype(bool), hist["kept"][b, :n].astype(bool)
    if sel is not None:
        px, mk, kp = px[:, sel], mk[:, sel], kp[:, sel]
    else:
        px, mk, kp = px[:, :N], mk[:, :N], kp[:, :N]
    if n:
        assert np.array_equal(px, tr["points2d"]), name + ": pixel indices"
        assert np.array_equal(mk, tr["mask"]), name + ": support mask"
        if use_ratio:
            assert np.array_equal(kp, tr["threshold_mask"]), name + ": ratio mask"
        assert_costs_close(hist["costs"][b, :n], tr["costs"], COST_RTOL, name)
        assert_costs_close(hist["costs"][b, :n], tr["costs"], TIGHT, name + " (tight)")
    R, t = np.array(res["R"]).reshape(3, 3), np.array(res["t"])
    assert rot_angle_deg(R, o["R"]) <= ROT_TOL_DEG and np.abs(t - o["t"]).max() <= TRANS_TOL_M, name + ": pose"
    assert np.abs(R - o["R"]).max() < 1e-8 and np.abs(t - o["t"]).max() < 1e-8, name + ": pose (tight)"
    assert int(res["status"]) == int(o["status"]), "%s: status %d vs oracle %d" % (name, res["status"], o["status"])


# ----------------------------------------------------------------------------------------------
# wide channel counts: rows longer than a staging slot are split into channel chunks
# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("C,levels", [(1024, None), (640, None), (1024, [(512, 1024), (128, 512), (0, 128)]),
                                      (2048, [(1024, 2048), (256, 1024), (0, 256)]),    # robotcar_feature_pyramid.gin:61
                                      (1664, [(640, 1664), (128, 640), (0, 128)]),      # default_robotcar.gin:75
                                      (1664, [(1152, 1664), (640, 1152), (384, 640), (128, 384), (0, 128)]),  # :76
                                      (516, None), (20, [(8, 20), (0, 8)])])
def test_wide_and_chunked_channels(C, levels, route):
    sc = synth().make_scene(21, N=200, C=C, Hf=16, Wf=16, im_width=128, im_height=128,
                            K=((50.0, 0.0, 63.5), (0.0, 50.0, 62.3), (0.0, 0.0, 1.0)), grad_scale=1.0 / 64,
                            outlier_frac=0.1)
    out = fm().refine_batch([to_problem(sc)], 6, loss="geman_mcclure_loss", lambda_=0.01, levels=levels, track=True)
    o = run_oracle(sc, 6, "geman_mcclure_loss", 0.01, None, levels)
    compare("C=%d" % C, out["results"][0], out["history"], o)


# ----------------------------------------------------------------------------------------------
# a ragged batch: different N per query, one launch
# ----------------------------------------------------------------------------------------------
def test_ragged_batch_matches_oracle_and_single_runs(route):
    Ns = [1, 7, 33, 150, 257, 64, 300, 96, 5, 211]
    scs = [synth().make_scene(40 + i, N=n, C=32, grad_scale=1.0 / 64, outlier_frac=0.1, **SMALL) for i, n in enumerate(Ns)]
    probs = [to_problem(sc) for sc in scs]
    out = fm().refine_batch(probs, 8, loss="cauchy_loss", lambda_=0.01, ratio_threshold=0.9, track=True)
    for i, sc in enumerate(scs):
        o = run_oracle(sc, 8, "cauchy_loss", 0.01, 0.9)
        compare("ragged[%d] N=%d" % (i, Ns[i]), out["results"][i], out["history"], o, b=i, use_ratio=True)
    # the same queries one by one are bit-identical to the batch (deterministic per query)
    for i in (0, 3, 9):
        single = fm().refine_batch([probs[i]], 8, loss="cauchy_loss", lambda_=0.01, ratio_threshold=0.9)["results"][0]
        assert same_records(single, out["results"][i])


def test_more_queries_than_resident_ctas():
    """B > 2 * #SM: the persistent grid loops over the work counter."""
    sc = synth().make_scene(77, N=40, C=8, grad_scale=1.0 / 64, **SMALL)
    p = to_problem(sc)
    B = 700
    out = fm().refine_batch([p] * B, 5, loss="squared_loss", lambda_=0.01)
    r = out["results"]
    assert len(r) == B
    assert same_records(r, np.repeat(r[:1], B))
    o = run_oracle(sc, 5, "squared_loss", 0.01, None)
    assert np.abs(np.array(r[0]["R"]).reshape(3, 3) - o["R"]).max() < 1e-8


# ----------------------------------------------------------------------------------------------
# thread-block clusters: 1..16 CTAs refine one query together (small batches / single queries)
# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("cfg", [dict(N=1000, C=256, loss="geman_mcclure_loss", ratio=None, levels=None),
                                 dict(N=777, C=128, loss="squared_loss", ratio=0.8, levels=None),
                                 dict(N=3000, C=1024, loss="cauchy_loss", ratio=0.9,
                                      levels=[(512, 1024), (128, 512), (0, 128)]),
                                 dict(N=45, C=20, loss="huber_loss", ratio=None, levels=[(8, 20), (0, 8)])])
def test_cluster_sizes_match_oracle_and_each_other(cfg):
    """The result does not depend on how many CTAs share a query, or on whether they are a cluster or
    workers of the task scheduler: the reduction order is fixed (128 point ranges, pairwise tree), so
    records are bit-identical for cluster sizes 1..16 and for the scheduler with 1..128 tasks per evaluation."""
    sc = synth().make_scene(61, N=cfg["N"], C=cfg["C"], Hf=16, Wf=16, im_width=128, im_height=128,
                            K=((50.0, 0.0, 63.5), (0.0, 50.0, 62.3), (0.0, 0.0, 1.0)), grad_scale=1.0 / 64,
                            outlier_frac=0.1)
    p = to_problem(sc)
    o = run_oracle(sc, 6, cfg["loss"], 0.01, cfg["ratio"], cfg["levels"])
    first = None
    for S in (1, 2, 4, 8, 16, 0):
        out = fm().refine_batch([p], 6, loss=cfg["loss"], lambda_=0.01, ratio_threshold=cfg["ratio"],
                                levels=cfg["levels"], track=True, cluster_size=S)
        g = fm().default_context().last_geometry
        if S:
            assert g["cluster_size"] == S and g["grid"] == S and g["tasks_per_evaluation"] == 0, g
        else:  # automatic: the task scheduler, as many tasks as keep >= 32 points each
            assert g["cluster_size"] == 1 and g["tasks_per_evaluation"] >= 1 and g["grid"] == g["tasks_per_evaluation"], g
        compare("cluster=%d" % S, out["results"][0], out["history"], o, use_ratio=cfg["ratio"] is not None)
        if first is None:
            first = out
        else:
            assert same_records(out["results"], first["results"]), "cluster=%d differs from cluster=1" % S
            for k in ("costs", "Rs", "ts", "points2d", "mask", "kept"):
                assert np.array_equal(out["history"][k], first["history"][k], equal_nan=k in ("costs", "Rs", "ts")), k
    # the task scheduler with 1..16 tasks per evaluation (FMPNP_UNITS is read when a context is created)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    for U in (1, 2, 4, 8, 16, 32, 64, 128):
        os.environ["FMPNP_UNITS"] = str(U)
        try:
            ctx = fm().Context()
        finally:
            del os.environ["FMPNP_UNITS"]
        out = fm().refine_batch([p, p, p], 6, loss=cfg["loss"], lambda_=0.01, ratio_threshold=cfg["ratio"],
                                levels=cfg["levels"], track=True, ctx=ctx)
        g = ctx.last_geometry
        assert g["cluster_size"] == 1 and g["tasks_per_evaluation"] == U and g["grid"] == min(3 * U, 2 * sms), g
        ctx.close()
        for i in range(3):
            assert same_records(out["results"][i], first["results"][0]), "units=%d differs from cluster=1" % U
        for k in ("costs", "Rs", "ts", "points2d", "mask", "kept"):
            assert np.array_equal(out["history"][k][0], first["history"][k][0], equal_nan=k in ("costs", "Rs", "ts")), k
            assert np.array_equal(out["history"][k][2], first["history"][k][0], equal_nan=k in ("costs", "Rs", "ts")), k


def test_automatic_geometry_and_batch_size_independence():
    """Auto geometry (task scheduler for every batch size): many tasks per evaluation for small batches, one for large ones;
    the records of a query are bit-identical whatever batch it is part of."""
    scs = [synth().make_scene(300 + i, N=500, C=1024, Hf=16, Wf=16, im_width=128, im_height=128,
                              K=((50.0, 0.0, 63.5), (0.0, 50.0, 62.3), (0.0, 0.0, 1.0)), grad_scale=1.0 / 64,
                              outlier_frac=0.1) for i in range(3)]
    probs = [to_problem(sc) for sc in scs]
    ctx = fm().default_context()
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    ref = None
    for B in (1, 3, 20, 70, sms, sms + 1, 2 * sms + 5, 6 * sms + 1):
        batch = [probs[i % 3] for i in range(B)]
        r = fm().refine_batch(batch, 5, loss="geman_mcclure_loss", lambda_=0.01)["results"]
        g = ctx.last_geometry
        U = g["tasks_per_evaluation"]
        assert g["cluster_size"] == 1 and U in (1, 2, 4, 8) and g["grid"] <= min(2 * sms, B * U), g
        # about 1.5 tasks per CTA slot with record reuse (light evaluations pay per task), 3 when every row is gathered
        assert U >= 4 or B * U >= 3 * sms, g
        full = fm().refine_batch(batch, 5, loss="geman_mcclure_loss", lambda_=0.01, gather_every_row=True)["results"]
        U2 = ctx.last_geometry["tasks_per_evaluation"]
        assert U2 >= U and (U2 >= 4 or B * U2 >= 6 * sms), (U, U2)  # a task keeps >= 64 of the 500 points
        assert same_records(r, full, skip=("gathered_rows", "reused_rows"))
        if B >= 6 * sms:
            assert U == 1
        if B == 1:
            ref = r
        for i in range(B):
            assert same_records(r[i], ref[0]) if i % 3 == 0 else True
        assert same_records(r[:3], r[3:6]) if B >= 6 else True
    inl = fm().find_inliers_batch(probs, 0.8, loss="cauchy_loss").cpu().numpy()
    for i, sc in enumerate(scs):
        inp = sc.reference_inputs()
        exp = oracle.find_inliers(inp["pts3D"].numpy(), inp["R_init"].numpy(), inp["t_init"].numpy(),
                                  inp["feature_map_query"].numpy(), inp["feature_ref"].numpy(), inp["K"].numpy(),
                                  sc.im_width, sc.im_height, 0.8, loss="cauchy_loss")
        assert np.array_equal(inl[i].astype(bool), exp)


# ----------------------------------------------------------------------------------------------
# exit paths
# ----------------------------------------------------------------------------------------------
def test_no_supported_point_returns_initial_pose(route):
    sc = synth().make_scene(50, N=30, C=8, **SMALL)
    t_far = sc.t0.clone()
    t_far[0] += 1000.0  # everything projects outside the image
    out = fm().refine_batch([to_problem(sc, t0=t_far)], 5, track=True)
    r = out["results"][0]
    assert int(r["status"]) == fm().STATUS_EMPTY_SUPPORT and int(out["history"]["len"][0]) == 0
    assert np.array_equal(np.array(r["R"]).reshape(3, 3), sc.R0.numpy()) and np.array_equal(np.array(r["t"]), t_far.numpy())
    assert int(r["best_num_inliers"]) == -1 and np.isnan(r["best_cost"])
    o = run_oracle(sc, 5, "squared_loss", 0.01, None, t0=t_far.numpy())
    assert o["status"] == oracle.EMPTY_SUPPORT and o["has_best"] == 0
    # multilevel: compute_cost is None -> initial pose, NO_INITIAL_COST
    out = fm().refine_batch([to_problem(sc, t0=t_far)], 5, levels=[(4, 8), (0, 4)])
    assert int(out["results"][0]["status"]) == fm().STATUS_NO_INITIAL_COST
    assert np.array_equal(np.array(out["results"][0]["t"]), t_far.numpy())
    c, ns = fm().compute_cost_batch([to_problem(sc, t0=t_far)])
    assert int(ns[0]) == 0 and torch.isnan(c[0])


def test_zero_points(route):
    sc = synth().make_scene(51, N=4, C=8, **SMALL)
    sc.pts3D = sc.pts3D[:0].contiguous()
    sc.ref = sc.ref[:0].contiguous()
    out = fm().refine_batch([to_problem(sc)], 3, track=True)
    assert int(out["results"][0]["status"]) == fm().STATUS_EMPTY_SUPPORT
    assert fm().find_inliers_batch([to_problem(sc)], 0.8).shape[0] == 1


def test_nan_step_breaks_and_returns_best(route):
    """Gauss-Newton (lambda_ = 0) with identically zero gradient maps: H = 0, g = 0, LU meets a
    zero pivot, the step is 0/0 = NaN, the loop is left and the best (= initial) pose returned
    (the intent of model.py:411-413)."""
    sc = synth().make_scene(52, N=20, C=4, grad_scale=0.0, **SMALL)
    assert float(sc.Gx.abs().max()) == 0.0
    o = run_oracle(sc, 5, "squared_loss", 0.0, None)
    assert o["status"] == oracle.NAN_STEP and len(o["track"]["costs"]) == 1
    for solver in ("lu", "cholesky"):
        out = fm().refine_batch([to_problem(sc)], 5, loss="squared_loss", lambda_=0.0, track=True, solver=solver)
        compare("nan-step/" + solver, out["results"][0], out["history"], o)
        assert int(out["results"][0]["status"]) == fm().STATUS_NAN_STEP
        assert np.array_equal(np.array(out["results"][0]["R"]).reshape(3, 3), sc.R0.numpy())
    # with damping the same problem is solvable (step = 0): runs to the end, pose unchanged
    out = fm().refine_batch([to_problem(sc)], 5, loss="squared_loss", lambda_=0.01, track=True)
    o2 = run_oracle(sc, 5, "squared_loss", 0.01, None)
    compare("zero-gradient damped", out["results"][0], out["history"], o2)


def test_ratio_test_that_removes_every_point(route):
    """One supported point: |rho| < thr * max|rho| is false for it -> zero survivors, cost = nan,
    H = g = 0; the reference then keeps 'accepting' nan costs. Must match the oracle exactly."""
    sc = synth().make_scene(53, N=1, C=8, **SMALL)
    o = run_oracle(sc, 4, "squared_loss", 0.01, 0.8)
    out = fm().refine_batch([to_problem(sc)], 4, loss="squared_loss", lambda_=0.01, ratio_threshold=0.8, track=True)
    compare("ratio-all-removed", out["results"][0], out["history"], o, use_ratio=True)
    assert np.isnan(out["history"]["costs"][0, 0]) and int(out["results"][0]["best_num_inliers"]) == 0


def test_step_that_leaves_the_image_returns_current_pose(route):
    """Tiny gradients -> tiny H -> huge undamped step -> no supported point at the trial pose:
    forward returns the CURRENT pose (model.py:441-445)."""
    sc = synth().make_scene(54, N=60, C=8, grad_scale=1e-9, **SMALL)
    o = run_oracle(sc, 6, "squared_loss", 0.0, None)
    out = fm().refine_batch([to_problem(sc)], 6, loss="squared_loss", lambda_=0.0, track=True, solver="lu")
    compare("fly-away", out["results"][0], out["history"], o)


def test_points_behind_the_camera_are_not_filtered(route):
    sc = synth().make_scene(55, N=80, C=8, grad_scale=1.0 / 64, **SMALL)
    pts = sc.pts3D.clone()
    pts[::3] *= -1.0  # mirrored through the camera centre: same pixel, negative depth
    sc.pts3D = pts
    o = run_oracle(sc, 5, "squared_loss", 0.01, None)
    out = fm().refine_batch([to_problem(sc)], 5, loss="squared_loss", lambda_=0.01, track=True)
    compare("behind-camera", out["results"][0], out["history"], o)
    assert o["track"]["mask"][0][::3].any()


def test_active_mask_equals_compaction(route):
    """fmpnp_problem.active == passing pts3D[inliers], feature_ref[inliers] (optimize_feature_pnp.py:43-44)."""
    sc = synth().make_scene(56, N=150, C=16, grad_scale=1.0 / 64, outlier_frac=0.2, **SMALL)
    inl = fm().find_inliers_batch([to_problem(sc)], 0.8, loss="squared_loss")[0].cpu().numpy().astype(bool)
    inp = sc.reference_inputs()
    inl_o = oracle.find_inliers(inp["pts3D"].numpy(), sc.R0.numpy(), sc.t0.numpy(), inp["feature_map_query"].numpy(),
                                inp["feature_ref"].numpy(), sc.K.numpy(), sc.im_width, sc.im_height, 0.8)
    assert np.array_equal(inl, inl_o) and 0 < inl.sum() < 150
    sel = np.nonzero(inl)[0]
    o = run_oracle(sc, 8, "squared_loss", 0.0, 0.8, sel=sel)
    out = fm().refine_batch([to_problem(sc, active=inl.astype(np.uint8))], 8, loss="squared_loss", lambda_=0.0,
                            ratio_threshold=0.8, track=True)
    compare("active-mask", out["results"][0], out["history"], o, use_ratio=True, sel=sel)


@pytest.mark.parametrize("cfg", [dict(C=16, levels=None, ratio=0.8, loss="squared_loss", lam=0.0),
                                 dict(C=64, levels=None, ratio=None, loss="geman_mcclure_loss", lam=0.01),
                                 dict(C=64, levels=[(32, 64), (8, 32), (0, 8)], ratio=0.9, loss="cauchy_loss", lam=0.01),
                                 dict(C=7, levels=None, ratio=None, loss="huber_loss", lam=0.01)])
def test_fused_inlier_pass_equals_two_calls(cfg, route):
    """fmpnp_options.inlier_threshold == fmpnp_find_inliers_batch + fmpnp_refine_batch(active = its mask): the body of
    feature_pnp_multi's loop (optimize_feature_pnp.py:42-46) in one launch; every result field and history entry
    has the same bits (only the row counters differ: the fused pass gathers once), and the oracle agrees."""
    Ns = [150, 37, 300, 64, 211, 1]
    scs = [synth().make_scene(560 + i, N=n, C=cfg["C"], grad_scale=1.0 / 64, outlier_frac=0.2, **SMALL) for i, n in enumerate(Ns)]
    kw = dict(loss=cfg["loss"], lambda_=cfg["lam"], ratio_threshold=cfg["ratio"], levels=cfg["levels"], track=True)
    inl = fm().find_inliers_batch([to_problem(sc) for sc in scs], 0.8, loss=cfg["loss"]).cpu().numpy()
    two = fm().refine_batch([to_problem(sc, active=inl[i, :Ns[i]]) for i, sc in enumerate(scs)], 6, **kw)
    for sub_ in (slice(0, 6), slice(0, 1)):  # a batch (task scheduler) and a single query
        probs = [to_problem(sc) for sc in scs[sub_]]
        one = fm().refine_batch(probs, 6, inlier_threshold=0.8, **kw)
        B = len(probs)
        assert same_records(one["results"], two["results"][sub_], skip=("gathered_rows", "reused_rows"))
        for b in range(B):
            n = int(one["history"]["len"][b])
            assert n == int(two["history"]["len"][b])
            for k in ("Rs", "ts", "costs", "points2d", "mask", "kept"):
                x, y = one["history"][k][b, :n], two["history"][k][b, :n]
                if k in ("points2d", "mask", "kept"):  # per-point arrays are padded to the longest query of their batch
                    x, y = x[:, :Ns[b]], y[:, :Ns[b]]
                assert np.array_equal(x, y, equal_nan=k in ("Rs", "ts", "costs")), k
        if cfg["levels"] is None:  # the pass's records serve the first evaluation: fewer rows than the two calls' refinement
            assert (one["results"]["gathered_rows"] + one["results"]["reused_rows"] >=
                    two["results"]["gathered_rows"][sub_] + two["results"]["reused_rows"][sub_]).all()
        every = fm().refine_batch(probs, 6, inlier_threshold=0.8, gather_every_row=True, **kw)
        assert same_records(one["results"], every["results"], skip=("gathered_rows", "reused_rows"))
        clus = fm().refine_batch(probs[:1], 6, inlier_threshold=0.8, cluster_size=4, **kw)
        assert same_records(clus["results"][0], one["results"][0])
    sc = scs[0]
    sel = np.nonzero(inl[0, :Ns[0]])[0]
    assert 0 < len(sel) < Ns[0]
    o = run_oracle(sc, 6, cfg["loss"], cfg["lam"], cfg["ratio"], cfg["levels"], sel=sel)
    compare("fused-inliers", one["results"][0], one["history"], o, use_ratio=cfg["ratio"] is not None, sel=sel)


def test_maps_in_pinned_host_memory_give_the_same_records():
    """fmpnp_problem.F / Gx / Gy / ref may live in pinned (mapped) host memory: the kernels read the rows they gather
    over PCIe.  Same bits as with device-resident maps."""
    scs = [synth().make_scene(580 + i, N=n, C=64, grad_scale=1.0 / 64, outlier_frac=0.1, **SMALL) for i, n in enumerate([200, 90, 33, 150, 64, 7])]
    kw = dict(loss="geman_mcclure_loss", lambda_=0.01, levels=[(32, 64), (0, 32)], track=True)
    dev = fm().refine_batch([to_problem(sc) for sc in scs], 5, **kw)
    pin = lambda t: t.cpu().pin_memory()
    probs = [fm().Problem(pin(sc.F), pin(sc.Gx), pin(sc.Gy), pin(sc.ref), sc.pts3D.cuda(), sc.K, sc.R0, sc.t0,
                          sc.im_width, sc.im_height) for sc in scs]
    zc = fm().refine_batch(probs, 5, **kw)
    assert same_records(zc["results"], dev["results"])
    assert np.array_equal(zc["history"]["costs"], dev["history"]["costs"], equal_nan=True)
    inl_d = fm().find_inliers_batch([to_problem(sc) for sc in scs], 0.8)
    inl_z = fm().find_inliers_batch(probs, 0.8)
    assert torch.equal(inl_d, inl_z)
    with pytest.raises(fm().FmpnpError):  # pageable host memory is not reachable from the device
        fm().Problem(scs[0].F.cpu(), pin(scs[0].Gx), pin(scs[0].Gy), pin(scs[0].ref), scs[0].pts3D.cuda(), scs[0].K,
                     scs[0].R0, scs[0].t0, scs[0].im_width, scs[0].im_height)


def test_unaligned_views_take_the_generic_path(route):
    """Descriptor / map tensors whose base address is not 16-byte aligned (a slice of a larger buffer)."""
    sc = synth().make_scene(57, N=64, C=16, grad_scale=1.0 / 64, **SMALL)
    dev = "cuda"

    def misalign(t):
        buf = torch.empty(t.numel() + 1, dtype=t.dtype, device=dev)
        v = buf[1:].view(t.shape)
        v.copy_(t)
        assert v.data_ptr() % 16 != 0
        return v

    p = fm().Problem(misalign(sc.F), misalign(sc.Gx), misalign(sc.Gy), misalign(sc.ref), sc.pts3D.to(dev), sc.K, sc.R0,
                     sc.t0, sc.im_width, sc.im_height)
    out = fm().refine_batch([p], 6, loss="geman_mcclure_loss", lambda_=0.01, track=True)
    o = run_oracle(sc, 6, "geman_mcclure_loss", 0.01, None)
    compare("unaligned", out["results"][0], out["history"], o)


def test_argument_errors_are_reported_not_crashed():
    sc = synth().make_scene(58, N=16, C=8, **SMALL)
    p = to_problem(sc)
    with pytest.raises(fm().FmpnpError):
        fm().refine_batch([p], 5, levels=[(0, 9)])  # slice outside the 8 channels
    with pytest.raises(fm().FmpnpError):
        fm().refine_batch([p], -1)
    with pytest.raises(fm().FmpnpError):
        fm().Problem(sc.F, sc.Gx, sc.Gy, sc.ref, sc.pts3D, sc.K, sc.R0, sc.t0, 256, 256)  # CPU tensors


# ----------------------------------------------------------------------------------------------
# BASELINE sizes: N = 4096, C = 1024, 128 x 128 maps, 1024 x 1024 image
# ----------------------------------------------------------------------------------------------
def torch_eval(sc, R, t, dev="cuda"):
    """Plain torch float64 restatement of ONE linearisation (squared loss, no ratio test) on the
    GPU: returns cost mean, H (6x6), g (6), pixel indices and support mask."""
    pts = sc.pts3D.to(dev)
    R, t, K = (torch.as_tensor(x, dtype=torch.float64, device=dev) for x in (R, t, sc.K))
    P = (R @ pts.T).T + t
    q = (K @ P.T).T
    px = torch.round(q[:, :2] / q[:, 2:3]).to(torch.int32) - 1
    m = (px[:, 0] >= 0) & (px[:, 1] >= 0) & (px[:, 0] < sc.im_width) & (px[:, 1] < sc.im_height)
    Hf, Wf, C = sc.F.shape
    row = torch.div(px[m, 1].long() * Hf, sc.im_height, rounding_mode="floor")
    col = torch.div(px[m, 0].long() * Wf, sc.im_width, rounding_mode="floor")
    cell = row * Wf + col
    F, Gx, Gy = (x.to(dev).reshape(Hf * Wf, C) for x in (sc.F, sc.Gx, sc.Gy))
    e = F[cell].double() - sc.ref.to(dev)[m].double()
    gx, gy = Gx[cell].double(), Gy[cell].double()
    Ps = P[m]
    X, Y, Z = Ps[:, 0], Ps[:, 1], Ps[:, 2]
    fx, fy = K[0, 0], K[1, 1]
    zero = torch.zeros_like(X)
    J1 = torch.stack([fx / Z, zero, -fx * X / Z / Z, zero, fy / Z, -fy * Y / Z / Z], -1).reshape(-1, 2, 3)
    sk = torch.stack([zero, -Z, Y, Z, zero, -X, -Y, X, zero], -1).reshape(-1, 3, 3)
    JpT = torch.cat([torch.eye(3, dtype=torch.float64, device=dev).expand(len(X), 3, 3), -sk], -1)
    A = J1 @ JpT  # n,2,6
    S = torch.stack([(gx * gx).sum(-1), (gx * gy).sum(-1), (gx * gy).sum(-1), (gy * gy).sum(-1)], -1).reshape(-1, 2, 2)
    b = torch.stack([(gx * e).sum(-1), (gy * e).sum(-1)], -1)
    H = (A.transpose(1, 2) @ S @ A).sum(0)
    g = (A.transpose(1, 2) @ b[..., None]).sum(0)[:, 0]
    cost = (0.5 * (e * e).sum(-1)).mean()
    return cost.item(), H.cpu().numpy(), g.cpu().numpy(), px.cpu().numpy(), m.cpu().numpy()


@pytest.fixture(scope="module")
def big_scenes():
    s = synth()
    return [s.make_scene(900 + q, N=4096, C=1024, Hf=128, Wf=128, device="cuda", grad_scale=1.0 / 64) for q in range(3)]


def test_full_size_first_step_against_torch_fp64(big_scenes):
    """At full size the oracle would take minutes; one LM step is checked against a vectorised
    torch float64 evaluation instead (cost, pixel indices, mask, and the pose after one step)."""
    for sc in big_scenes:
        p = fm().Problem.from_scene(sc)
        out = fm().refine_batch([p], 1, loss="squared_loss", lambda_=0.01, track=True)
        h, r = out["history"], out["results"][0]
        cost, H, g, px, m = torch_eval(sc, sc.R0, sc.t0)
        assert np.array_equal(h["points2d"][0, 0], px) and np.array_equal(h["mask"][0, 0].astype(bool), m)
        assert abs(h["costs"][0, 0] - cost) <= 1e-11 * abs(cost)
        Hd = H + 0.01 * np.diag(np.diag(H) + 1e-9)
        delta = -np.linalg.solve(Hd, g)
        dr = oracle.so3exp(delta[3:])
        Rn, tn = dr @ sc.R0.numpy(), dr @ sc.t0.numpy() + delta[:3]
        np.testing.assert_allclose(h["Rs"][0, 1], Rn, rtol=0, atol=1e-10)
        np.testing.assert_allclose(h["ts"][0, 1], tn, rtol=0, atol=1e-9)
        cost1, _, _, px1, m1 = torch_eval(sc, Rn, tn)
        assert np.array_equal(h["points2d"][0, 1], px1) and np.array_equal(h["mask"][0, 1].astype(bool), m1)
        assert abs(h["costs"][0, 1] - cost1) <= 1e-11 * abs(cost1)
        assert int(r["gathered_rows"]) + int(r["reused_rows"]) == (int(m.sum()) + int(m1.sum())) * 1024 * 4
        cell = lambda px_: (px_[:, 1] * 128 // 1024) * 128 + px_[:, 0] * 128 // 1024
        stay = m & m1 & (cell(px) == cell(px1))  # supported twice, same cell: the record is kept
        assert int(r["reused_rows"]) == int(stay.sum()) * 1024 * 4
        full = fm().refine_batch([p], 1, loss="squared_loss", lambda_=0.01, track=True, gather_every_row=True)
        assert_reuse_is_invisible(out, full, 1)


def test_full_size_fixed_point_and_determinism(big_scenes):
    s = synth()
    sc = s.make_scene(950, N=4096, C=1024, Hf=128, Wf=128, device="cuda", noise_sigma=0.0, grad_scale=1.0 / 64)
    eye, zero = torch.eye(3, dtype=torch.float64), torch.zeros(3, dtype=torch.float64)
    p = fm().Problem.from_scene(sc, R0=eye, t0=zero)
    out = fm().refine_batch([p], 20, loss="geman_mcclure_loss", lambda_=0.01, track=True)
    r, h = out["results"][0], out["history"]
    # descriptors equal the map at the ground-truth cells and the pose is the ground truth:
    # zero residual -> zero gradient -> zero step: cost stays 0 and the pose does not move at all
    assert int(h["len"][0]) == 21 and np.all(h["costs"][0, :21] == 0.0)
    assert np.array_equal(np.array(r["R"]).reshape(3, 3), np.eye(3)) and np.array_equal(np.array(r["t"]), np.zeros(3))
    # determinism + batch independence at full size, LM with rejections possible
    probs = [fm().Problem.from_scene(x) for x in big_scenes]
    a = fm().refine_batch(probs, 20, loss="geman_mcclure_loss", lambda_=0.01)["results"]
    b = fm().refine_batch(probs[::-1], 20, loss="geman_mcclure_loss", lambda_=0.01)["results"][::-1]
    assert same_records(a, b)
    assert np.all(a["status"] == 0) and np.all(a["n_evaluations"] == 21)
    assert np.all(a["best_cost"] <= a["initial_cost"])
    # pyramid at full size: every level's evaluations are counted, compute_cost included
    lv = [(512, 1024), (128, 512), (0, 128)]
    c = fm().refine_batch(probs, 20, loss="squared_loss", lambda_=0.01, levels=lv)["results"]
    assert np.all(c["n_evaluations"] == 1 + 3 * 21)
    costs, nsup = fm().compute_cost_batch(probs)
    assert np.allclose(c["initial_cost"], costs.cpu().numpy(), rtol=0, atol=0)


def test_two_contexts_on_two_streams_at_once():
    """Two host threads, each with its own context and CUDA stream, refine different batches at the same time:
    the CTAs of the second launch only become resident as those of the first retire, which the scheduler must
    tolerate (queries are started from a counter by whichever CTA is running).  Records equal the solo runs."""
    import threading
    scs = [synth().make_scene(900 + i, N=700, C=256, Hf=16, Wf=16, im_width=128, im_height=128,
                              K=((50.0, 0.0, 63.5), (0.0, 50.0, 62.3), (0.0, 0.0, 1.0)), grad_scale=1.0 / 64,
                              outlier_frac=0.1) for i in range(6)]
    probs = [to_problem(sc) for sc in scs]
    batches = [[probs[i % 6] for i in range(400)], [probs[(i + 3) % 6] for i in range(333)]]
    kw = dict(loss="geman_mcclure_loss", lambda_=0.01, ratio_threshold=0.8)
    solo = [fm().refine_batch(b, 8, **kw)["results"] for b in batches]
    out, errs = [None, None], []

    def worker(j):
        try:
            ctx = fm().Context()
            with torch.cuda.stream(torch.cuda.Stream()):
                for _ in range(3):
                    out[j] = fm().refine_batch(batches[j], 8, ctx=ctx, **kw)["results"]
            ctx.close()
        except Exception as e:  # pragma: no cover
            errs.append(e)

    ts = [threading.Thread(target=worker, args=(j,)) for j in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=100)
    assert not errs, errs
    for j in range(2):
        assert out[j] is not None and same_records(out[j], solo[j]), "batch %d differs when run concurrently" % j


@pytest.mark.parametrize("cfg", ["gn_ratio_inlier", "lm_geman", "pyramid_ratio"])
def test_light_route_equals_task_route_with_many_misses(cfg):
    """The light-evaluation route (csrc/fmpnp_light.cuh: one CTA runs a whole evaluation, projection and per-point
    phase per warp, misses streamed from a per-warp list) against the task route, on queries whose points DO change
    cell (per-pixel gradients, 1-3 degree errors: more than 32 misses per warp and evaluation, so the list overflows
    and is flushed in mid-loop), ragged point counts, with history: every record, counter and history entry identical."""
    from conftest import _scheduler_context
    s = synth()
    Ks = ((400.0, 0.0, 255.5), (0.0, 400.0, 254.3), (0.0, 0.0, 1.0))
    scs = [s.make_scene(4200 + i, N=n, C=128, Hf=64, Wf=64, im_width=512, im_height=512, K=Ks, grad_scale=1.0 / 64,
                        rot_deg=(1.0, 3.0), outlier_frac=0.1) for i, n in enumerate((2048, 1500, 4096, 777))]
    probs = [to_problem(sc) for sc in scs]
    kw = {"gn_ratio_inlier": dict(loss="squared_loss", lambda_=0.0, ratio_threshold=0.8, inlier_threshold=0.8),
          "lm_geman": dict(loss="geman_mcclure_loss", lambda_=0.01),
          "pyramid_ratio": dict(loss="huber_loss", lambda_=0.01, ratio_threshold=0.9, levels=[(64, 128), (16, 64), (0, 16)])}[cfg]
    track = cfg != "gn_ratio_inlier"
    a = fm().refine_batch(probs, 12, track=track, **kw)
    b = fm().refine_batch(probs, 12, track=track, ctx=_scheduler_context("light"), **kw)
    for f in a["results"].dtype.names:
        assert a["results"][f].tobytes() == b["results"][f].tobytes(), "field %s differs on the light route" % f
    # the same route of the one-CTA-per-query kernel (cluster_size = 1: fmpnp::refine_kernel runs light_evaluate too)
    b1 = fm().refine_batch(probs, 12, track=track, ctx=_scheduler_context("light"), cluster_size=1, **kw)
    for f in a["results"].dtype.names:
        assert a["results"][f].tobytes() == b1["results"][f].tobytes(), "field %s differs on the per-query light route" % f
    assert int(a["results"]["gathered_rows"].min()) > 0 and float(a["results"]["reused_rows"].sum()) > 0
    if track:
        for k in ("len", "costs", "Rs", "ts", "points2d", "mask", "kept"):
            assert np.array_equal(a["history"][k], b["history"][k], equal_nan=True) if a["history"][k].dtype.kind == "f" \
                else np.array_equal(a["history"][k], b["history"][k]), "history %s differs on the light route" % k


def test_keep_records_starts_from_the_previous_launch():
    """fmpnp_options.keep_records: a second launch on the same arrays (other poses) reuses the records the first one
    left -- same results as without, next to nothing gathered; any other call on the context in between, or other
    arrays, makes it a normal call."""
    s = synth()
    Ks = ((400.0, 0.0, 255.5), (0.0, 400.0, 254.3), (0.0, 0.0, 1.0))
    scs = [s.make_scene(5100 + i, N=n, C=256, Hf=64, Wf=64, im_width=512, im_height=512, K=Ks, outlier_frac=0.1)
           for i, n in enumerate((1500, 2048, 900))]
    first = [to_problem(sc) for sc in scs]
    kw = dict(loss="squared_loss", lambda_=0.0, ratio_threshold=0.8, inlier_threshold=0.8)
    ca, cb = fm().Context(), fm().Context()
    r1a = fm().refine_batch(first, 6, ctx=ca, **kw)["results"]
    r1b = fm().refine_batch(first, 6, ctx=cb, **kw)["results"]
    assert same_records(r1a, r1b)

    def again(res):  # round 2 of feature_pnp_multi: the same arrays at the pose round 1 returned
        return [fm().Problem(p.F, p.Gx, p.Gy, p.ref, p.pts3D, p.K, np.array(r["R"]), np.array(r["t"]), p.im_width, p.im_height)
                for p, r in zip(first, res)]

    second = again(r1a)
    r2a = fm().refine_batch(second, 6, ctx=ca, keep_records=True, **kw)["results"]
    r2b = fm().refine_batch(second, 6, ctx=cb, **kw)["results"]
    counters = ("gathered_rows", "reused_rows")
    assert same_records(r2a, r2b, skip=counters)
    assert np.array_equal(r2a["gathered_rows"] + r2a["reused_rows"], r2b["gathered_rows"] + r2b["reused_rows"])
    assert r2a["gathered_rows"].sum() < 0.1 * r2b["gathered_rows"].sum(), (r2a["gathered_rows"], r2b["gathered_rows"])
    # another call of the context in between: its records are gone, the option is ignored
    fm().compute_cost_batch(second, ctx=ca)
    r3a = fm().refine_batch(second, 6, ctx=ca, keep_records=True, **kw)["results"]
    assert same_records(r3a, r2b)
    # other arrays (copies): ignored as well
    r4a = fm().refine_batch(second, 6, ctx=ca, keep_records=True, **kw)["results"]  # (this one does reuse: same arrays again)
    assert same_records(r4a, r2b, skip=counters) and r4a["gathered_rows"].sum() < 0.1 * r2b["gathered_rows"].sum()
    copies = [fm().Problem(p.F.clone(), p.Gx.clone(), p.Gy.clone(), p.ref.clone(), p.pts3D.clone(), p.K, p.R0, p.t0,
                           p.im_width, p.im_height) for p in second]
    r5a = fm().refine_batch(copies, 6, ctx=ca, keep_records=True, **kw)["results"]
    assert same_records(r5a, r2b)
    ca.close()
    cb.close()
