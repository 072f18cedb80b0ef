"""GPU: the rows either side of the hot path -- device Sobel + pack, descriptor
gather, per-level resize / blur, and the device-resident adapter -- against torch float64 on the host and
against fixtures produced by the UNMODIFIED reference adapter (oracle/gen_golden_adapter.py)."""
import functools
import os
from collections import namedtuple

import numpy as np
import pytest
import torch

from util import COST_RTOL, GOLDEN_DIR, ROT_TOL_DEG, TRANS_TOL_M, assert_costs_close, rot_angle_deg

pytestmark = pytest.mark.gpu
TIGHT = 1e-9
Prediction = namedtuple("Prediction", ["points_3d", "reference_inliers", "matrix", "inlier_mask", "quaternion",
                                       "reference_filename"])


def fm():
    import fmpnp_b200
    return fmpnp_b200


def load(name):
    d = dict(np.load(os.path.join(GOLDEN_DIR, name + ".npz"), allow_pickle=False))
    for k in ("loss", "kind"):
        if k in d:
            d[k] = str(d[k])
    return d


# ----------------------------------------------------------------------------------------------
# kernels against torch float64
# ----------------------------------------------------------------------------------------------
def torch_sobel(x64, border, normalized):
    c, h, w = x64.shape
    kx = torch.tensor([[-1.0, 0.0, 1.0], [-2.0, 0.0, 2.0], [-1.0, 0.0, 1.0]], dtype=torch.float64)
    if normalized:
        kx = kx / 8.0
    k = torch.stack([kx, kx.t()])[:, None]
    xp = torch.nn.functional.pad(x64.reshape(c, 1, h, w), (1, 1, 1, 1), mode="replicate" if border == "replicate" else "constant")
    g = torch.nn.functional.conv2d(xp, k)
    return g[:, 0], g[:, 1]


@pytest.mark.parametrize("shape", [(32, 16, 16), (70, 37, 45), (1, 128, 128), (33, 8, 33)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("border,normalized", [("replicate", False), ("zero", False), ("replicate", True)])
def test_sobel_pack_matches_torch_float64(shape, dtype, border, normalized):
    g = torch.Generator().manual_seed(sum(shape))
    x = torch.rand(shape, generator=g, dtype=torch.float32).to(dtype)
    F, Gx, Gy = fm().sobel_pack(x.cuda(), border=border, normalized=normalized)
    gx, gy = torch_sobel(x.double(), border, normalized)
    hwc = lambda m: m.permute(1, 2, 0).contiguous().float()  # noqa: E731  (the storage rounding)
    assert torch.equal(F.cpu(), hwc(x.double()))
    assert torch.equal(Gx.cpu(), hwc(gx)) and torch.equal(Gy.cpu(), hwc(gy))
    # and the drop-in helpers.utils.sobel_filter + pack path gives the same maps
    import helpers.utils as hu
    g2x, g2y = hu.sobel_filter(x.double(), normalized=normalized) if border == "replicate" else (gx, gy)
    assert torch.equal(fm().pack_chw_to_hwc(g2x.contiguous().cuda()).cpu(), Gx.cpu())
    assert torch.equal(fm().pack_chw_to_hwc(g2y.contiguous().cuda()).cpu(), Gy.cpu())


def test_gather_descriptors():
    g = torch.Generator().manual_seed(5)
    src = torch.rand((37, 19, 23), generator=g)
    rows = torch.randint(-19, 19, (301,), generator=g)
    cols = torch.randint(-23, 23, (301,), generator=g)
    for s in (src, src.double()):
        out = fm().gather_descriptors(s.cuda(), rows, cols)
        exp = torch.stack([s[:, i, j] for i, j in zip(rows, cols)]).float()
        assert torch.equal(out.cpu(), exp)
    assert tuple(fm().gather_descriptors(src.cuda(), rows[:0], cols[:0]).shape) == (0, 37)
    with pytest.raises(IndexError):
        fm().gather_descriptors(src.cuda(), torch.tensor([19]), torch.tensor([0]))


@pytest.mark.parametrize("size,out", [((5, 32, 32), (16, 16)), ((3, 17, 29), (40, 11)), ((2, 8, 8), (8, 8)),
                                      ((4, 16, 16), (48, 48))])
def test_resize_matches_interpolate(size, out):
    x = torch.rand(size, generator=torch.Generator().manual_seed(9)).double()
    exp = torch.nn.functional.interpolate(x[None], size=out, mode="bilinear")[0]
    got = fm().resize_bilinear_chw(x.cuda(), *out).cpu()
    # torch's own CPU kernel is not specified to the last bit (its fma contraction depends on the build):
    # agreement to a few ulp of the [0, 1) values is the bar here
    assert torch.allclose(got, exp, rtol=0, atol=2e-15)
    got32 = fm().resize_bilinear_chw(x.float().cuda(), *out).cpu()
    exp32 = torch.nn.functional.interpolate(x.float().double()[None], size=out, mode="bilinear")[0]
    assert torch.allclose(got32, exp32, rtol=0, atol=2e-15)


@pytest.mark.parametrize("k", [3, 5])
def test_blur_matches_conv2d(k):
    import helpers.utils as hu
    x = torch.rand((6, 20, 27), generator=torch.Generator().manual_seed(k)).double()
    kern = hu.gaussian_kernel2d(k, 1.0)
    assert kern.dtype == torch.float64 and abs(kern.sum().item() - 1.0) < 1e-6
    w = kern[None, None].repeat(6, 1, 1, 1)
    exp = torch.nn.functional.conv2d(x[None], w, groups=6, padding=1)[0]
    got = fm().blur_chw(x.cuda(), kern, pad=1).cpu()
    assert tuple(got.shape) == tuple(exp.shape) == (6, 20 + 3 - k, 27 + 3 - k)
    assert torch.allclose(got, exp, rtol=0, atol=1e-15)


# ----------------------------------------------------------------------------------------------
# the adapter against the reference adapter's fixtures
# ----------------------------------------------------------------------------------------------
def prediction_of(g):
    im = g["inlier_mask"]
    return Prediction(points_3d=g["points_3d"], reference_inliers=g["reference_inliers"], matrix=g["matrix"],
                      inlier_mask=None if (im.size == 1 and im[0] < 0) else im, quaternion=np.array([1.0, 0, 0, 0]),
                      reference_filename="ref.png")


def check_track(name, model, g, tol):
    tr = model.track_
    n = len(g["costs"])
    assert len(tr["costs"]) == n, "%s: %d entries, expected %d" % (name, len(tr["costs"]), n)
    assert_costs_close(tr["costs"], g["costs"], tol, name)
    npts = g["n_points"]
    assert [int(p.shape[0]) for p in tr["points2d"]] == [int(v) for v in npts]
    got_px = np.concatenate([p.numpy().reshape(-1, 2) for p in tr["points2d"]])
    got_mask = np.concatenate([m.numpy().astype(np.uint8) for m in tr["mask"]])
    assert np.array_equal(got_px, g["points2d_flat"]), name + ": projected pixel indices differ"
    assert np.array_equal(got_mask, g["mask_flat"]), name + ": support masks differ"
    np.testing.assert_allclose(np.stack([r.numpy() for r in tr["Rs"]]), g["Rs"], rtol=0, atol=1e-7)
    np.testing.assert_allclose(np.stack([t.numpy() for t in tr["ts"]]), g["ts"], rtol=0, atol=1e-7)
    assert model.best_num_inliers_ == int(g["best_num_inliers"])
    assert abs(model.best_cost_.item() - float(g["best_cost"])) <= tol * abs(float(g["best_cost"]))
    assert abs(model.initial_cost_.item() - float(g["initial_cost"])) <= tol * abs(float(g["initial_cost"]))


@pytest.mark.parametrize("name,tol", [("adapter_pnp_exact", TIGHT), ("adapter_pnp_pyramid", TIGHT),
                                      ("adapter_pnp_float", COST_RTOL)])
@pytest.mark.parametrize("device", ["cuda", "cpu"])
def test_feature_pnp_matches_reference_adapter(name, tol, device, monkeypatch):
    import helpers.utils as hu
    import model
    from pose_prediction import optimize_feature_pnp as adapter
    g = load(name)
    rt = float(g["ratio_threshold"])
    monkeypatch.setattr(adapter, "sparseFeaturePnP", functools.partial(
        model.sparseFeaturePnP, n_iters=int(g["n_iters"]), loss_fn=getattr(hu, g["loss"]), lambda_=float(g["lambda_"]),
        ratio_threshold=None if np.isnan(rt) else rt))
    pyr = [(int(s), int(e), None, None) for s, e in g["levels"]] or None
    qh, rh = torch.from_numpy(g["qh"]).to(device), torch.from_numpy(g["rh"]).to(device)
    R, t, m = adapter.feature_pnp(qh, rh, prediction_of(g), torch.from_numpy(g["K"]), [int(v) for v in g["image_shape"]],
                                  track=True, feature_pyramid=pyr)
    assert R.dtype == torch.float64 and R.device.type == "cpu"
    check_track(name, m, g, tol)
    assert rot_angle_deg(R.numpy(), g["out_R"]) <= ROT_TOL_DEG and np.abs(t.numpy() - g["out_t"]).max() <= TRANS_TOL_M
    T = np.eye(4)
    T[:3, :3], T[3, :3] = R.numpy(), t.numpy()
    from pose_prediction import matrix_utils
    assert np.abs(matrix_utils.matrix_quaternion(T) - g["quaternion"]).max() < 1e-7


def test_feature_pnp_with_normalised_sobel_matches_reference_adapter(monkeypatch):
    """The Sobel normalisation of the reference's kornia is unpinned (helpers.utils.sobel_normalization): the fixture
    was produced by the reference adapter on a kornia stand-in that scales the kernels by 1/8; the adapter follows
    when `sobel_normalization.normalized = True` is bound, and does NOT match with the un-normalised default."""
    import helpers.utils as hu
    import model
    from pose_prediction import optimize_feature_pnp as adapter
    g = load("adapter_pnp_norm")
    assert int(g["sobel_normalized"]) == 1
    monkeypatch.setattr(adapter, "sparseFeaturePnP", functools.partial(
        model.sparseFeaturePnP, n_iters=int(g["n_iters"]), loss_fn=getattr(hu, g["loss"]), lambda_=float(g["lambda_"])))
    args = (torch.from_numpy(g["qh"]).cuda(), torch.from_numpy(g["rh"]).cuda(), prediction_of(g), torch.from_numpy(g["K"]),
            [int(v) for v in g["image_shape"]])
    monkeypatch.setattr(adapter, "sobel_normalization", lambda: True)
    R, t, m = adapter.feature_pnp(*args, track=True)
    check_track("adapter_pnp_norm", m, g, TIGHT)
    assert np.abs(R.numpy() - g["out_R"]).max() < 1e-7 and np.abs(t.numpy() - g["out_t"]).max() < 1e-7
    monkeypatch.setattr(adapter, "sobel_normalization", lambda: False)
    R2, t2, m2 = adapter.feature_pnp(*args, track=True)
    assert np.abs(t2.numpy() - g["out_t"]).max() > 1e-6  # an 8x gradient scale is not a detail
    prob = adapter.build_problem(*args, sobel_normalized=True)
    prob_raw = adapter.build_problem(*args)
    assert torch.equal(prob.Gx * 8.0, prob_raw.Gx) and torch.equal(prob.F, prob_raw.F)


def test_feature_pnp_with_resampled_levels_matches_reference_adapter(monkeypatch):
    """feature_pnp with target_size / kernel_size pyramid levels (optimize_feature_pnp.py:63-69 -> model.py:193-210):
    rebuilt per level on the device from the CHW hypercolumn.  (The Gaussian window is kornia's: parity unpinned.)"""
    import helpers.utils as hu
    import model
    from pose_prediction import optimize_feature_pnp as adapter
    model.clear_pack_cache()
    g = load("adapter_pnp_resampled")
    pyr = [(int(s), int(e), None if ts < 0 else int(ts), None if ks < 0 else int(ks)) for s, e, ts, ks in g["pyramid4"]]
    assert any(ts is not None or ks is not None for _, _, ts, ks in pyr)
    monkeypatch.setattr(adapter, "sparseFeaturePnP", functools.partial(
        model.sparseFeaturePnP, n_iters=int(g["n_iters"]), loss_fn=getattr(hu, g["loss"]), lambda_=float(g["lambda_"])))
    R, t, m = adapter.feature_pnp(torch.from_numpy(g["qh"]).cuda(), torch.from_numpy(g["rh"]).cuda(), prediction_of(g),
                                  torch.from_numpy(g["K"]), [int(v) for v in g["image_shape"]], track=True,
                                  feature_pyramid=pyr)
    check_track("adapter_pnp_resampled", m, g, COST_RTOL)
    assert rot_angle_deg(R.numpy(), g["out_R"]) <= ROT_TOL_DEG and np.abs(t.numpy() - g["out_t"]).max() <= TRANS_TOL_M


@pytest.mark.parametrize("name", ["adapter_multi_mask", "adapter_multi_nomask"])
def test_feature_pnp_multi_matches_reference_adapter(name, monkeypatch):
    import helpers.utils as hu
    import model
    from pose_prediction import optimize_feature_pnp as adapter
    g = load(name)
    monkeypatch.setattr(adapter, "sparseFeaturePnP", functools.partial(
        model.sparseFeaturePnP, n_iters=int(g["n_iters"]), loss_fn=getattr(hu, g["loss"]), lambda_=float(g["lambda_"])))
    R, t, m = adapter.feature_pnp_multi(torch.from_numpy(g["qh"]).cuda(), torch.from_numpy(g["rh"]).cuda(), prediction_of(g),
                                        torch.from_numpy(g["K"]), [int(v) for v in g["image_shape"]], track=True,
                                        threshold=float(g["threshold"]))
    check_track(name, m, g, TIGHT)
    assert np.abs(R.numpy() - g["out_R"]).max() < 1e-7 and np.abs(t.numpy() - g["out_t"]).max() < 1e-7


@pytest.mark.parametrize("name", ["adapter_multi_mask", "adapter_multi_nomask"])
def test_feature_pnp_multi_untracked_runs_selection_and_refinement_in_one_launch(name, monkeypatch):
    """track=False: every {find_inliers, forward} pair of feature_pnp_multi is one launch
    (fmpnp_options.inlier_threshold); same pose as the reference adapter's fixture and as the tracked path."""
    import fmpnp_b200 as fm
    import helpers.utils as hu
    import model
    from pose_prediction import optimize_feature_pnp as adapter
    g = load(name)
    monkeypatch.setattr(adapter, "sparseFeaturePnP", functools.partial(
        model.sparseFeaturePnP, n_iters=int(g["n_iters"]), loss_fn=getattr(hu, g["loss"]), lambda_=float(g["lambda_"])))
    args = (torch.from_numpy(g["qh"]).cuda(), torch.from_numpy(g["rh"]).cuda(), prediction_of(g), torch.from_numpy(g["K"]),
            [int(v) for v in g["image_shape"]])
    n0 = fm.default_context().launches
    R, t, m = adapter.feature_pnp_multi(*args, track=False, threshold=float(g["threshold"]), loss_fn=getattr(hu, g["loss"]))
    launches = fm.default_context().launches - n0
    # build_problem: Sobel + pack, descriptor gather; then one launch per round (the first round of a prediction
    # that brings its own inlier mask has nothing to select)
    assert launches == 2 + 3, launches
    assert np.abs(R.numpy() - g["out_R"]).max() < 1e-7 and np.abs(t.numpy() - g["out_t"]).max() < 1e-7
    Rt, tt, mt = adapter.feature_pnp_multi(*args, track=True, threshold=float(g["threshold"]), loss_fn=getattr(hu, g["loss"]))
    assert torch.equal(R, Rt) and torch.equal(t, tt) and m.best_num_inliers_ == mt.best_num_inliers_
    assert float(m.best_cost_) == float(mt.best_cost_)


def test_optimize_feature_pnp_entry_point(monkeypatch):
    """The top-level entry of the adapter: reference hypercolumn from `net`, (list(t), list(quaternion), model) out."""
    import helpers.utils as hu
    import model
    from pose_prediction import optimize_feature_pnp as adapter
    g = load("adapter_pnp_exact")
    monkeypatch.setattr(adapter, "sparseFeaturePnP", functools.partial(
        model.sparseFeaturePnP, n_iters=int(g["n_iters"]), loss_fn=getattr(hu, g["loss"]), lambda_=float(g["lambda_"])))

    class Net:
        def compute_hypercolumn(self, files, to_cpu=False, resize=True):
            assert files == ["ref.png"] and to_cpu is False and resize is True
            return torch.from_numpy(g["rh"]).cuda(), None

    t, q, m = adapter.optimize_feature_pnp(torch.from_numpy(g["qh"]).cuda(), Net(), prediction_of(g), g["K"],
                                           [int(v) for v in g["image_shape"]])
    assert isinstance(t, list) and isinstance(q, list) and len(t) == 3 and len(q) == 4
    assert np.abs(np.array(t) - g["out_t"]).max() < 1e-7 and np.abs(np.array(q) - g["quaternion"]).max() < 1e-7
    assert m.best_num_inliers_ == int(g["best_num_inliers"])


# ----------------------------------------------------------------------------------------------
# pyramid levels with resize / blur (model.py:193-210) through the drop-in model.py
# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["pyramid_resampled_a", "pyramid_resampled_b"])
def test_resampled_pyramid_matches_reference(name):
    import helpers.utils as hu
    import model
    model.clear_pack_cache()
    g = load(name)
    pyr = [(int(s), int(e), None if ts < 0 else int(ts), None if ks < 0 else int(ks)) for s, e, ts, ks in g["pyramid"]]
    F = torch.from_numpy(g["F_chw"]).double()
    gx, gy = hu.sobel_filter(F)
    m = model.sparseFeaturePnP(n_iters=int(g["n_iters"]), loss_fn=getattr(hu, g["loss"]), lambda_=float(g["lambda_"]))
    R, t = m.multilevel_optimization(pyr, torch.from_numpy(g["pts3D"]), torch.from_numpy(g["ref"]).double(), F, gx, gy,
                                     torch.from_numpy(g["K"]), int(g["im_width"]), int(g["im_height"]),
                                     R_init=torch.from_numpy(g["R0"]), t_init=torch.from_numpy(g["t0"]), track=True)
    # resized / blurred maps are float64 in the reference and float32 in HBM: BASELINE bars, not the tight ones
    check_track(name, m, g, COST_RTOL)
    assert rot_angle_deg(R.numpy(), g["out_R"]) <= ROT_TOL_DEG and np.abs(t.numpy() - g["out_t"]).max() <= TRANS_TOL_M


# ----------------------------------------------------------------------------------------------
# --only_optimization replay: one batched launch over the cached predictions == per-query feature_pnp
# ----------------------------------------------------------------------------------------------
def test_replay_batch_equals_per_query_adapter(tmp_path, monkeypatch):
    import model
    from pose_prediction import optimize_feature_pnp as adapter
    from pose_prediction import replay
    names = ["adapter_pnp_exact", "adapter_pnp_float", "adapter_multi_mask"]
    gs = {n: load(n) for n in names}
    records, hcs = {}, {}
    for n, g in gs.items():
        N = g["points_3d"].shape[0]
        records["q_%s.jpg:r_%s.jpg" % (n, n)] = {
            "reference_filename": "r_%s.jpg" % n, "success": True, "query_2D": np.zeros((N, 2)),
            "reference_2D": g["reference_inliers"], "points_3D": g["points_3d"], "num_matches": N, "num_inliers": N,
            "inlier_mask": np.arange(N), "quaternion": np.array([1.0, 0.0, 0.0, 0.0]), "matrix": g["matrix"]}
        hcs["q_%s.jpg" % n], hcs["r_%s.jpg" % n] = torch.from_numpy(g["qh"]).cuda(), torch.from_numpy(g["rh"]).cuda()
    failed = dict(records["q_adapter_pnp_exact.jpg:r_adapter_pnp_exact.jpg"], success=False)
    records["q_failed.jpg"] = failed
    out = str(tmp_path) + os.sep
    replay.write_npz(out + "cached_matches/slice.npz", records)
    back = replay.read_npz(out)
    rows, poses = replay.refine_cached(back, lambda f: hcs[f], lambda q: gs[names[0]]["K"], [256, 256], n_iters=7,
                                       loss_fn="geman_mcclure_loss", lambda_=0.01)
    assert len(rows) == 4 and len(poses) == 8
    monkeypatch.setattr(adapter, "sparseFeaturePnP", functools.partial(
        model.sparseFeaturePnP, n_iters=7, loss_fn=__import__("helpers.utils").utils.geman_mcclure_loss, lambda_=0.01))
    for i, n in enumerate(names):
        g = gs[n]
        R, t, m = adapter.feature_pnp(hcs["q_%s.jpg" % n], hcs["r_%s.jpg" % n],
                                      replay.prediction_from_record(back["q_%s.jpg:r_%s.jpg" % (n, n)]),
                                      torch.from_numpy(g["K"]), [256, 256])
        row, pose = rows[i], poses[2 * i + 1]
        assert row[0] == "r_%s.jpg" % n and row[1] == "q_%s.jpg" % n and row[2] == g["points_3d"].shape[0]
        assert row[3] == m.best_num_inliers_ and row[4] == m.initial_cost_.item() and row[5] == m.best_cost_.item()
        assert np.array_equal(np.array(pose[5:]), t.numpy())  # same kernel, same reduction order: bit-identical
        T = np.eye(4)
        T[:3, :3] = R.numpy()
        from pose_prediction import matrix_utils
        assert np.array_equal(np.array(pose[1:5]), matrix_utils.matrix_quaternion(T))
    assert rows[3][2:] == [None] * 5 and poses[6] == poses[7]
    replay.write_summary(out + "featurePnP_summary.csv", rows)
    replay.write_poses(out + "poses.txt", poses)
    assert len(open(out + "featurePnP_summary.csv").read().splitlines()) == 5
    assert len(open(out + "poses.txt").read().splitlines()) == 4
