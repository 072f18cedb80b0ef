"""Golden fixtures for the rows either side of the hot path, produced by
running the UNMODIFIED reference adapter (`s2dhm/pose_prediction/optimize_feature_pnp.py`) and the
reference's resize / blur pyramid branch (`featurePnP/model.py:193-210`) on CPU in the build container:

    python oracle/gen_golden_adapter.py        # writes tests/golden/adapter_*.npz, pyramid_resampled_*.npz

Two things the reference needs that this container cannot give it are supplied here and nowhere else:
  * gin injection: `sparseFeaturePnP()` / `find_inliers(...)` are called without arguments in the adapter
    (their parameters come from .gin files); the generator binds them with functools.partial;
  * `find_inliers` ends with `mask[mask] = ...` (model.py:151), which torch >= 2 rejects as a
    self-overlapping write; the generator substitutes the same statements with the write going to a
    clone (identical to oracle/gen_golden.py's find_inliers_ref).
"""
import functools
import importlib
import os
import sys
import warnings
from collections import namedtuple

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "featuremetric-pnp_b200"))

import pyref  # noqa: E402
from fmpnp_b200 import synth  # noqa: E402

warnings.filterwarnings("ignore")
GOLDEN = os.path.join(ROOT, "tests", "golden")
Prediction = namedtuple("Prediction", ["success", "num_matches", "num_inliers", "reference_inliers", "query_inliers",
                                       "points_3d", "quaternion", "matrix", "reference_filename", "reference_keypoints",
                                       "inlier_mask"])
KS = ((100.0, 0.0, 127.055733), (0.0, 100.0, 124.546844), (0.0, 0.0, 1.0))


def load_reference_adapter():
    """Imports the reference's adapter module under a private name, with the reference's own model /
    helpers / matrix_utils and the gin / kornia stand-ins of oracle/refshim."""
    ref_model, ref_utils = pyref.load()
    saved_path = list(sys.path)
    saved = {k: sys.modules.get(k) for k in ("model", "helpers", "helpers.utils", "gin", "kornia", "kornia.filters",
                                             "pose_prediction", "pose_prediction.matrix_utils",
                                             "pose_prediction.optimize_feature_pnp")}
    try:
        for k in saved:
            sys.modules.pop(k, None)
        sys.modules["model"] = ref_model
        sys.modules["helpers.utils"] = ref_utils
        import types
        hp = types.ModuleType("helpers")
        hp.utils = ref_utils
        sys.modules["helpers"] = hp
        sys.path.insert(0, os.path.join(HERE, "refshim"))
        # s2dhm/pose_prediction has no __init__.py (namespace package): load its two files by path
        import importlib.util
        pdir = os.path.join(pyref.REFERENCE_ROOT, "s2dhm", "pose_prediction")
        pkg = types.ModuleType("pose_prediction")
        pkg.__path__ = [pdir]
        sys.modules["pose_prediction"] = pkg
        spec = importlib.util.spec_from_file_location("pose_prediction.matrix_utils", os.path.join(pdir, "matrix_utils.py"))
        mu = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mu)
        sys.modules["pose_prediction.matrix_utils"] = mu
        pkg.matrix_utils = mu
        spec = importlib.util.spec_from_file_location("_fmpnp_ref_adapter", os.path.join(pdir, "optimize_feature_pnp.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.path[:] = saved_path
        for k, v in saved.items():
            sys.modules.pop(k, None)
            if v is not None:
                sys.modules[k] = v
    return mod, ref_model, ref_utils


def fixed_find_inliers(ref_model, ref_utils, pts3D, R, t, feature_map_query, feature_ref, K, im_width, im_height,
                       threshold=None, loss_fn=None, mode="ratio_max"):
    """model.py:131-152 with its last write going to a clone (see the module docstring)."""
    loss_fn = loss_fn or ref_utils.squared_loss
    points_3d = torch.mm(R, pts3D.T).T + t
    points_2d = torch.round(ref_utils.from_homogeneous(torch.mm(K, points_3d.T).T)).type(torch.IntTensor) - 1
    mask_supported = ref_model.points_within_image(points_2d, im_width, im_height)
    p2s = points_2d[mask_supported, :]
    error = ref_model.indexing_(feature_map_query, torch.flip(p2s, (1,)), im_width, im_height) - feature_ref[mask_supported]
    cost = 0.5 * (error ** 2).sum(-1)
    cost_full, _, _ = loss_fn(cost)
    threshold_mask = ref_model.ratio_threshold_feature_errors(cost_full, threshold=threshold)
    out = mask_supported.clone()
    out[mask_supported] = threshold_mask
    return out


def make_adapter_case(seed, N=120, C=32, H=32, W=32, image=256, quantised=True, outlier_frac=0.1, with_mask=True):
    """Query hypercolumn (from a synthetic scene), a reference hypercolumn whose cells under the reference
    keypoints hold the descriptors of the scene, and a solve_pnp-style Prediction."""
    sc = synth.make_scene(seed, N=N, C=C, Hf=H, Wf=W, im_width=image, im_height=image, K=KS, device="cpu",
                          outlier_frac=outlier_frac, grad_scale=1.0)
    F, ref = sc.F.clone(), sc.ref.clone()
    if quantised:  # multiples of 2^-12: every Sobel sum is then float32-representable
        F = torch.round(F * 4096.0) / 4096.0
        ref = torch.round(ref * 4096.0) / 4096.0
    g = torch.Generator().manual_seed(7000 + seed)
    qh = F.permute(2, 0, 1).contiguous()[None]  # 1 x C x H x W float32
    rh = torch.rand((C, H, W), generator=g)
    if quantised:
        rh = torch.round(rh * 4096.0) / 4096.0
    cells = torch.randperm(H * W, generator=g)[:N]
    rows, cols = cells // W, cells % W
    rh[:, rows, cols] = ref.t()
    rh = rh[None].contiguous()
    step = image // H
    ref_xy = torch.stack([cols * step + 3, rows * step + 5], dim=1).numpy().astype(np.float64)  # (x, y) pixels
    T = np.eye(4)
    T[:3, :3], T[:3, 3] = sc.R0.numpy(), sc.t0.numpy()
    inlier_mask = None
    if with_mask:
        inlier_mask = np.sort(torch.randperm(N, generator=g)[: (3 * N) // 4].numpy())
    pred = Prediction(True, N, N, ref_xy, None, sc.pts3D.numpy().copy(), np.array([1.0, 0.0, 0.0, 0.0]), T, "ref.png",
                      None, inlier_mask)
    return dict(qh=qh, rh=rh, pred=pred, K=np.array(KS), image_shape=[image, image])


def track_arrays(model, N):
    tr = model.track_
    L = len(tr["costs"])
    out = dict(costs=np.array(tr["costs"], dtype=np.float64),
               Rs=np.stack([x.numpy() for x in tr["Rs"]]), ts=np.stack([x.numpy() for x in tr["ts"]]))
    # feature_pnp_multi passes a different number of points every round: keep ragged lists flat
    out["n_points"] = np.array([int(x.shape[0]) for x in tr["points2d"]], dtype=np.int32)
    out["points2d_flat"] = np.concatenate([x.numpy().astype(np.int32).reshape(-1, 2) for x in tr["points2d"]])
    out["mask_flat"] = np.concatenate([x.numpy().astype(np.uint8) for x in tr["mask"]])
    out["best_cost"] = np.float64(model.best_cost_.item())
    out["initial_cost"] = np.float64(model.initial_cost_.item())
    out["best_num_inliers"] = np.int32(model.best_num_inliers_)
    assert L == len(out["n_points"])
    return out


def save(name, case, cfg, out):
    p = case["pred"]
    d = dict(qh=case["qh"].numpy(), rh=case["rh"].numpy(), K=case["K"], image_shape=np.array(case["image_shape"], np.int32),
             points_3d=p.points_3d, reference_inliers=p.reference_inliers, matrix=p.matrix,
             inlier_mask=np.array([-1]) if p.inlier_mask is None else p.inlier_mask.astype(np.int64))
    d.update({k: (np.array(v) if isinstance(v, str) else v) for k, v in cfg.items()})
    d.update(out)
    path = os.path.join(GOLDEN, name + ".npz")
    np.savez_compressed(path, **d)
    print("%-30s entries=%3d cost %.9g -> %.9g (%d KB)" % (name, len(out["costs"]), out["costs"][0], out["costs"][-1],
                                                            os.path.getsize(path) // 1024))


def run_feature_pnp(name, case, n_iters, loss, lambda_, ratio, pyramid=None, sobel_normalized=False):
    """`sobel_normalized`: run the reference with a kornia whose spatial_gradient scales the Sobel kernels by 1/8
    (what newer kornia releases do by default; the reference's pinned 0.2.2 cannot be checked offline)."""
    mod, ref_model, ref_utils = load_reference_adapter()
    if sobel_normalized:
        plain = ref_utils.kornia.filters.spatial_gradient
        ref_utils.kornia.filters.spatial_gradient = functools.partial(plain, normalized=True)
    try:
        _run_feature_pnp(mod, ref_model, ref_utils, name, case, n_iters, loss, lambda_, ratio, pyramid, sobel_normalized)
    finally:
        if sobel_normalized:
            ref_utils.kornia.filters.spatial_gradient = plain


def _run_feature_pnp(mod, ref_model, ref_utils, name, case, n_iters, loss, lambda_, ratio, pyramid, sobel_normalized):
    mod.sparseFeaturePnP = functools.partial(ref_model.sparseFeaturePnP, n_iters=n_iters, loss_fn=getattr(ref_utils, loss),
                                             lambda_=lambda_, ratio_threshold=ratio)
    K = torch.from_numpy(case["K"])
    R, t, model = mod.feature_pnp(case["qh"], case["rh"], case["pred"], K, case["image_shape"], track=True,
                                  feature_pyramid=pyramid)
    T = np.eye(4)
    T[:3, :3], T[3, :3] = R.numpy(), t.numpy()
    out = track_arrays(model, case["pred"].points_3d.shape[0])
    out.update(out_R=R.numpy().copy(), out_t=t.numpy().copy(), quaternion=mod.matrix_utils.matrix_quaternion(T))
    cfg = dict(kind="feature_pnp", n_iters=np.int32(n_iters), loss=loss, lambda_=np.float64(lambda_),
               ratio_threshold=np.float64(np.nan if ratio is None else ratio),
               levels=np.array([[s, e] for s, e, _, _ in pyramid] if pyramid else np.zeros((0, 2)), dtype=np.int32),
               pyramid4=np.array([[s, e, -1 if ts is None else ts, -1 if ks is None else ks] for s, e, ts, ks in pyramid]
                                 if pyramid else np.zeros((0, 4)), dtype=np.int32),
               sobel_normalized=np.int32(1 if sobel_normalized else 0))
    save(name, case, cfg, out)


def run_feature_pnp_multi(name, case, n_iters, loss, lambda_, threshold):
    mod, ref_model, ref_utils = load_reference_adapter()
    mod.sparseFeaturePnP = functools.partial(ref_model.sparseFeaturePnP, n_iters=n_iters, loss_fn=getattr(ref_utils, loss),
                                             lambda_=lambda_)
    mod.find_inliers = functools.partial(fixed_find_inliers, ref_model, ref_utils, threshold=threshold)
    K = torch.from_numpy(case["K"])
    R, t, model = mod.feature_pnp_multi(case["qh"], case["rh"], case["pred"], K, case["image_shape"], track=True)
    out = track_arrays(model, case["pred"].points_3d.shape[0])
    out.update(out_R=R.numpy().copy(), out_t=t.numpy().copy())
    cfg = dict(kind="feature_pnp_multi", n_iters=np.int32(n_iters), loss=loss, lambda_=np.float64(lambda_),
               threshold=np.float64(threshold))
    save(name, case, cfg, out)


def run_resampled_pyramid(name, seed, pyramid, n_iters=8, loss="squared_loss", lambda_=0.01, C=24, H=32):
    """multilevel_optimization with target_size / kernel_size levels (model.py:193-210), reference tensors in."""
    ref_model, ref_utils = pyref.load()
    sc = synth.make_scene(seed, N=150, C=C, Hf=H, Wf=H, im_width=256, im_height=256, K=KS, device="cpu",
                          outlier_frac=0.1, grad_scale=1.0)
    inp = sc.reference_inputs()
    gx, gy = ref_utils.sobel_filter(inp["feature_map_query"])  # what the adapter would pass
    model = ref_model.sparseFeaturePnP(n_iters=n_iters, loss_fn=getattr(ref_utils, loss), lambda_=lambda_)
    R, t = model.multilevel_optimization(pyramid, inp["pts3D"], inp["feature_ref"], inp["feature_map_query"], gx, gy,
                                         inp["K"], inp["im_width"], inp["im_height"], R_init=inp["R_init"],
                                         t_init=inp["t_init"], track=True)
    out = track_arrays(model, 150)
    out.update(out_R=R.numpy().copy(), out_t=t.numpy().copy())
    d = dict(F_chw=inp["feature_map_query"].float().numpy(), ref=sc.ref.numpy(), pts3D=sc.pts3D.numpy(), K=sc.K.numpy(),
             R0=sc.R0.numpy(), t0=sc.t0.numpy(), im_width=np.int32(256), im_height=np.int32(256),
             pyramid=np.array([[s, e, -1 if ts is None else ts, -1 if ks is None else ks] for s, e, ts, ks in pyramid],
                              dtype=np.int32),
             n_iters=np.int32(n_iters), loss=np.array(loss), lambda_=np.float64(lambda_))
    d.update(out)
    path = os.path.join(GOLDEN, name + ".npz")
    np.savez_compressed(path, **d)
    print("%-30s entries=%3d cost %.9g -> %.9g (%d KB)" % (name, len(out["costs"]), out["costs"][0], out["costs"][-1],
                                                            os.path.getsize(path) // 1024))


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    run_feature_pnp("adapter_pnp_exact", make_adapter_case(1), 12, "geman_mcclure_loss", 0.01, None)
    run_feature_pnp("adapter_pnp_pyramid", make_adapter_case(2), 8, "squared_loss", 0.01, 0.8,
                    pyramid=[(16, 32, None, None), (4, 16, None, None), (0, 4, None, None)])
    run_feature_pnp("adapter_pnp_float", make_adapter_case(3, quantised=False), 12, "cauchy_loss", 0.01, None)
    run_feature_pnp("adapter_pnp_norm", make_adapter_case(6), 12, "geman_mcclure_loss", 0.01, None, sobel_normalized=True)
    run_feature_pnp("adapter_pnp_resampled", make_adapter_case(7, C=24), 8, "squared_loss", 0.01, None,
                    pyramid=[(16, 24, 16, None), (8, 16, None, 3), (0, 8, None, None)])
    run_feature_pnp_multi("adapter_multi_mask", make_adapter_case(4), 6, "squared_loss", 0.01, 0.8)
    run_feature_pnp_multi("adapter_multi_nomask", make_adapter_case(5, with_mask=False), 6, "squared_loss", 0.01, 0.8)
    run_resampled_pyramid("pyramid_resampled_a", 21, [(16, 24, 16, None), (8, 16, None, 3), (0, 8, None, None)])
    run_resampled_pyramid("pyramid_resampled_b", 22, [(12, 24, 48, 5), (0, 12, 24, 3)], loss="geman_mcclure_loss")


if __name__ == "__main__":
    main()
