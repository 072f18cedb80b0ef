"""Generates the committed golden fixtures under tests/golden/ by running the UNMODIFIED
reference (`/root/reference/featurePnP/model.py`, imported through oracle/pyref.py) on CPU.

Run in the build container only (the GPU box has no /root/reference):

    python oracle/gen_golden.py            # writes tests/golden/*.npz

Each fixture stores the exact inputs (fp32 maps in HWC, fp32 descriptors, fp64 points and
poses) and everything the reference produced with track=True: per-entry cost, pose, projected
pixel indices, support masks, ratio-test masks, and the returned pose / best_cost_ /
initial_cost_ / best_num_inliers_.  The fixtures pin (a) the C oracle (CPU tests) and (b) the
CUDA path (GPU tests).
"""
import math
import os
import sys
import warnings

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


def run_reference(inp, n_iters, loss, lambda_, ratio_threshold, levels=None, barron_alpha=None):
    """inp: dict from Scene.reference_inputs(). Returns dict of numpy outputs."""
    ref_model, ref_utils = pyref.load()
    if loss == "barron_loss":
        a = torch.tensor([barron_alpha], dtype=torch.float64)
        loss_fn = lambda x: ref_utils.barron_loss(x, a)  # noqa: E731
    else:
        loss_fn = getattr(ref_utils, loss)
    model = ref_model.sparseFeaturePnP(n_iters=n_iters, loss_fn=loss_fn, lambda_=lambda_,
                                       ratio_threshold=ratio_threshold)
    args = (inp["pts3D"], inp["feature_ref"], inp["feature_map_query"], inp["feature_grad_x"], inp["feature_grad_y"],
            inp["K"], inp["im_width"], inp["im_height"])
    if levels is None:
        R, t = model(*args, R_init=inp["R_init"], t_init=inp["t_init"], track=True)
    else:
        pyr = [(int(s), int(e), None, None) for s, e in levels]
        R, t = model.multilevel_optimization(pyr, *args, R_init=inp["R_init"], t_init=inp["t_init"], track=True)
    tr = model.track_
    N = inp["pts3D"].shape[0]
    L = len(tr["costs"])
    mask = np.zeros((L, N), dtype=np.uint8)
    kept = np.zeros((L, N), dtype=np.uint8)
    for i in range(L):
        m = tr["mask"][i].numpy().astype(bool)
        mask[i] = m
        if tr["threshold_mask"][i] is None:
            kept[i] = m
        else:
            k = np.zeros(N, dtype=bool)
            k[m] = tr["threshold_mask"][i].numpy().astype(bool)
            kept[i] = k

    def item(x):
        if x is None:
            return np.nan
        return float(x.item()) if hasattr(x, "item") else float(x)

    return dict(
        out_R=R.numpy().copy(), out_t=t.numpy().copy(),
        costs=np.array(tr["costs"], dtype=np.float64),
        Rs=np.stack([x.numpy() for x in tr["Rs"]]) if L else np.zeros((0, 3, 3)),
        ts=np.stack([x.numpy() for x in tr["ts"]]) if L else np.zeros((0, 3)),
        points2d=np.stack([x.numpy() for x in tr["points2d"]]).astype(np.int32) if L else np.zeros((0, N, 2), np.int32),
        mask=mask, kept=kept,
        best_cost=item(getattr(model, "best_cost_", None)),
        initial_cost=item(model.initial_cost_),
        best_num_inliers=int(getattr(model, "best_num_inliers_", -1)),
    )


def scene_arrays(sc):
    return dict(F=sc.F.numpy(), Gx=sc.Gx.numpy(), Gy=sc.Gy.numpy(), ref=sc.ref.numpy(), pts3D=sc.pts3D.numpy(),
                K=sc.K.numpy(), R0=sc.R0.numpy(), t0=sc.t0.numpy(),
                im_width=np.int32(sc.im_width), im_height=np.int32(sc.im_height))


def save_case(name, sc, n_iters, loss, lambda_, ratio_threshold, levels=None, barron_alpha=None, extra=None):
    out = run_reference(sc.reference_inputs(), n_iters, loss, lambda_, ratio_threshold, levels, barron_alpha)
    cfg = dict(n_iters=np.int32(n_iters), loss=np.array(loss), lambda_=np.float64(lambda_),
               ratio_threshold=np.float64(np.nan if ratio_threshold is None else ratio_threshold),
               levels=np.array(levels if levels is not None else np.zeros((0, 2)), dtype=np.int32),
               barron_alpha=np.float64(np.nan if barron_alpha is None else barron_alpha))
    d = {}
    d.update(scene_arrays(sc))
    d.update(cfg)
    d.update(out)
    if extra:
        d.update(extra)
    path = os.path.join(GOLDEN, name + ".npz")
    np.savez_compressed(path, **d)
    print("%-28s entries=%3d cost %.9g -> %.9g  inliers=%d  (%d KB)" % (
        name, len(out["costs"]), out["costs"][0] if len(out["costs"]) else np.nan,
        out["costs"][-1] if len(out["costs"]) else np.nan, out["best_num_inliers"], os.path.getsize(path) // 1024))


def toy_scene(idx=6, angle_deg=4.0, tz=-0.1, map_size=128):
    """featurePnP/toy_example/FeatureBA_ToyExample.ipynb cells 14-20, from the shipped fixtures
    (`*_data.p` is not shipped: rebuilt from the CSV rows with found == True, toy_example.py:122-123)."""
    import cv2
    import pandas as pd
    _, ref_utils = pyref.load()
    root = os.path.join(pyref.REFERENCE_ROOT, "featurePnP", "toy_example", "data", "toyexample_%d" % idx)
    df = pd.read_csv(root + "_data.csv", sep=";")
    df = df[df["found"] == True]  # noqa: E712
    pts = df[["X", "Y", "Z"]].values.astype(np.float64)
    p2 = df[["x", "y"]].values.astype(np.float64)
    K = np.load(root + "_K.npy")
    img = cv2.imread(root + ".png", 0).astype("double")
    img_t = torch.from_numpy(img)[None]
    gx, gy = ref_utils.sobel_filter(img_t)
    coords = np.around(p2).astype(int) - 1
    ref2d = np.flip(coords, axis=1)
    fref = np.stack([img[i, j] for i, j in ref2d])[:, None]
    a = angle_deg * math.pi / 180
    T = np.array([[math.cos(a), -math.sin(a), 0, 0], [math.sin(a), math.cos(a), 0, 0], [0, 0, 1, tz]])
    itp = lambda x: torch.nn.functional.interpolate(x.unsqueeze(0), size=(map_size, map_size), mode="bilinear").squeeze(0)  # noqa: E731
    F, Gx, Gy = itp(img_t), itp(gx), itp(gy)
    for m in (F, Gx, Gy):
        assert bool((m.float().double() == m).all()), "toy maps must be fp32-representable"
    hwc = lambda m: m.permute(1, 2, 0).contiguous().float()  # noqa: E731
    h, w = img.shape
    return synth.Scene(F=hwc(F), Gx=hwc(Gx), Gy=hwc(Gy), ref=torch.from_numpy(fref).float(),
                       pts3D=torch.from_numpy(pts), K=torch.from_numpy(K), R0=torch.from_numpy(T[:, :3].copy()),
                       t0=torch.from_numpy(T[:, 3].copy()), im_width=w, im_height=h)


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    torch.set_num_threads(1)
    # 1. the reference's own known answer (notebook cell 28: 27497.41105769231 -> 276.125)
    save_case("toy6_squared_lm", toy_scene(6), 50, "squared_loss", 0.01, None)
    save_case("toy5_geman_lm", toy_scene(5, angle_deg=3.0, tz=0.05), 30, "geman_mcclure_loss", 0.01, None)
    # 2. synthetic RobotCar-shaped scenes, small: 256 x 256 image over a 32 x 32 map keeps the
    #    reference's 8 px per cell; K scaled accordingly.
    Ks = ((100.0, 0.0, 127.055733), (0.0, 100.0, 124.546844), (0.0, 0.0, 1.0))
    base = dict(Hf=32, Wf=32, im_width=256, im_height=256, K=Ks, noise_sigma=0.01)
    mk = lambda q, **kw: synth.make_scene(q, device="cpu", **{**base, **kw})  # noqa: E731
    GS = 1.0 / 64  # un-normalised Sobel (x8) on 8 px cells -> d/dpixel: real Gauss-Newton sized steps
    # raw reference behaviour (Jacobian 64x too large -> tiny steps, cost hardly moves)
    save_case("syn_raw_squared_lm", mk(0, N=96, C=16), 20, "squared_loss", 0.01, None)
    save_case("syn_raw_geman_lm", mk(1, N=96, C=16, outlier_frac=0.1), 20, "geman_mcclure_loss", 0.01, None)
    # the four gin-registered losses + barron, with real LM dynamics (accept / reject)
    save_case("syn_squared_lm", mk(2, N=96, C=16, grad_scale=GS), 20, "squared_loss", 0.01, None)
    save_case("syn_geman_lm", mk(3, N=96, C=16, grad_scale=GS, outlier_frac=0.1), 20, "geman_mcclure_loss", 0.01, None)
    save_case("syn_cauchy_lm", mk(4, N=96, C=16, grad_scale=GS, outlier_frac=0.1), 20, "cauchy_loss", 0.01, None)
    save_case("syn_huber_lm", mk(5, N=96, C=16, grad_scale=GS, outlier_frac=0.1, noise_sigma=0.3), 20, "huber_loss",
              0.01, None)
    save_case("syn_barron_m1", mk(6, N=96, C=16, grad_scale=GS, outlier_frac=0.1), 12, "barron_loss", 0.01, None,
              barron_alpha=-1.0)
    # ratio test (GN: lambda_ = 0, the robotcar_inlier_GN.gin shape) and with a robust loss
    save_case("syn_gn_ratio08", mk(7, N=96, C=16, grad_scale=GS), 20, "squared_loss", 0.0, 0.8)
    save_case("syn_geman_ratio05", mk(8, N=96, C=16, grad_scale=GS, outlier_frac=0.1), 20, "geman_mcclure_loss", 0.01,
              0.5)
    # larger perturbation
    save_case("syn_dyn_squared", mk(9, N=128, C=16, grad_scale=GS, rot_deg=(1.0, 3.0), trans_sigma=0.15), 30,
              "squared_loss", 0.01, None)
    save_case("syn_dyn_geman", mk(10, N=128, C=16, grad_scale=GS, rot_deg=(1.0, 3.0), trans_sigma=0.15,
                                  outlier_frac=0.1), 30, "geman_mcclure_loss", 0.01, 0.8)
    # pyramids (channel slices, deep -> shallow as the shipped configs order them)
    save_case("syn_pyramid_squared", mk(11, N=96, C=32, grad_scale=GS), 10, "squared_loss", 0.01, None,
              levels=[(16, 32), (4, 16), (0, 4)])
    save_case("syn_pyramid_ratio06", mk(12, N=96, C=32, outlier_frac=0.1, grad_scale=GS), 10, "squared_loss", 0.01,
              0.6, levels=[(16, 32), (4, 16), (0, 4)])
    # channel counts that are not multiples of 4 (generic load path), non-square map and image
    odd = dict(Hf=24, Wf=40, im_width=320, im_height=192, K=((110.0, 0.0, 158.3), (0.0, 90.0, 97.6), (0.0, 0.0, 1.0)))
    save_case("syn_oddC_squared", mk(13, N=80, C=5, grad_scale=GS, **odd), 15, "squared_loss", 0.01, None)
    save_case("syn_oddC_pyramid", mk(14, N=80, C=11, grad_scale=GS, **odd), 8, "cauchy_loss", 0.01, None,
              levels=[(6, 11), (1, 6), (0, 1)])
    # wider channel count, closer to the bench shape
    save_case("syn_c128_geman", mk(15, N=160, C=128, Hf=16, Wf=16, im_width=128, im_height=128,
                                   K=((50.0, 0.0, 63.5), (0.0, 50.0, 62.3), (0.0, 0.0, 1.0)), grad_scale=GS,
                                   outlier_frac=0.1), 20, "geman_mcclure_loss", 0.01, None)

    # find_inliers / compute_cost known answers on one scene
    ref_model, ref_utils = pyref.load()
    sc = mk(16, N=96, C=16, outlier_frac=0.15)
    inp = sc.reference_inputs()
    inl = {}

    def find_inliers_ref(loss_fn, threshold):
        """model.py:131-152 composed from the reference's OWN building blocks, in its order.  The
        function itself cannot run under torch 2.x: its last statement
        `mask_supported[mask_supported] = threshold_mask` (:151) is rejected as a self-overlapping
        write (it was accepted by the torch 1.4 the reference pins); the only change here is that
        the write goes to a clone of the mask."""
        pts3D, R, t, K = inp["pts3D"], inp["R_init"], inp["t_init"], inp["K"]
        points_3d = torch.mm(R, pts3D.T).T + t
        points_2d = torch.round(ref_utils.from_homogeneous(torch.mm(K, points_3d.T).T)).type(torch.IntTensor) - 1
        mask_supported = ref_model.points_within_image(points_2d, inp["im_width"], inp["im_height"])
        p2s = points_2d[mask_supported, :]
        error = ref_model.indexing_(inp["feature_map_query"], torch.flip(p2s, (1,)), inp["im_width"],
                                    inp["im_height"]) - inp["feature_ref"][mask_supported]
        cost = 0.5 * (error ** 2).sum(-1)
        cost_full, _, _ = loss_fn(cost)
        threshold_mask = ref_model.ratio_threshold_feature_errors(cost_full, threshold=threshold)
        out = mask_supported.clone()
        out[mask_supported] = threshold_mask
        return out

    for name in ("squared_loss", "geman_mcclure_loss", "cauchy_loss"):
        inl["inliers_" + name] = find_inliers_ref(getattr(ref_utils, name), 0.8).numpy().astype(np.uint8)
    model = ref_model.sparseFeaturePnP(n_iters=1)
    cc = model.compute_cost(inp["pts3D"], inp["R_init"], inp["t_init"], inp["feature_map_query"], inp["feature_ref"],
                            inp["K"], inp["im_width"], inp["im_height"])
    model2 = ref_model.sparseFeaturePnP(n_iters=1, ratio_threshold=0.7)
    cc2 = model2.compute_cost(inp["pts3D"], inp["R_init"], inp["t_init"], inp["feature_map_query"],
                              inp["feature_ref"], inp["K"], inp["im_width"], inp["im_height"])
    inl["compute_cost"] = np.float64(cc.item())
    inl["compute_cost_ratio07"] = np.float64(cc2.item())
    save_case("syn_inliers_cost", sc, 5, "squared_loss", 0.01, None, extra=inl)


if __name__ == "__main__":
    main()
