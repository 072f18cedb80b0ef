"""Drop-in for the reference's `s2dhm/pose_prediction/optimize_feature_pnp.py` (the caller of the hot
path): same function names and argument meaning, but the optimizer inputs are built ON THE DEVICE.

Reference (optimize_feature_pnp.py:50-71): copies both hypercolumns to the CPU as float64, gathers the
reference descriptors in a Python loop, runs a CPU conv3d Sobel, then the CPU optimizer.
Here: the hypercolumns stay where the network left them (CUDA float32); `fmpnp_sobel_pack` turns the
query hypercolumn into the HWC F / Gx / Gy maps in one pass, `fmpnp_gather_descriptors` reads the
reference descriptors, and the refinement kernel runs on those buffers.  Only the 3D points, K and the
initial pose cross PCIe (a few hundred KB per query).

    from pose_prediction.optimize_feature_pnp import optimize_feature_pnp, feature_pnp, feature_pnp_multi
"""
import numpy as np
import torch

try:
    import gin
except ImportError:  # pragma: no cover
    from helpers import _gin_shim as gin

import fmpnp_b200 as _fm
from helpers.utils import sobel_filter, sobel_normalization
from model import _device, find_inliers, sparseFeaturePnP  # noqa: F401
from pose_prediction import matrix_utils


def _chw_on_device(hypercolumns):
    """(1, C, H, W) or (C, H, W) tensor / array, any device -> contiguous CUDA (C, H, W) float32/float64."""
    h = torch.as_tensor(hypercolumns)
    if h.dim() == 4:
        h = h.squeeze(0)
    if h.dim() != 3:
        raise ValueError("hypercolumns must be (1, C, H, W) or (C, H, W); got %s" % (tuple(h.shape),))
    if h.dtype not in (torch.float32, torch.float64):
        h = h.float()
    return h.to(_device()).contiguous()


def _reference_cells(reference_hypercolumns, reference_inliers, image_shape):
    """(row, col) cell of every reference keypoint, exactly as the reference computes it
    (optimize_feature_pnp.py:51-54): scale in float64, truncate toward zero, flip (x, y) -> (row, col)."""
    shp = reference_hypercolumns.shape
    relative_shape = np.array([shp[-2] / image_shape[0], shp[-1] / image_shape[1]])
    ref2d = torch.from_numpy(relative_shape).view(1, 2) * torch.from_numpy(np.asarray(reference_inliers))
    ref2d = torch.flip(ref2d.type(torch.IntTensor), (1,))
    return ref2d[:, 0].contiguous(), ref2d[:, 1].contiguous()


def build_problem(query_hypercolumns, reference_hypercolumns, prediction, K, image_shape, sobel_normalized=None):
    """Device-resident optimizer inputs of one query (an fmpnp_b200.Problem) from the two hypercolumns
    and a solve_pnp `Prediction` (fields used: points_3d, reference_inliers, matrix).
    `sobel_normalized`: 1/8-scaled Sobel kernels or not; None = helpers.utils.sobel_normalization() (gin-bindable;
    the reference's kornia default is unpinned, see there)."""
    q = _chw_on_device(query_hypercolumns)
    r = _chw_on_device(reference_hypercolumns)
    if sobel_normalized is None:
        sobel_normalized = sobel_normalization()
    # sobel_filter(feature_map_query) fused with the HWC pack (:57, :61)
    F, Gx, Gy = _fm.sobel_pack(q, normalized=bool(sobel_normalized))
    rows, cols = _reference_cells(r, prediction.reference_inliers, image_shape)
    ref = _fm.gather_descriptors(r, rows, cols)  # feature_ref (:56)
    pts = torch.from_numpy(np.ascontiguousarray(prediction.points_3d, dtype=np.float64).reshape(-1, 3)).to(q.device)
    T_init = np.asarray(prediction.matrix, dtype=np.float64)
    return _fm.Problem(F, Gx, Gy, ref, pts, torch.as_tensor(K).double().cpu(), T_init[:3, :3].copy(), T_init[:3, 3].copy(),
                       int(image_shape[0]), int(image_shape[1]))


def _with(prob, R=None, t=None, active=None):
    return _fm.Problem(prob.F, prob.Gx, prob.Gy, prob.ref, prob.pts3D, prob.K, prob.R0 if R is None else R,
                       prob.t0 if t is None else t, prob.im_width, prob.im_height, active=active)


def feature_pnp(query_hypercolumns, reference_hypercolumns, prediction, K, image_shape, track=False,
                feature_pyramid=None):
    """Reference optimize_feature_pnp.py:50-71.  Returns (R, t, model) with R, t CPU float64 tensors."""
    prob = build_problem(query_hypercolumns, reference_hypercolumns, prediction, K, image_shape)
    model = sparseFeaturePnP()  # parameters from gin
    if feature_pyramid is None:
        R, t = model.forward_packed(prob, track=track)
    else:
        if any(ts is not None or ks is not None for _, _, ts, ks in feature_pyramid):
            # resized / blurred levels (model.py:196-207) are rebuilt from the CHW hypercolumn level by level: the
            # drop-in multilevel_optimization does that on the device (resize / blur / Sobel + pack kernels)
            q = _chw_on_device(query_hypercolumns)
            gx, gy = sobel_filter(q)
            R, t = model.multilevel_optimization(
                feature_pyramid, prob.pts3D, prob.ref, q, gx, gy, torch.from_numpy(prob.K.reshape(3, 3).copy()),
                prob.im_width, prob.im_height, R_init=torch.from_numpy(prob.R0.reshape(3, 3).copy()),
                t_init=torch.from_numpy(prob.t0.copy()), track=track)
        else:
            R, t = model.multilevel_packed([(s, e) for s, e, _, _ in feature_pyramid], prob, track=track)
    return R, t, model


def feature_pnp_multi(query_hypercolumns, reference_hypercolumns, prediction, K, image_shape, track=False,
                      threshold=None, loss_fn=None):
    """Reference optimize_feature_pnp.py:20-47: three rounds of {refine on the current inliers, re-select
    inliers at the refined pose}.  `threshold` / `loss_fn` are forwarded to find_inliers when given (the
    reference takes them from gin: `find_inliers.threshold = 0.8`).

    A selection and the refinement that follows it run as ONE launch (fmpnp_options.inlier_threshold) when the
    mask itself is not needed on the host (track=False) and find_inliers uses the optimizer's loss; the
    selection after the last round, whose result the reference discards, is not computed."""
    prob = build_problem(query_hypercolumns, reference_hypercolumns, prediction, K, image_shape)
    kw = {}
    if threshold is not None:
        kw["threshold"] = threshold
    if loss_fn is not None:
        kw["loss_fn"] = loss_fn

    def selection_args():
        thr = kw.get("threshold", _gin_default(find_inliers, "threshold"))
        if thr is None:
            raise TypeError("find_inliers needs a threshold (gin: find_inliers.threshold = ...)")
        return float(thr), kw.get("loss_fn", _gin_default(find_inliers, "loss_fn") or "squared_loss")

    def select(R, t):
        thr, loss = selection_args()
        return _fm.find_inliers_batch([_with(prob, R, t)], thr, loss=loss)[0, :prob.N].contiguous()

    R, t = torch.from_numpy(prob.R0.reshape(3, 3).copy()), torch.from_numpy(prob.t0.copy())
    if getattr(prediction, "inlier_mask", None) is not None:
        m = np.zeros(prob.N)
        m[prediction.inlier_mask] = 1
        inliers = torch.from_numpy(m.astype(np.uint8)).to(prob.pts3D.device)
    else:
        print("No initial inliers found.")
        inliers = None  # selected at (R, t) by the first round
    model = sparseFeaturePnP()  # parameters from gin
    fused_before = False  # the previous round was such a launch on these very arrays: its records still apply
    for _ in range(3):
        if inliers is None:
            thr, loss = selection_args()
            if not track and _fm.loss_id_of(loss) == _fm.loss_id_of(model.loss_fn):
                # selection + refinement, one launch; from the second such round on the launch starts from the records
                # the previous one left (fmpnp_options.keep_records): a point that is still in the cell it was last
                # gathered for -- nearly all of them, the round starts where the previous one ended -- reads nothing
                R, t = model.forward_packed(_with(prob, R, t), inlier_threshold=thr, keep_records=fused_before)
                fused_before = True
                continue
            inliers = select(R, t)
        fused_before = False
        R, t = model.forward_packed(_with(prob, R, t, active=inliers), track=track,
                                    select=inliers.bool().cpu().numpy() if track else None)
        inliers = None  # re-selected at the refined pose by the next round
    return R, t, model


def _gin_default(fn, name):
    """Value gin (when installed) would inject for `fn.name`, else the Python default."""
    try:
        import gin as real_gin
        return real_gin.query_parameter("%s.%s" % (getattr(fn, "__name__", "find_inliers"), name))
    except Exception:
        import inspect
        target = getattr(fn, "__wrapped__", fn)
        try:
            d = inspect.signature(target).parameters[name].default
            return None if d is inspect.Parameter.empty else d
        except Exception:
            return None


@gin.configurable
def optimize_feature_pnp(query_hypercolumns, net, prediction, K, image_shape, track=False, feature_pyramid=None,
                         features="s2dhm"):
    """Reference optimize_feature_pnp.py:73-91: computes the reference hypercolumn with `net`, refines, and
    returns (list(t), list(quaternion), model)."""
    K = torch.from_numpy(np.asarray(K))
    print("Initial : {}".format(list(prediction.quaternion) + list(prediction.matrix[:3, 3])))
    if features == "d2-net":
        from s2dhm.d2net_utils.extract_dense_features import extract_dense_features
        reference_hypercolumns = extract_dense_features([prediction.reference_filename], to_cpu=False, resize=True)
    else:
        reference_hypercolumns, _ = net.compute_hypercolumn([prediction.reference_filename], to_cpu=False, resize=True)
    R, t, model = feature_pnp(query_hypercolumns, reference_hypercolumns, prediction, K, image_shape, track=track,
                              feature_pyramid=feature_pyramid)
    T = np.eye(4)
    T[:3, :3], T[3, :3] = R.numpy(), t.numpy()  # (sic) the reference stores t in the bottom row; only R is used
    quaternion = matrix_utils.matrix_quaternion(T)
    print("Final : {}".format(list(quaternion) + list(t.numpy())))
    return list(t.numpy()), list(quaternion), model
