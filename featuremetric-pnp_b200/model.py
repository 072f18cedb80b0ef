"""Drop-in for the reference's `featurePnP/model.py`: same module name, same gin-configurable
`sparseFeaturePnP` constructor and call signatures, same returned poses, attributes and `track_`
history -- but the whole Levenberg-Marquardt loop runs in ONE hand-written sm_100a CUDA kernel
(libfmpnp.so, reached through the C ABI of include/fmpnp.h).

Usage is unchanged for the reference's callers (s2dhm/pose_prediction/optimize_feature_pnp.py:17):

    sys.path.append('featuremetric-pnp_b200/')      # instead of 'featurePnP/'
    from model import sparseFeaturePnP, find_inliers

Differences that are deliberate and documented in DESIGN.md:
  * inputs may live on any device / float dtype; they are moved to the GPU once, packed to
    channel-contiguous HWC float32 (cached per source tensor) and the arithmetic is float64.
    The reference's callers feed float32-representable maps (`hypercolumn.type(DoubleTensor)`);
  * `useGPU` is accepted and ignored: there is no CPU path (a missing CUDA device raises);
  * a NaN step leaves the loop and returns the best pose -- the reference's intended behaviour
    (its `logging.warning` at model.py:412 is a NameError because `logging` is never imported);
  * `loss_fn` must be one of the gin-registered losses (or barron_loss with an alpha).
"""
from __future__ import division, print_function

import collections

import numpy as np
import torch
from torch import nn

try:
    import gin
except ImportError:  # pragma: no cover
    from helpers import _gin_shim as gin

import fmpnp_b200 as _fm
from helpers.utils import (barron_loss, cauchy_loss, gaussian_kernel2d, geman_mcclure_loss, huber_loss,  # noqa: F401
                           scaled_loss, sobel_filter, sobel_normalization, squared_loss)

# ------------------------------------------------------------------------------------------------
# packing cache: CHW (reference layout) -> HWC fp32 on the device, keyed by the source tensor
# ------------------------------------------------------------------------------------------------
_PACK_CACHE = collections.OrderedDict()
_PACK_CACHE_MAX = 8


def _device():
    if not torch.cuda.is_available():
        raise _fm.FmpnpError("sparseFeaturePnP needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _pack_map(m):
    """(C, Hf, Wf) tensor, any device / float dtype -> contiguous CUDA HWC float32."""
    if m.dim() != 3:
        raise ValueError("feature maps must be (C, Hf, Wf); got %s" % (tuple(m.shape),))
    key = (m.data_ptr(), tuple(m.shape), tuple(m.stride()), m.dtype, str(m.device), m._version)
    hit = _PACK_CACHE.get(key)
    if hit is not None:
        _PACK_CACHE.move_to_end(key)
        return hit[1]
    src = m
    if src.dtype not in (torch.float32, torch.float64):
        src = src.double()
    if src.device.type == "cpu":
        src = _upload_as_float32(src)
    else:
        src = src.to(_device(), non_blocking=True).contiguous()
    out = torch.ops.fmpnp.pack_chw_to_hwc(src)
    _PACK_CACHE[key] = (m, out)  # keeping `m` alive keeps data_ptr unique
    while len(_PACK_CACHE) > _PACK_CACHE_MAX:
        _PACK_CACHE.popitem(last=False)
    return out


_STAGING = {}  # numel -> [next slot, [(pinned float32 buffer, event or None), ...]]


def _upload_as_float32(src):
    """CPU float64 / float32 map -> CUDA float32, same values as packing the float64 map on the device
    (round-to-nearest either way).  The conversion runs on the host threads into a pinned buffer, so half the
    bytes cross PCIe and the copy is a DMA instead of a pageable staged copy; three buffers per size rotate."""
    dev = _device()
    n = src.numel()
    ring = _STAGING.get(n)
    if ring is None:
        if len(_STAGING) >= 4:
            _STAGING.clear()
        ring = _STAGING[n] = [0, [[torch.empty(n, dtype=torch.float32).pin_memory(), None] for _ in range(3)]]
    slot = ring[1][ring[0]]
    ring[0] = (ring[0] + 1) % len(ring[1])
    if slot[1] is not None:
        slot[1].synchronize()  # the upload that last used this buffer has left it
    buf = slot[0].view(src.shape)
    buf.copy_(src)
    out = buf.to(dev, non_blocking=True)
    slot[1] = torch.cuda.Event()
    slot[1].record()
    return out


def clear_pack_cache():
    _PACK_CACHE.clear()


def _dev_tensor(x, dtype):
    return torch.as_tensor(x).to(device=_device(), dtype=dtype).contiguous()


def _problem(pts3D, feature_ref, feature_map_query, feature_grad_x, feature_grad_y, K, im_width, im_height, R, t,
             c_range=None):
    F = _pack_map(feature_map_query)
    Gx = _pack_map(feature_grad_x) if feature_grad_x is not None else None
    Gy = _pack_map(feature_grad_y) if feature_grad_y is not None else None
    ref = _dev_tensor(feature_ref, torch.float32)
    if ref.dim() == 1:
        ref = ref[:, None].contiguous()
    pts = _dev_tensor(pts3D, torch.float64).reshape(-1, 3)
    if c_range is not None:  # a channel slice of the descriptors was passed against full maps
        raise ValueError("internal: c_range is handled through pyramid levels")
    return _fm.Problem(F, Gx, Gy, ref, pts, torch.as_tensor(K).double().cpu(), torch.as_tensor(R).double().cpu(),
                       torch.as_tensor(t).double().cpu(), int(im_width), int(im_height))


@gin.configurable
def find_inliers(pts3D, R, t, feature_map_query, feature_ref, K, im_width, im_height, threshold=None,
                 loss_fn=squared_loss, mode="ratio_max"):
    """Boolean mask over all points: projected inside the image AND robust cost below
    threshold * max robust cost (reference model.py:131-152).  Any mode other than "ratio_max"
    returns None, as in the reference."""
    if mode != "ratio_max":
        return None
    prob = _problem(pts3D, feature_ref, feature_map_query, None, None, K, im_width, im_height, R, t)
    out = _fm.find_inliers_batch([prob], threshold, loss=loss_fn)
    return out[0, :prob.N].bool().cpu()


@gin.configurable
class sparseFeaturePnP(nn.Module):
    def __init__(self, n_iters, loss_fn=squared_loss, lambda_=0.01, verbose=False, ratio_threshold=None,
                 useGPU=False):
        super().__init__()
        self.iterations = n_iters
        self.loss_fn = loss_fn
        self.verbose = verbose
        self.lambda_ = lambda_
        self.track_ = {"Rs": [], "ts": [], "costs": [], "points2d": [], "mask": [], "threshold_mask": []}
        self.use_ratio_test_ = ratio_threshold is not None
        self.ratio_threshold_ = ratio_threshold
        self.initial_cost_ = None
        self.useGPU = useGPU  # accepted for compatibility; the optimizer always runs on the GPU
        self.solver = "cholesky"
        _fm.loss_id_of(loss_fn)  # fail early on a loss the kernel cannot evaluate

    def track(self, R, t, cost, points_2d, mask, threshold_mask):
        self.track_["Rs"].append(R)
        self.track_["ts"].append(t)
        self.track_["costs"].append(cost)
        self.track_["points2d"].append(points_2d)
        self.track_["mask"].append(mask)
        self.track_["threshold_mask"].append(threshold_mask)

    # -- shared tail of forward / multilevel_optimization ---------------------------------------
    def _run(self, prob, levels, track, select=None, inlier_threshold=None, keep_records=False):
        """One query through torch.ops.fmpnp.refine_batch (PyTorch custom op -> C ABI -> CUDA kernel).
        `select`: boolean mask over the points; the history is reported for those points only (a
        caller that passes pts3D[inliers] to the reference sees exactly these rows).
        `inlier_threshold`: the launch starts with find_inliers(threshold, loss_fn=self.loss_fn) at the initial
        pose and refines the points it keeps (fmpnp_options.inlier_threshold).
        `keep_records`: the same arrays as the previous call, possibly at another pose (fmpnp_options.keep_records)."""
        loss_id, alpha = _fm.loss_id_of(self.loss_fn)
        rt = float("nan") if self.ratio_threshold_ is None else float(self.ratio_threshold_)
        solver = {"cholesky": 0, "lu": 1}[self.solver]
        opts = torch.tensor([float(self.iterations), float(loss_id), alpha, float(self.lambda_), rt, float(solver),
                             0.0 if inlier_threshold is None else float(inlier_threshold), 0.0,
                             1.0 if keep_records else 0.0], dtype=torch.float64)
        lv = torch.tensor(levels if levels else [], dtype=torch.int32).reshape(-1, 2)
        outs = torch.ops.fmpnp.refine_batch(
            [prob.F], [prob.Gx], [prob.Gy], [prob.ref], [prob.pts3D], torch.from_numpy(prob.K.reshape(1, 3, 3)),
            torch.from_numpy(prob.R0.reshape(1, 3, 3)), torch.from_numpy(prob.t0.reshape(1, 3)),
            torch.tensor([[prob.im_width, prob.im_height]], dtype=torch.int32), opts, lv, bool(track),
            [prob.active] if prob.active is not None else [])
        r = outs[0].cpu().numpy().view(_fm.binding.RESULT_DTYPE).reshape(-1)[0]
        if track:
            h = {k: v.cpu().numpy() for k, v in zip(_fm.ops.HISTORY_KEYS, outs[1:])}
            n = int(h["len"][0])
            N = prob.N
            sel = slice(None) if select is None else np.asarray(select, dtype=bool)
            for e in range(n):
                m = torch.from_numpy(h["mask"][0, e, :N].astype(bool)[sel])
                thr = None
                if self.use_ratio_test_:
                    thr = torch.from_numpy(h["kept"][0, e, :N].astype(bool)[sel])[m]
                self.track(torch.from_numpy(h["Rs"][0, e].copy()), torch.from_numpy(h["ts"][0, e].copy()),
                           float(h["costs"][0, e]), torch.from_numpy(h["points2d"][0, e, :N][sel].copy()), m, thr)
            if self.verbose:
                for e in range(n):
                    print("entry", e, float(h["costs"][0, e]))
        if int(r["best_num_inliers"]) >= 0:  # some level reached its i == 0 block
            self.best_cost_ = torch.tensor(float(r["best_cost"]), dtype=torch.float64)
            self.best_num_inliers_ = int(r["best_num_inliers"])
        self.last_status_ = int(r["status"])
        self.last_gathered_rows_, self.last_reused_rows_ = int(r["gathered_rows"]), int(r["reused_rows"])
        R = torch.from_numpy(np.array(r["R"], dtype=np.float64).reshape(3, 3).copy())
        t = torch.from_numpy(np.array(r["t"], dtype=np.float64).copy())
        return R, t, r

    @gin.configurable
    def multilevel_optimization(self, feature_pyramid, pts3D, feature_ref, feature_map_query, feature_grad_x,
                                feature_grad_y, *args, **kwargs):
        """Coarse-to-fine refinement over channel slices (reference model.py:179-213).
        args = (K, im_width, im_height); kwargs carry R_init, t_init and optionally track."""
        K, im_width, im_height = args[0], args[1], args[2]
        R_init, t_init = kwargs["R_init"], kwargs["t_init"]
        track = kwargs.get("track", False)
        if feature_pyramid is None:
            self.initial_cost_ = self.compute_cost(pts3D, R_init, t_init, feature_map_query, feature_ref, K, im_width,
                                                   im_height)
            if self.initial_cost_ is None:
                return R_init, t_init
            return self.forward(pts3D, feature_ref, feature_map_query, feature_grad_x, feature_grad_y, K, im_width,
                                im_height, R_init=R_init, t_init=t_init, track=track)
        if any(ts is not None or ks is not None for _, _, ts, ks in feature_pyramid):
            return self._multilevel_resampled(feature_pyramid, pts3D, feature_ref, feature_map_query, feature_grad_x,
                                              feature_grad_y, K, im_width, im_height, R_init, t_init, track)
        levels = [(int(start), int(end)) for start, end, _, _ in feature_pyramid]
        prob = _problem(pts3D, feature_ref, feature_map_query, feature_grad_x, feature_grad_y, K, im_width, im_height,
                        R_init, t_init)
        return self.multilevel_packed(levels, prob, track=track)

    def _multilevel_resampled(self, feature_pyramid, pts3D, feature_ref, feature_map_query, feature_grad_x,
                              feature_grad_y, K, im_width, im_height, R_init, t_init, track):
        """Pyramid with per-level resize (`target_size`) and / or Gaussian blur (`kernel_size`), reference
        model.py:193-210.  Each such level gets its own maps: channel slice -> bilinear resize -> depthwise
        blur (sigma = 1, zero padding 1 whatever the kernel size, as the reference) -> Sobel, all by the CUDA
        kernels of libfmpnp (float64 arithmetic, float32 storage); then one forward() per level."""
        self.initial_cost_ = self.compute_cost(pts3D, R_init, t_init, feature_map_query, feature_ref, K, im_width,
                                               im_height)
        if self.initial_cost_ is None:
            return R_init, t_init
        dev = _device()
        src = feature_map_query
        if src.dtype not in (torch.float32, torch.float64):
            src = src.double()
        src = src.to(dev).contiguous()
        ref = _dev_tensor(feature_ref, torch.float32)
        pts = _dev_tensor(pts3D, torch.float64).reshape(-1, 3)
        R, t = R_init, t_init
        for start, end, target_size, kernel_size in feature_pyramid:
            start, end = int(start), int(end)
            ref_l = ref[:, start:end].contiguous()
            if target_size is None and kernel_size is None:
                F = _pack_map(feature_map_query)[:, :, start:end].contiguous()
                Gx = _pack_map(feature_grad_x)[:, :, start:end].contiguous()
                Gy = _pack_map(feature_grad_y)[:, :, start:end].contiguous()
            else:
                local = src[start:end]
                if target_size is not None:
                    local = _fm.resize_bilinear_chw(local, int(target_size), int(target_size))
                if kernel_size is not None:
                    local = _fm.blur_chw(local, gaussian_kernel2d(int(kernel_size), 1.0), pad=1)
                F, Gx, Gy = _fm.sobel_pack(local, normalized=sobel_normalization())
            prob = _fm.Problem(F, Gx, Gy, ref_l, pts, torch.as_tensor(K).double().cpu(), torch.as_tensor(R).double().cpu(),
                               torch.as_tensor(t).double().cpu(), int(im_width), int(im_height))
            R, t = self.forward_packed(prob, track=track)
        return R, t

    def multilevel_packed(self, levels, prob, track=False):
        """multilevel_optimization over channel slices [(start, end), ...] on device-resident inputs
        (an fmpnp_b200.Problem: HWC float32 maps); what the device-side adapter calls."""
        R, t, r = self._run(prob, [(int(s), int(e)) for s, e in levels], track)
        if int(r["status"]) == _fm.STATUS_NO_INITIAL_COST:
            self.initial_cost_ = None
            return torch.from_numpy(prob.R0.reshape(3, 3).copy()), torch.from_numpy(prob.t0.copy())
        self.initial_cost_ = torch.tensor(float(r["initial_cost"]), dtype=torch.float64)
        return R, t

    def forward_packed(self, prob, track=False, select=None, inlier_threshold=None, keep_records=False):
        """forward() on device-resident inputs (an fmpnp_b200.Problem); with `inlier_threshold`, on the points
        find_inliers(threshold, loss_fn=self.loss_fn) keeps at the initial pose (same launch)."""
        R_out, t_out, r = self._run(prob, None, track, select, inlier_threshold, keep_records)
        if self.initial_cost_ is None and int(r["best_num_inliers"]) >= 0:
            self.initial_cost_ = torch.tensor(float(r["initial_cost"]), dtype=torch.float64)
        return R_out, t_out

    def forward_many(self, probs, inlier_threshold=None, levels=None):
        """forward() / multilevel_optimization (levels = channel slices) of MANY queries in one launch: `probs` is a list
        of fmpnp_b200.Problem (device-resident inputs, e.g. from pose_prediction.optimize_feature_pnp.build_problem).
        Not in the reference, whose API is one query per call: called one at a time a query costs 1.4-1.8 ms (launch
        latency: a single query cannot fill the GPU), in batches of 256 it costs 17 us.  Returns a list of (R, t) CPU
        float64 tensors and leaves the records (numpy, fmpnp_b200.binding.RESULT_DTYPE) in self.last_results_."""
        out = _fm.refine_batch(probs, int(self.iterations), loss=self.loss_fn, lambda_=float(self.lambda_),
                               ratio_threshold=self.ratio_threshold_, solver=self.solver,
                               levels=[(int(s), int(e)) for s, e in levels] if levels else None,
                               inlier_threshold=inlier_threshold)
        self.last_results_ = out["results"]
        return [(torch.from_numpy(np.array(r["R"], dtype=np.float64).reshape(3, 3).copy()),
                 torch.from_numpy(np.array(r["t"], dtype=np.float64).copy())) for r in out["results"]]

    def compute_cost(self, pts3D, R, t, feature_map_query, feature_ref, K, im_width, im_height):
        """Mean of 0.5 * ||F[cell] - ref||^2 over the supported points (after the raw-SSD ratio
        test when ratio_threshold is set); None when no point is supported (reference
        model.py:216-243)."""
        prob = _problem(pts3D, feature_ref, feature_map_query, None, None, K, im_width, im_height, R, t)
        costs, nsup = _fm.compute_cost_batch([prob], ratio_threshold=self.ratio_threshold_)
        if int(nsup[0].item()) == 0:
            return None
        return costs[0].cpu()

    @gin.configurable
    def forward(self, pts3D, feature_ref, feature_map_query, feature_grad_x, feature_grad_y, K, im_width, im_height,
                R_init=None, t_init=None, track=False, confidence=None, scale=None):
        """Refines (R, t); returns the best pose as CPU float64 tensors (reference model.py:246-494).
        `confidence` and `scale` are accepted and ignored, as in the reference."""
        pts3D = torch.as_tensor(pts3D)
        R = torch.eye(3, dtype=torch.float64) if R_init is None else R_init
        t = torch.tensor([1.0, 1.0, 0.0], dtype=torch.float64) if t_init is None else t_init
        prob = _problem(pts3D, feature_ref, feature_map_query, feature_grad_x, feature_grad_y, K, im_width, im_height,
                        R, t)
        return self.forward_packed(prob, track=track)
