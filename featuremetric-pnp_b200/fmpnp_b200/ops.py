"""PyTorch custom ops over the C ABI:  torch.ops.fmpnp.{refine_batch, compute_cost_batch,
find_inliers_batch, pack_chw_to_hwc}.

The ops take tensors only (the dispatcher's requirement); each one marshals `data_ptr()`s into
the `fmpnp_problem` descriptors and calls libfmpnp.so on the current CUDA stream.  They are
registered for the CUDA backend alone -- calling them with CPU tensors raises, there is no CPU
kernel behind them.

    out = torch.ops.fmpnp.refine_batch(F, Gx, Gy, ref, pts3D, K, R0, t0, im_size, opts_f64, levels, track, active)
      F, Gx, Gy : lists of (Hf, Wf, C) float32 CUDA tensors (one per query)
      ref       : list of (N, C) float32;  pts3D: list of (N, 3) float64
      K, R0, t0 : (B, 3, 3), (B, 3, 3), (B, 3) float64 CPU tensors
      im_size   : (B, 2) int32 CPU tensor (im_width, im_height)
      opts_f64  : (6,) float64 CPU tensor [n_iters, loss_id, barron_alpha, lambda_, ratio_threshold (nan = None), solver],
                  optionally followed by [inlier_threshold (0 / nan = off), gather_every_row (0 / 1), keep_records (0 / 1)]
      levels    : (L, 2) int32 CPU tensor of channel slices (L = 0: plain forward)
      track     : bool; True also returns the per-iteration history (the reference's track_)
      active    : [] or one (N,) uint8 CUDA mask per query: points with 0 are left out (== pts3D[inliers])
      returns   : [results] or [results, len, Rs, ts, costs, points2d, mask, kept]; results is a
                  (B, 176) uint8 CUDA tensor of fmpnp_result records (view with binding.RESULT_DTYPE)
"""
from typing import List

import torch

from . import binding as _b

_LOSS_NAMES = {v: k for k, v in _b.LOSS_IDS.items()}


def _problems(F, Gx, Gy, ref, pts3D, K, R0, t0, im_size, grads=True, active=None):
    B = len(F)
    return [_b.Problem(F[i], Gx[i] if grads else None, Gy[i] if grads else None, ref[i], pts3D[i], K[i], R0[i], t0[i],
                       int(im_size[i, 0]), int(im_size[i, 1]), active=active[i] if active else None) for i in range(B)]


@torch.library.custom_op("fmpnp::refine_batch", mutates_args=(), device_types="cuda")
def refine_batch(F: List[torch.Tensor], Gx: List[torch.Tensor], Gy: List[torch.Tensor], ref: List[torch.Tensor],
                 pts3D: List[torch.Tensor], K: torch.Tensor, R0: torch.Tensor, t0: torch.Tensor, im_size: torch.Tensor,
                 opts_f64: torch.Tensor, levels: torch.Tensor, track: bool,
                 active: List[torch.Tensor]) -> List[torch.Tensor]:
    probs = _problems(F, Gx, Gy, ref, pts3D, K, R0, t0, im_size, active=active)
    o = [float(v) for v in opts_f64.tolist()]
    rt = None if o[4] != o[4] else o[4]
    lv = [(int(a), int(b)) for a, b in levels.tolist()] if levels.numel() else None
    loss = _LOSS_NAMES[int(o[1])]
    inl = o[6] if len(o) > 6 and o[6] == o[6] and o[6] != 0.0 else None
    out = _b.refine_batch(probs, int(o[0]), loss=loss, lambda_=o[3], ratio_threshold=rt, levels=lv,
                          solver=int(o[5]), barron_alpha=o[2] if loss == "barron_loss" else None, sync=False,
                          track=track, inlier_threshold=inl, gather_every_row=len(o) > 7 and o[7] != 0.0,
                          keep_records=len(o) > 8 and o[8] != 0.0)
    res = [out["results_dev"][: len(probs)].clone()]
    if track:
        h = out["history_dev"]
        res += [h[k] for k in HISTORY_KEYS]
    return res


HISTORY_KEYS = ("len", "Rs", "ts", "costs", "points2d", "mask", "kept")


@refine_batch.register_fake
def _(F, Gx, Gy, ref, pts3D, K, R0, t0, im_size, opts_f64, levels, track, active):
    B, dev = len(F), F[0].device
    res = [torch.empty((B, _b.RESULT_DTYPE.itemsize), dtype=torch.uint8, device=dev)]
    if track:
        cap = max(1, int(levels.shape[0])) * (int(opts_f64[0]) + 1)
        n = max(int(p.shape[0]) for p in pts3D)
        res += [torch.empty(B, dtype=torch.int32, device=dev), torch.empty((B, cap, 3, 3), dtype=torch.float64, device=dev),
                torch.empty((B, cap, 3), dtype=torch.float64, device=dev), torch.empty((B, cap), dtype=torch.float64, device=dev),
                torch.empty((B, cap, n, 2), dtype=torch.int32, device=dev), torch.empty((B, cap, n), dtype=torch.uint8, device=dev),
                torch.empty((B, cap, n), dtype=torch.uint8, device=dev)]
    return res


@torch.library.custom_op("fmpnp::compute_cost_batch", mutates_args=(), device_types="cuda")
def compute_cost_batch(F: List[torch.Tensor], ref: List[torch.Tensor], pts3D: List[torch.Tensor], K: torch.Tensor,
                       R0: torch.Tensor, t0: torch.Tensor, im_size: torch.Tensor,
                       ratio_threshold: float) -> torch.Tensor:
    probs = _problems(F, None, None, ref, pts3D, K, R0, t0, im_size, grads=False)
    rt = None if ratio_threshold != ratio_threshold else ratio_threshold
    costs, _ = _b.compute_cost_batch(probs, ratio_threshold=rt)
    return costs.clone()


@compute_cost_batch.register_fake
def _(F, ref, pts3D, K, R0, t0, im_size, ratio_threshold):
    return torch.empty((len(F),), dtype=torch.float64, device=F[0].device)


@torch.library.custom_op("fmpnp::find_inliers_batch", mutates_args=(), device_types="cuda")
def find_inliers_batch(F: List[torch.Tensor], ref: List[torch.Tensor], pts3D: List[torch.Tensor], K: torch.Tensor,
                       R0: torch.Tensor, t0: torch.Tensor, im_size: torch.Tensor, loss_id: int, barron_alpha: float,
                       threshold: float) -> torch.Tensor:
    probs = _problems(F, None, None, ref, pts3D, K, R0, t0, im_size, grads=False)
    loss = _LOSS_NAMES[int(loss_id)]
    return _b.find_inliers_batch(probs, threshold, loss=loss,
                                 barron_alpha=barron_alpha if loss == "barron_loss" else None).clone()


@find_inliers_batch.register_fake
def _(F, ref, pts3D, K, R0, t0, im_size, loss_id, barron_alpha, threshold):
    n = max(int(p.shape[0]) for p in pts3D)
    return torch.empty((len(F), n), dtype=torch.uint8, device=F[0].device)


@torch.library.custom_op("fmpnp::pack_chw_to_hwc", mutates_args=(), device_types="cuda")
def pack_chw_to_hwc(src: torch.Tensor) -> torch.Tensor:
    return _b.pack_chw_to_hwc(src)


@pack_chw_to_hwc.register_fake
def _(src):
    c, h, w = src.shape
    return torch.empty((h, w, c), dtype=torch.float32, device=src.device)
