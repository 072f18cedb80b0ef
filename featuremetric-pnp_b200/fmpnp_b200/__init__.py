"""fmpnp_b200 -- Python host side of the B200-native feature-metric PnP refinement.

Thin marshalling layer over the C ABI of `include/fmpnp.h` (libfmpnp.so, hand-written sm_100a
CUDA).  PyTorch is used for device memory and streams only.  There is NO CPU fallback: every
entry point raises if the CUDA library or a CUDA device is missing.

Public surface
  lib()                      the ctypes handle (loads featuremetric-pnp_b200/lib/libfmpnp.so)
  Context                    owns one `fmpnp_ctx`
  refine_batch(...)          batched sparseFeaturePnP.forward / multilevel_optimization
  compute_cost_batch(...)    batched sparseFeaturePnP.compute_cost
  find_inliers_batch(...)    batched find_inliers
  pack_chw_to_hwc(...)       reference layout (CHW f32/f64) -> device layout (HWC f32)
  sobel_pack(...)            sobel_filter fused with that pack: CHW map -> HWC F, Gx, Gy
  gather_descriptors(...)    feature_ref[n] = hypercolumn[:, i_n, j_n] on the device
  resize_bilinear_chw / blur_chw   per-level resize / blur of multilevel_optimization
  torch.ops.fmpnp.*          the same calls registered as PyTorch custom ops (see ops.py)
The drop-in replacement of the reference's `model.py` lives one directory up.
"""
from .binding import (  # noqa: F401
    LOSS_IDS,
    STATUS_EMPTY_SUPPORT,
    STATUS_NAN_STEP,
    STATUS_NO_INITIAL_COST,
    STATUS_OK,
    Context,
    FmpnpError,
    Problem,
    ProblemBatch,
    SharedRecords,
    compute_cost_batch,
    compute_cost_batch_noalloc,
    default_context,
    find_inliers_batch,
    lib,
    lib_path,
    loss_id_of,
    blur_chw,
    gather_descriptors,
    pack_chw_to_hwc,
    refine_batch,
    resize_bilinear_chw,
    sobel_pack,
)

from . import ops  # noqa: E402,F401  (registers torch.ops.fmpnp.*)

__all__ = [
    "LOSS_IDS", "STATUS_OK", "STATUS_EMPTY_SUPPORT", "STATUS_NAN_STEP", "STATUS_NO_INITIAL_COST", "Context",
    "FmpnpError", "Problem", "ProblemBatch", "SharedRecords", "compute_cost_batch", "compute_cost_batch_noalloc", "default_context", "find_inliers_batch", "lib", "lib_path",
    "loss_id_of", "pack_chw_to_hwc", "refine_batch", "sobel_pack", "gather_descriptors", "resize_bilinear_chw",
    "blur_chw",
]
