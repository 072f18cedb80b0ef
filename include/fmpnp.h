/*
 * fmpnp.h -- C ABI of the B200-native feature-metric PnP refinement (libfmpnp.so).
 *
 * This is the drop-in boundary for ONE hot path of aunagar/FeatureMetric-PnP: the damped
 * Gauss-Newton / Levenberg-Marquardt loop of `sparseFeaturePnP` (featurePnP/model.py) and the
 * two cost-only helpers around it.  The reference has no FFI of its own (it is pure Python);
 * each entry point below names the reference interface it replaces.  The Python side of the
 * boundary (`featuremetric-pnp_b200/model.py`, same class / call signature as the reference)
 * reaches these symbols through ctypes -- see INTEGRATION.md.
 *
 * Conventions
 *   - plain C types, raw pointers and sizes only; no torch / C++ types in any signature;
 *   - every function returns an `int` status (FMPNP_SUCCESS == 0, negative on error) and never
 *     throws; `fmpnp_last_error()` gives the text of the last failure on that context;
 *   - pointers documented "device" are CUDA device pointers on the context's device; the caller
 *     owns every buffer it passes; the library owns only the workspace inside `fmpnp_ctx`;
 *     fmpnp_problem.F / Gx / Gy / ref may instead point to PINNED HOST memory that is mapped into
 *     the device's address space (cudaHostAlloc / cudaHostRegister; with unified addressing the
 *     host pointer is the device pointer): the kernels then fetch only the rows they gather over
 *     PCIe ("zero-copy") and the caller does not upload whole maps;
 *   - launches are asynchronous on the `stream` argument (a `cudaStream_t` passed as void*);
 *     a context must not be used from two host threads at once.  A context owns ONE
 *     workspace: a call issued on a different stream than the previous call is ordered behind it on the device
 *     (cudaStreamWaitEvent), so launches of one context never overlap -- use one context per stream for
 *     concurrency;
 *   - there is NO CPU fallback: without a CUDA device fmpnp_create() fails.
 *
 * Data layout (differs from the reference on purpose): dense maps are
 * channel-contiguous HWC float32, [Hf][Wf][C]; descriptors are [N][C] float32; points and poses
 * are float64.  `fmpnp_pack_chw_to_hwc` converts the reference's CHW float32/float64 maps.
 * All arithmetic is float64.
 */
#ifndef FMPNP_H
#define FMPNP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FMPNP_VERSION 103
#define FMPNP_MAX_LEVELS 8

typedef struct fmpnp_ctx fmpnp_ctx;

/* return codes */
enum {
    FMPNP_SUCCESS = 0,
    FMPNP_ERR_INVALID_ARGUMENT = -1,
    FMPNP_ERR_CUDA = -2,
    FMPNP_ERR_NO_DEVICE = -3,
    FMPNP_ERR_OUT_OF_MEMORY = -4
};

/* Robust losses: the gin-registered functions of featurePnP/helpers/utils.py:15-38
 * (`@squared_loss`, `@huber_loss`, `@cauchy_loss`, `@geman_mcclure_loss`) and
 * `barron_loss(x, alpha)` (:40-78). */
enum {
    FMPNP_LOSS_SQUARED = 0,
    FMPNP_LOSS_HUBER = 1,
    FMPNP_LOSS_CAUCHY = 2,
    FMPNP_LOSS_GEMAN_MCCLURE = 3,
    FMPNP_LOSS_BARRON = 4
};

/* Per-query exit path of the optimisation loop (fmpnp_result.status). */
enum {
    FMPNP_STATUS_OK = 0,            /* ran all iterations; BEST pose returned (model.py:491-494)            */
    FMPNP_STATUS_EMPTY_SUPPORT = 1, /* no point inside the image at some pose; CURRENT pose returned
                                       (model.py:316-320, 441-445)                                          */
    FMPNP_STATUS_NAN_STEP = 2,      /* NaN in the step; loop left early, BEST pose returned (the intended
                                       behaviour of model.py:411-413)                                       */
    FMPNP_STATUS_NO_INITIAL_COST = 3, /* multilevel only: compute_cost() was None, initial pose returned
                                       (model.py:183-187)                                                   */
    FMPNP_STATUS_INCOMPLETE = -1    /* the kernel gave up before this query finished (scheduler watchdog):
                                       the record is not a result                                           */
};

/* 6x6 solver for the damped step (replaces torch.lu + torch.lu_solve, model.py:51,61) */
enum {
    FMPNP_SOLVER_CHOLESKY = 0, /* warp-level Cholesky; falls back to pivoted LU if H is not SPD */
    FMPNP_SOLVER_LU = 1        /* warp-level LU with partial pivoting (LAPACK getrf/getrs order)  */
};

/* One refinement problem == the positional arguments of sparseFeaturePnP.forward
 * (model.py:246-248): pts3D, feature_ref, feature_map_query, feature_grad_x, feature_grad_y, K,
 * im_width, im_height, R_init, t_init.  The struct itself lives in HOST memory. */
typedef struct {
    const float *F;        /* device, [Hf][Wf][C]  feature_map_query                               */
    const float *Gx;       /* device, [Hf][Wf][C]  feature_grad_x  (may be NULL for cost-only calls) */
    const float *Gy;       /* device, [Hf][Wf][C]  feature_grad_y                                   */
    const float *ref;      /* device, [N][C]       feature_ref                                      */
    const double *pts3D;   /* device, [N][3]                                                        */
    const uint8_t *active; /* device, [N] or NULL: points with 0 are left out, which equals passing
                              pts3D[inliers], feature_ref[inliers] (optimize_feature_pnp.py:43-44) */
    double K[9];           /* row-major 3x3 */
    double R0[9];          /* R_init, row-major */
    double t0[3];          /* t_init */
    int32_t N, C, Hf, Wf;
    int32_t im_width, im_height;
} fmpnp_problem;

/* == sparseFeaturePnP.__init__ arguments (model.py:157-158) + the feature pyramid of
 * multilevel_optimization (model.py:179-213, channel slices; target_size / kernel_size = None). */
typedef struct {
    int32_t n_iters;        /* n_iters */
    int32_t loss_id;        /* loss_fn */
    double barron_alpha;    /* only for FMPNP_LOSS_BARRON */
    double lambda0;         /* lambda_ (0 = Gauss-Newton on the first iteration, model.py:46) */
    double ratio_threshold; /* ratio_threshold; NaN = None (ratio test off) */
    int32_t n_levels;       /* 0: forward() on all channels; >0: multilevel_optimization */
    int32_t levels[FMPNP_MAX_LEVELS][2]; /* [start, end) channel slices, in execution order */
    int32_t solver;         /* FMPNP_SOLVER_* */
    int32_t cluster_size;   /* 0 = automatic: the task scheduler (an evaluation is cut into tasks that any CTA
                               can run; every CTA stays busy for any batch size).  1, 2, 4, 8 or 16 = refine
                               each query with one thread-block cluster of that many CTAs instead */
    int32_t gather_every_row; /* 0 (default): a point that projects into the same cell as at its previous evaluation
                               keeps its record -- the reference's lookup is nearest-cell (indexing_,
                               model.py:74-97), so the record is a function of (point, cell, channel slice) and
                               the result is bit-identical; its rows are not read again.  1: gather the rows of
                               every supported point at every evaluation (4 N C 4 bytes per iteration) */
    int32_t pad_;
    double inlier_threshold;  /* 0 or NaN = off.  Otherwise the launch starts with find_inliers(pts3D, R_init, t_init,
                               ..., threshold = inlier_threshold, loss_fn = loss_id) (model.py:131-152) over the
                               active points and refines pts3D[inliers], feature_ref[inliers]: one round of
                               feature_pnp_multi's loop (optimize_feature_pnp.py:42-46) in a single launch.  Same
                               results as fmpnp_find_inliers_batch followed by fmpnp_refine_batch with its mask as
                               `active`; without a pyramid the pass's records serve the first evaluation, so the
                               rows are gathered once instead of twice */
    int32_t keep_records;     /* 0 (default): nothing carries over from one call to the next.  1: this call refines the SAME
                               problems as the previous fmpnp_refine_batch call of this context -- same F / Gx / Gy / ref /
                               pts3D arrays with unchanged contents, same sizes, same order; poses, K and masks may differ --
                               so the records that call left may be reused where a point still projects into the cell it
                               was last gathered for: rounds 2 and 3 of feature_pnp_multi (optimize_feature_pnp.py:20-47)
                               start where round 1 ended and gather next to nothing.  Single level only (n_levels == 0),
                               record reuse on, task scheduler; ignored (a normal call) when the pointers, sizes or batch
                               differ from the previous call's, or when any other call of the context came in between.
                               Results are bit-identical either way; only gathered_rows / reused_rows differ. */
    int32_t reserved;
} fmpnp_options;

/* What the callers read back (sparse_to_dense_predictor.py:257): the returned pose,
 * model.best_cost_, model.initial_cost_, model.best_num_inliers_. Lives in DEVICE memory. */
typedef struct {
    double R[9];
    double t[3];
    double best_cost;          /* best_cost_ of the last level that reached its i == 0 block (NaN if none) */
    double initial_cost;       /* forward(): mean robust cost at i == 0 of the first level;
                                  multilevel: compute_cost() = mean SSD on all channels (model.py:182) */
    int32_t best_num_inliers;  /* best_num_inliers_ (-1 if never set) */
    int32_t status;            /* FMPNP_STATUS_* of the last level */
    int32_t n_evaluations;     /* pose evaluations performed (1 + trial steps per level, + 1 for compute_cost) */
    int32_t level_status[FMPNP_MAX_LEVELS];
    int64_t gathered_rows;     /* sum over evaluations of (points whose rows were read x channels x maps read):
                                  x 4 B = bytes this query gathered (bench roofline) */
    int64_t reused_rows;       /* the same count for the supported points whose record was reused instead;
                                  gathered_rows + reused_rows = sum of (supported points x channels x maps) */
} fmpnp_result;

/* Optional per-iteration history == the reference's `track_` dict (model.py:170-176):
 * entry 0 of each level is the i == 0 state, then one entry per trial step (accepted or not).
 * All arrays are DEVICE memory owned by the caller. mask / kept are over all N points:
 * track_["mask"][e] = mask[e], track_["threshold_mask"][e] = kept[e][mask[e]]. */
typedef struct {
    int32_t cap;        /* entries per query; needs >= max(1, n_levels) * (n_iters + 1) */
    int32_t n_stride;   /* points per entry (>= max N) */
    int32_t *len;       /* [B]                      entries written */
    double *Rs;         /* [B][cap][9]  */
    double *ts;         /* [B][cap][3]  */
    double *costs;      /* [B][cap]     */
    int32_t *points2d;  /* [B][cap][n_stride][2]   (x, y), all points, unmasked */
    uint8_t *mask;      /* [B][cap][n_stride]      supported */
    uint8_t *kept;      /* [B][cap][n_stride]      supported and survived the ratio test */
} fmpnp_history;

int fmpnp_version(void);

/* Creates a context on CUDA device `device`. Fails (FMPNP_ERR_NO_DEVICE) without a GPU. */
int fmpnp_create(fmpnp_ctx **ctx, int device);
int fmpnp_destroy(fmpnp_ctx *ctx);
const char *fmpnp_last_error(const fmpnp_ctx *ctx);

/* Replaces sparseFeaturePnP.forward (model.py:246-494) and .multilevel_optimization
 * (model.py:179-213) for a batch of B independent queries: the whole LM loop (projection,
 * gathers, robust weights, normal equations, solve, se(3) update, accept/reject, damping
 * schedule, best-pose tracking) runs on the device without host synchronisation.
 *   probs   : HOST array of B problems (device pointers inside)
 *   results : DEVICE array of B results
 *   hist    : NULL, or history buffers (track=True) */
int fmpnp_refine_batch(fmpnp_ctx *ctx, const fmpnp_problem *probs, int32_t B, const fmpnp_options *opts,
                       fmpnp_result *results, const fmpnp_history *hist, void *stream);

/* Replaces sparseFeaturePnP.compute_cost (model.py:216-243) at pose (R0, t0) of each problem:
 * mean over supported points (after the raw-SSD ratio test if opts->ratio_threshold is set) of
 * 0.5 * ||F[cell] - ref||^2.  costs: DEVICE [B] (NaN where the reference returns None);
 * n_supported: DEVICE [B] or NULL. */
int fmpnp_compute_cost_batch(fmpnp_ctx *ctx, const fmpnp_problem *probs, int32_t B, const fmpnp_options *opts,
                             double *costs, int32_t *n_supported, void *stream);

/* No counterpart in the reference.  fmpnp_compute_cost_batch for
 * callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc): same arguments, same results (bit for bit), but the call performs NO cudaMalloc / cudaFree /
 * cudaMallocHost and uses none of the context's workspace.  The problem descriptors travel as a kernel parameter
 * and the kernel's scratch (per-point records and flags, the query counter) is a file-scope `static __device__`
 * array of sizeof = FMPNP_NOALLOC_MAX_QUERIES x FMPNP_NOALLOC_MAX_POINTS x 65 bytes (about 2.1 MB) that every GPU
 * which loads the library holds until the process exits.  Limits: B <= FMPNP_NOALLOC_MAX_QUERIES problems per call
 * and N <= FMPNP_NOALLOC_MAX_POINTS points per problem; anything larger is rejected with
 * FMPNP_ERR_INVALID_ARGUMENT (and a message naming the limit), never truncated.  ALL calls on a device share the
 * one array, whatever their context or stream: two calls that overlap on the GPU race, so order them on one stream
 * (or with events).  The array keeps the last call's contents; the library clears what has to start at zero. */
#define FMPNP_NOALLOC_MAX_QUERIES 8
#define FMPNP_NOALLOC_MAX_POINTS 4096
int fmpnp_compute_cost_batch_noalloc(fmpnp_ctx *ctx, const fmpnp_problem *probs, int32_t B, const fmpnp_options *opts,
                                     double *costs, int32_t *n_supported, void *stream);

/* Replaces find_inliers (model.py:131-152), mode "ratio_max", at pose (R0, t0):
 * inliers[b][n] = supported(n) && |rho_n| < threshold * max |rho|.  inliers: DEVICE [B][n_stride]. */
int fmpnp_find_inliers_batch(fmpnp_ctx *ctx, const fmpnp_problem *probs, int32_t B, int32_t loss_id,
                             double barron_alpha, double threshold, uint8_t *inliers, int32_t n_stride,
                             void *stream);

/* Layout pack for the reference's maps (feature_map_query etc. are C x Hf x Wf, model.py:253-255):
 * src DEVICE CHW (float32 if src_is_f64 == 0, else float64) -> dst DEVICE HWC float32. */
int fmpnp_pack_chw_to_hwc(fmpnp_ctx *ctx, const void *src, int32_t src_is_f64, float *dst, int32_t C, int32_t Hf,
                          int32_t Wf, void *stream);

/* Replaces `sobel_filter` (featurePnP/helpers/utils.py:81-104 = kornia.filters.spatial_gradient, called at
 * s2dhm/pose_prediction/optimize_feature_pnp.py:31,61 and model.py:207) fused with the layout pack: one pass
 * over a DEVICE CHW map (float32, or float64 if src_is_f64) writes the three DEVICE HWC float32 maps the
 * refinement gathers from.  Gx = [[-1,0,1],[-2,0,2],[-1,0,1]] (x) F, Gy = transposed kernel, summed in
 * float64.  border: FMPNP_BORDER_REPLICATE (kornia) or FMPNP_BORDER_ZERO (helpers/sobel_pytorch.py:41);
 * normalized != 0 scales by 1/8 (newer kornia default; the reference's stored known answer needs 0). */
enum { FMPNP_BORDER_REPLICATE = 0, FMPNP_BORDER_ZERO = 1 };
int fmpnp_sobel_pack(fmpnp_ctx *ctx, const void *src, int32_t src_is_f64, float *F, float *Gx, float *Gy, int32_t C,
                     int32_t Hf, int32_t Wf, int32_t border, int32_t normalized, void *stream);

/* Replaces the reference-descriptor lookup of the adapter (optimize_feature_pnp.py:56):
 * dst[n][c] = src[c][rows[n]][cols[n]].  src DEVICE CHW (float32 / float64), rows / cols DEVICE int32 [N]
 * (negative indices wrap once, as torch indexing does), dst DEVICE [N][C] float32. */
int fmpnp_gather_descriptors(fmpnp_ctx *ctx, const void *src, int32_t src_is_f64, const int32_t *rows,
                             const int32_t *cols, float *dst, int32_t C, int32_t H, int32_t W, int32_t N, void *stream);

/* Pyramid-level preparation of multilevel_optimization (model.py:196-200), DEVICE CHW in (float32 / float64),
 * DEVICE CHW float64 out:
 *   fmpnp_resize_bilinear_chw == nn.functional.interpolate(size=(Ho, Wo), mode='bilinear'), align_corners=False;
 *   fmpnp_blur_chw            == depthwise conv2d with a DEVICE k x k float64 kernel and zero padding `pad`
 *                                (the reference passes pad = 1 for every k): dst is (H+2pad-k+1) x (W+2pad-k+1). */
int fmpnp_resize_bilinear_chw(fmpnp_ctx *ctx, const void *src, int32_t src_is_f64, double *dst, int32_t C, int32_t H,
                              int32_t W, int32_t Ho, int32_t Wo, void *stream);
int fmpnp_blur_chw(fmpnp_ctx *ctx, const void *src, int32_t src_is_f64, const double *kernel, double *dst, int32_t C,
                   int32_t H, int32_t W, int32_t k, int32_t pad, void *stream);

/* Number of kernels this context has launched so far (bench.py's `gpu_launches`). */
int64_t fmpnp_launch_count(const fmpnp_ctx *ctx);

/* Launch geometry of the most recent kernel of this context: CTAs in the grid, CTAs per query
 * (thread-block cluster size; 1 under the task scheduler), bytes of shared-memory staging ring per CTA and
 * tasks per pose evaluation (0 = one CTA / cluster per query, no task queue).  Diagnostics. */
int fmpnp_last_geometry(const fmpnp_ctx *ctx, int32_t *grid, int32_t *cluster_size, int32_t *ring_bytes,
                        int32_t *tasks_per_evaluation);

/* Result records written straight into another GPU's memory (queries are sharded over the
 * GPUs of one box, the only exchange is the final gather of the fixed-size result records; the reference shards
 * a slice over independent jobs the same way, submit_cmu_auto.sh:17-43).  The gathering rank creates one buffer
 * for all queries and passes the 64-byte handle to the other ranks (any transport); they open it and hand
 * `opened + first_query * sizeof(fmpnp_result)` to fmpnp_refine_batch as `results`: the kernel's own stores of
 * the records travel over NVLink / PCIe peer access, no collective and no extra kernel.  The records of a launch
 * are complete once its stream has been synchronised on the WRITING rank.
 *   create : cudaMalloc + zero + cudaIpcGetMemHandle on the context's device
 *   open   : cudaIpcOpenMemHandle (peer access enabled lazily); close with fmpnp_shared_buffer_close
 *   fmpnp_copy_to_host : synchronous D2H of any device range (reading the gathered records without torch) */
#define FMPNP_IPC_HANDLE_BYTES 64
int fmpnp_shared_buffer_create(fmpnp_ctx *ctx, size_t bytes, void **dev_ptr, unsigned char *handle);
int fmpnp_shared_buffer_open(fmpnp_ctx *ctx, const unsigned char *handle, void **dev_ptr);
int fmpnp_shared_buffer_close(fmpnp_ctx *ctx, void *dev_ptr);
int fmpnp_shared_buffer_destroy(fmpnp_ctx *ctx, void *dev_ptr);
int fmpnp_copy_to_host(fmpnp_ctx *ctx, void *dst_host, const void *src_dev, size_t bytes, void *stream);

/* sizeof() of the ABI structs (0 problem, 1 options, 2 result, 3 history), so that a binding
 * in another language can verify its mirrors; -1 for an unknown index. */
int fmpnp_struct_size(int which);

#ifdef __cplusplus
}
#endif
#endif /* FMPNP_H */
