// fmpnp_kernels.cuh -- launch-parameter structs shared by the kernels (fmpnp_kernels.cu) and
// the C ABI (fmpnp_capi.cu).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/fmpnp.h"

namespace fmpnp {

constexpr int kWarps = 8;                         // each streams its own points through its slice of the ring
constexpr int kThreads = kWarps * 32;             // 256
constexpr int kRecDoubles = 8;                    // per-point record: c, Sxx, Sxy, Syy, bx, by, reuse tag (+1 pad) = 64 B
constexpr int kRecTagSlot = 6;                    // the record's tag: (generation << 32 | cell) it was gathered for
constexpr int kRanges = 128;                      // max fixed reduction ranges per query == max tasks per evaluation
#ifndef FMPNP_RANGE_POINTS
#define FMPNP_RANGE_POINTS 128
#endif
constexpr int kRangePoints = FMPNP_RANGE_POINTS;  // points a fixed reduction range aims at (one transposition + lane tree per range)
constexpr int kMinRanges = 16;                    // min (small queries) == max cluster size
constexpr int kMaxCluster = 16;                   // CTAs of a thread-block cluster (non-portable size)
constexpr int kLaneStride = 34;                   // reduction scratch: doubles per transposed row (16-byte aligned rows, conflict-free 128-bit reads)
constexpr int kHalfSums = 14;                     // the 28 sums of a range are transposed in two rounds of 14
constexpr int kWarpScratchDoubles = kHalfSums * kLaneStride + 8 * 28;  // per warp: transposition + pairwise-tree stack
constexpr int kMinRingPerWarp = (kWarpScratchDoubles * 8 + 127) & ~127;  // bytes of ring a warp needs as reduction scratch
constexpr int kMaxDepth = 8;                      // max slots per sub-ring

enum KernelMode { kModeRefine = 0, kModeComputeCost = 1, kModeFindInliers = 2 };

// Device copy of one fmpnp_problem.
struct QueryDev {
    const float* F;
    const float* Gx;
    const float* Gy;
    const float* ref;
    const double* pts;
    const uint8_t* active;
    double K[9], R0[9], t0[3];
    int32_t N, C, Hf, Wf, im_w, im_h;
};

struct HistoryDev {
    int32_t enabled, cap, n_stride;
    int32_t* len;
    double* Rs;
    double* ts;
    double* costs;
    int32_t* points2d;
    uint8_t* mask;
    uint8_t* kept;
};

// ---------------------------------------------------------------------------------------------
// Task scheduler.  An evaluation of a query is cut into `units` point ranges; CTAs pull units from one
// global FIFO, the CTA that finishes the last unit of an evaluation reduces the per-point records,
// advances the query's LM state machine and queues the units of its next evaluation.
// ---------------------------------------------------------------------------------------------
enum TaskPhase { kPhaseCost = 0, kPhaseFirst = 1, kPhaseTrial = 2, kPhaseInliers = 3 };

struct __align__(16) QueryState {
    double R[9], t[3];    // current pose (last accepted)
    double Rn[9], tn[3];  // trial pose
    double Rb[9], tb[3];  // best pose
    double lin[27];       // accepted linearisation
    double prev_cost, best_cost, lambda, lr, initial_cost;
    double evR[9], evt[3];  // pose of the evaluation in flight
    double ev_limit;        // its ratio-test limit (published between round 0 and round 1)
    long long gathered, reused;
    int32_t best_n, has_best, status, flag, hist_len, n_evals, initial_set;
    int32_t level_status[FMPNP_MAX_LEVELS];
    int32_t phase, level, it;
    int32_t ev_grad, ev_c0, ev_Cl, ev_ratio_mode, ev_loss, ev_hist;  // the evaluation in flight (ev_grad: rows streamed)
    int32_t pending;      // its units still running
    int32_t ev_round;     // 0: streaming round, 1: summing round of an evaluation with a ratio test
};

struct SchedDev {
    unsigned long long* queue;  // [cap] entries (position << 32 | unit + 1), 0xff.. = empty
    unsigned int* head;         // next ticket
    unsigned int* tail;         // next position to fill
    unsigned int* done;         // finished queries
    unsigned int* error;        // watchdog / protocol error flag
    unsigned int* started;      // queries whose first evaluation has been queued
    unsigned long long wait_limit_ns;  // a CTA that waits this long for a task raises `error` (0 = wait forever)
    uint32_t cap;               // power of two
    int32_t units;              // tasks per evaluation: a power of two, 1..128
    QueryState* states;         // [B]
    double* part;               // [B][units][32] partial sums (+ max |rho|, NaN flag) of the tasks of the evaluation in flight
};

struct LaunchParams {
    const QueryDev* queries;
    int32_t B;
    int32_t mode;
    // options
    int32_t n_iters, loss_id, use_ratio, n_levels, solver;
    double alpha, lambda0, ratio_thr;
    int32_t levels[FMPNP_MAX_LEVELS][2];
    // outputs
    fmpnp_result* results;
    HistoryDev hist;
    double* cost_out;      // compute_cost mode
    int32_t* nsup_out;     // compute_cost mode (may be null)
    uint8_t* inliers_out;  // find_inliers mode
    int32_t inliers_stride;
    // per-CTA scratch
    double* rec;    // [grid][n_max][kRecDoubles]
    uint8_t* supp;  // [grid][n_max]
    uint8_t* inl;   // [grid][n_max] inlier mask of the fused find_inliers pass (inlier_on)
    int32_t n_max;
    unsigned int* work_counter;
    // staging ring geometry (slot size / depth are derived per evaluation, see set_geometry)
    int32_t ring_per_warp;  // bytes of the shared-memory ring owned by one warp
    int32_t min_depth;      // minimum slots per warp
    int32_t cluster_size;   // CTAs per query (thread-block cluster); grid = clusters * cluster_size
    SchedDev sched;         // task scheduler (scheduler == 1): rec / supp are then per QUERY, not per CTA
    int32_t scheduler;      // 0: one CTA (or cluster) per query, 1: task queue
    // Record reuse.  The reference looks features up at the NEAREST cell (indexing_, model.py:74-97), so a point's
    // record (its six channel sums) is a function of (point, cell, channel slice) only: while a point stays in
    // its cell from one evaluation to the next the record in `rec` is still the right one, bit for bit, and its
    // rows are not gathered again.  A record carries the tag (generation << 32 | cell); the generation is unique
    // per (launch, query, level) -- reuse_base + query * levels + level -- so nothing stale can ever match.
    int32_t reuse;          // 0: gather every row at every evaluation
    // Fused inlier selection (fmpnp_options.inlier_threshold): find_inliers (model.py:131-152) at (R0, t0) runs as
    // the first evaluation of the launch and the refinement continues on the points it keeps -- one iteration of
    // feature_pnp_multi's loop (optimize_feature_pnp.py:42-46) without a second launch, and (without a pyramid)
    // without a second gather: the pass leaves full records that the i == 0 evaluation reuses.
    int32_t inlier_on;
    double inlier_thr;
    uint32_t reuse_base;    // first generation of this launch (never 0)
    // Light evaluations (fmpnp_light.cuh): with record reuse, an evaluation after the first one of a level finds nearly
    // every record in place.  1: such an evaluation is run as a whole by the CTA that completed the previous evaluation
    // of the query -- the LM state stays in shared memory, nothing goes through the task queue -- with the latency-
    // oriented light_evaluate(); 0: every evaluation takes the general route.  Same bits either way.
    int32_t light;
    int32_t light_miss_den;  // ... while the previous evaluation gathered at most 1 / light_miss_den of its points
    int32_t l2_policy;      // 1: gathered rows evict_first, staged records evict_last in L2; 0: no preference
};

// Host-side launchers (defined in fmpnp_kernels.cu)
cudaError_t launch_refine(const LaunchParams& p, int grid, int ring_bytes, cudaStream_t stream);
cudaError_t launch_pack_chw_to_hwc(const void* src, int src_is_f64, float* dst, int C, int Hf, int Wf,
                                   cudaStream_t stream);
cudaError_t launch_sobel_pack(const void* src, int src_is_f64, float* F, float* Gx, float* Gy, int C, int H, int W,
                              int zero_pad, int normalized, cudaStream_t stream);
cudaError_t launch_gather_chw_rows(const void* src, int src_is_f64, const int32_t* rows, const int32_t* cols, float* dst,
                                   int C, int H, int W, int N, cudaStream_t stream);
cudaError_t launch_bilinear_resize_chw(const void* src, int src_is_f64, double* dst, int Cs, int H, int W, int Ho, int Wo,
                                       cudaStream_t stream);
cudaError_t launch_blur_chw(const void* src, int src_is_f64, const double* kern, double* dst, int Cs, int H, int W, int k,
                            int pad, cudaStream_t stream);
size_t refine_static_smem_bytes();
// compute_cost with its scratch in a file-scope device array (fmpnp_compute_cost_batch_noalloc)
constexpr int kStaticMaxQueries = FMPNP_NOALLOC_MAX_QUERIES;
constexpr int kStaticMaxPoints = FMPNP_NOALLOC_MAX_POINTS;
struct StaticQueries {
    QueryDev q[kStaticMaxQueries];
};
cudaError_t static_scratch_setup(int device, int ring_bytes, void** scratch);
cudaError_t launch_compute_cost_static(LaunchParams p, const StaticQueries& pack, void* scratch, int grid, int ring_bytes,
                                       cudaStream_t stream);
int max_active_clusters(int cluster_size, int ring_bytes);
cudaError_t refine_allow_nonportable_clusters(int allow);
cudaError_t refine_set_max_dynamic_smem(int bytes);

}  // namespace fmpnp
