// fmpnp_kernels.cuh -- launch-parameter structs shared by the kernels (fmpnp_kernels.cu) and
// the C ABI (fmpnp_capi.cu).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/fmpnp.h"

namespace fmpnp {

constexpr int kWarps = 8;                         // each streams its own points through its slice of the ring
constexpr int kThreads = kWarps * 32;             // 256
constexpr int kRecDoubles = 8;                    // per-point record: c, Sxx, Sxy, Syy, bx, by (+2 pad) = 64 B
constexpr int kRanges = 16;                       // fixed reduction ranges per query == max cluster size
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
    int32_t n_max;
    unsigned int* work_counter;
    // staging ring geometry (slot size / depth are derived per evaluation, see set_geometry)
    int32_t ring_per_warp;  // bytes of the shared-memory ring owned by one warp
    int32_t min_depth;      // minimum slots per warp
    int32_t cluster_size;   // CTAs per query (thread-block cluster); grid = clusters * cluster_size
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
int max_active_clusters(int cluster_size, int ring_bytes);
cudaError_t refine_allow_nonportable_clusters(int allow);
cudaError_t refine_set_max_dynamic_smem(int bytes);

}  // namespace fmpnp
