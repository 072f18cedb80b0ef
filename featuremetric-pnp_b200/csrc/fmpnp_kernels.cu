// fmpnp_kernels.cu -- the feature-metric PnP refinement as ONE persistent sm_100a kernel per batch.
//
// Replaces sparseFeaturePnP.forward / multilevel_optimization / compute_cost and find_inliers
// of featurePnP/model.py (reference lines cited inline).  The whole Levenberg-Marquardt loop of
// every query runs on the device: no host synchronisation, no per-iteration launch, no
// collective.  Queries reach CTAs in one of two ways (same device code below them):
//   task scheduler   (*_tasks_kernel, default) a pose evaluation is cut into tasks on a global
//                    FIFO; any CTA runs any task; the CTA that completes an evaluation advances
//                    the query's LM state (kept in global memory) and queues the next evaluation;
//   CTA / cluster    (*_kernel) a CTA, or a thread-block cluster of up to 16 CTAs exchanging
//                    partial sums through distributed shared memory, owns one query at a time.
//
// Per pose evaluation (the only part that touches HBM in bulk):
//   every warp      projects 32 points at a time in float64 (bit-exact pixel indices), then
//                   streams their channel-contiguous rows F[cell], Gx[cell], Gy[cell], ref[n]
//                   through its own slice of a shared-memory ring with cp.async (SASS
//                   LDGSTS.E.BYPASS.128, one commit group per item, several items in flight),
//                   the next batch projected while the last rows of the current one are still
//                   landing.  The staged rows are read as float4 and accumulated in float64
//                   into the per-point c = sum e^2, the 2x2 structure tensor S = G^T G and
//                   b = G^T e; short rows share a warp (L lanes per point), long rows are cut
//                   into channel chunks.  One 48-byte record per point;
//   all threads     apply the robust loss / ratio test, contract (S, b) with the 2x6
//                   projection Jacobian and reduce the 6x6 normal equations in a fixed order
//                   (128 point ranges, pairwise tree: the same bits for any launch geometry);
//   warp 0          solves the damped system (warp-level Cholesky, pivoted-LU fallback),
//                   applies the se(3) update and runs the accept / reject / damping schedule.
// The trial evaluation of iteration i doubles as the linearisation of iteration i+1 when the
// step is accepted, so every LM iteration gathers each map exactly once: 4*N*C*4 bytes.
//
// Channel ranges that are not 16-byte aligned (C, slice start or slice length not a multiple
// of 4) take a generic path with direct coalesced global loads instead of the staged one.
#include <cmath>
#include <cstdint>

#include <cooperative_groups.h>

#include "fmpnp_device.cuh"
#include "fmpnp_kernels.cuh"

namespace cg = cooperative_groups;

namespace fmpnp {

struct __align__(16) Shared {
    // --- LM state (thread 0 / warp 0 write, everybody reads after a barrier)
    double R[9], t[3];    // current pose (last accepted)
    double Rn[9], tn[3];  // trial pose
    double Rb[9], tb[3];  // best pose
    double lin[27];       // accepted linearisation: upper triangle of H (21) + g (6)
    double ev[28];        // last evaluation: H (21), g (6), sum of rho
    double A[36], b[8];   // solver workspace
    double red[kWarps][32];
    double wtot[kWarps][28];  // per-warp totals of the ranges a warp owns (reduce_points), group totals (combine_parts)
    double climit;
    double cpart[2][32];  // this CTA's partial sums, read by the other CTAs of the cluster (double-buffered)
    double prev_cost, best_cost, lambda, lr, ev_cost, limit, initial_cost;
    long long gathered;
    int32_t red_i[kWarps][4];
    int32_t cpart_i[2][4];
    int32_t climit_nan;
    int32_t ev_count, ev_nsup, best_n, has_best, status, flag, hist_len, query, n_evals, initial_set;
    int32_t level_status[FMPNP_MAX_LEVELS];
    // --- the supported points of the two batches a warp works on (double-buffered): cell and index
    int32_t bl_cell[kWarps][2][32];
    int32_t bl_n[kWarps][2][32];
};

// Position inside a warp's slice of the ring.  The ring is empty between evaluations, so an evaluation
// with a different depth simply starts again at slot 0.
struct RingPos {
    int32_t pos;   // next slot to consume
    int32_t cpar;  // which half of Shared::cpart the next cluster reduction uses
};

struct EvalCtx {
    const QueryDev* q;
    int32_t c0, Cl;        // channel slice [c0, c0 + Cl)
    bool aligned;          // staged (16-byte) path allowed
    int32_t loss_id;
    double alpha;
    int32_t ratio_mode;    // 0 none, 1 on rho (forward / find_inliers), 2 on raw SSD (compute_cost)
    double thr;
    double* rec;           // per-point records of this CTA
    uint8_t* supp;         // per-point support flags of this CTA
    int32_t* h_px;         // history slices for the entry being evaluated (or null)
    uint8_t* h_mask;
    uint8_t* h_kept;
    uint8_t* inl_out;      // find_inliers output (or null)
    unsigned char* ring;
    int32_t ring_per_warp, min_depth;  // launch geometry: bytes of staging ring per warp
    int32_t depth, slot_bytes;         // per-evaluation geometry (set_geometry)
    int32_t CH, nCh;                   // channels per staged item, items per point
    int32_t L;                         // lanes sharing one point when a whole point is one item (nCh == 1)
    int32_t n0, n1;                    // this CTA's share of the points: [n0, n1)
    int32_t range_len;                 // points per fixed reduction range (kRanges ranges cover N)
    int32_t csize, crank;              // thread-block cluster working on this query (or: tasks of this evaluation)
    double* part_out;                  // task scheduler: this task's partial sums go here (global), no combine
    bool limit_given;                  // task scheduler: sh.limit already holds the ratio-test limit
};

// Staging geometry of one evaluation.
//   * Short rows: L = max(4, next power of two >= Cl / 8) lanes share a point (2 float4 of every
//     row per lane and pass), an item is the 32 / L points a warp accumulates together.
//   * Rows of more than 256 channels: the whole warp works on one point; the point is one item
//     if `min_depth` of them fit the warp's ring slice, else it is cut into equal chunks of a
//     multiple of 128 channels.
// The lane <-> channel mapping (and so the rounding of the per-point sums) depends on the slice
// length only, never on the ring size: results are bit-identical for every launch geometry.
__device__ __forceinline__ void set_geometry(EvalCtx& c, int rows) {
    const int W = c.ring_per_warp;
    const int nvec = c.Cl >> 2;
    int L = 4;
    while (L < 32 && L * 2 < nvec) L <<= 1;
    c.L = L;
    c.CH = c.Cl;
    c.nCh = 1;
    int item = (32 / L) * rows * c.Cl * 4;
    if (L == 32 && ((item + 127) & ~127) * c.min_depth > W) {
        int maxCH = ((W / c.min_depth) / (rows * 4)) & ~127;
        if (maxCH < 128) maxCH = 128;
        c.nCh = (c.Cl + maxCH - 1) / maxCH;
        const int per = (c.Cl + c.nCh - 1) / c.nCh;
        c.CH = min(maxCH, (per + 127) & ~127);
        c.nCh = (c.Cl + c.CH - 1) / c.CH;
        item = rows * c.CH * 4;
    }
    c.slot_bytes = (item + 127) & ~127;
    c.depth = max(1, min(kMaxDepth, W / c.slot_bytes));
}

// ---------------------------------------------------------------------------------------------
// channel accumulation: c += e^2, Sxx += gx^2, Sxy += gx gy, Syy += gy^2, bx += gx e, by += gy e
// (model.py:322, 338, 386-405 after the J = G A factorisation)
// ---------------------------------------------------------------------------------------------
template <bool GRAD>
__device__ __forceinline__ void acc1(float f, float r, float gx, float gy, double (&a)[6]) {
    const double e = __dadd_rn((double)f, -(double)r);
    a[0] = fma(e, e, a[0]);
    if (GRAD) {
        const double dx = (double)gx, dy = (double)gy;
        a[1] = fma(dx, dx, a[1]);
        a[2] = fma(dx, dy, a[2]);
        a[3] = fma(dy, dy, a[3]);
        a[4] = fma(dx, e, a[4]);
        a[5] = fma(dy, e, a[5]);
    }
}

// projects point n, records support / history, returns the feature-map cell or -1
__device__ __forceinline__ int32_t project_and_flag(const EvalCtx& c, const double* R, const double* t, int n) {
    const QueryDev& q = *c.q;
    const bool act = (q.active == nullptr) || (q.active[n] != 0);
    const double x = q.pts[3 * n + 0], y = q.pts[3 * n + 1], z = q.pts[3 * n + 2];
    const Proj pr = project_point(R, t, q.K, x, y, z, q.im_w, q.im_h);
    const bool sup = act && pr.supported;
    c.supp[n] = sup ? 1 : 0;
    if (c.h_px) {
        c.h_px[2 * n + 0] = pr.px;
        c.h_px[2 * n + 1] = pr.py;
        c.h_mask[n] = sup ? 1 : 0;
    }
    return sup ? cell_index(pr.px, pr.py, q.Hf, q.Wf, q.im_w, q.im_h) : -1;
}

// ---------------------------------------------------------------------------------------------
// Staged streaming (16-byte aligned slices).  Every warp is its own producer and consumer: it
// copies the rows of its next items into its slice of the shared-memory ring with cp.async
// (SASS LDGSTS.BYPASS.128: L2 -> shared memory, no registers, one commit group per item) and
// accumulates the oldest item once it has landed.  LOG2L = log2 of the lanes sharing one point.
// ---------------------------------------------------------------------------------------------
template <bool GRAD, int LOG2L>
__device__ __forceinline__ void reduce_lanes(double (&a)[6]) {
    constexpr int NV = GRAD ? 6 : 1;
#pragma unroll
    for (int s = LOG2L - 1; s >= 0; --s) {
#pragma unroll
        for (int i = 0; i < NV; ++i) a[i] = __dadd_rn(a[i], shfl_xor_f64(a[i], 1 << s));
    }
}

// Full-warp sum of the six per-point values with a halving butterfly: at distance 16 each lane keeps
// three values and hands the other three to its partner, at 8 and 4 it keeps two, then one, and the
// last two steps are a plain butterfly on that one value.  Every pairwise addition is the one the
// straight butterfly (reduce_lanes<GRAD, 5>) performs, so the totals are bit-identical to it; it
// just needs 8 additions and 16 shuffles instead of 30 and 60.  Afterwards the lanes with
// (lane & 3) == 0 hold value number v = {0: a0, 4: a1, 8: a2, 12: 0, 16: a3, 20: a4, 24: a5, 28: 0}[lane].
__device__ __forceinline__ double reduce_six(const double (&a)[6], int lane) {
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4;
    double k0 = h16 ? a[3] : a[0], k1 = h16 ? a[4] : a[1], k2 = h16 ? a[5] : a[2];
    const double s0 = h16 ? a[0] : a[3], s1 = h16 ? a[1] : a[4], s2 = h16 ? a[2] : a[5];
    k0 = __dadd_rn(k0, shfl_xor_f64(s0, 16));
    k1 = __dadd_rn(k1, shfl_xor_f64(s1, 16));
    k2 = __dadd_rn(k2, shfl_xor_f64(s2, 16));
    // distance 8: lanes with bit 3 clear keep (k0, k1), the others keep (k2, -)
    double m0 = h8 ? k2 : k0, m1 = h8 ? 0.0 : k1;
    const double t0 = h8 ? k0 : k2, t1 = h8 ? k1 : 0.0;
    m0 = __dadd_rn(m0, shfl_xor_f64(t0, 8));
    m1 = __dadd_rn(m1, shfl_xor_f64(t1, 8));
    // distance 4: bit 2 clear keeps m0, set keeps m1
    double v = h4 ? m1 : m0;
    const double u = h4 ? m0 : m1;
    v = __dadd_rn(v, shfl_xor_f64(u, 4));
    v = __dadd_rn(v, shfl_xor_f64(v, 2));
    v = __dadd_rn(v, shfl_xor_f64(v, 1));
    return v;
}

// The same halving butterfly over a group of L = 1 << LOG2L lanes (2 <= LOG2L <= 4) that share one point:
// six values -> three -> two -> one, then plain butterfly steps.  Lane sl of the group ends up with the
// totals number s0 (in v0) and, for groups of four lanes only, s1 (in v1); -1 = nothing to write.
template <int LOG2L>
__device__ __forceinline__ void reduce_group(const double (&a)[6], int sl, double& v0, double& v1, int& s0, int& s1) {
    constexpr int D = 1 << (LOG2L - 1);
    const bool hA = sl & D, hB = sl & (D >> 1);
    double k0 = hA ? a[3] : a[0], k1 = hA ? a[4] : a[1], k2 = hA ? a[5] : a[2];
    const double x0 = hA ? a[0] : a[3], x1 = hA ? a[1] : a[4], x2 = hA ? a[2] : a[5];
    k0 = __dadd_rn(k0, shfl_xor_f64(x0, D));
    k1 = __dadd_rn(k1, shfl_xor_f64(x1, D));
    k2 = __dadd_rn(k2, shfl_xor_f64(x2, D));
    double m0 = hB ? k2 : k0, m1 = hB ? 0.0 : k1;
    const double t0 = hB ? k0 : k2, t1 = hB ? k1 : 0.0;
    m0 = __dadd_rn(m0, shfl_xor_f64(t0, D >> 1));
    m1 = __dadd_rn(m1, shfl_xor_f64(t1, D >> 1));
    const int base = hA ? 3 : 0;
    if (LOG2L == 2) {
        v0 = m0;
        v1 = m1;
        s0 = base + (hB ? 2 : 0);
        s1 = hB ? -1 : base + 1;
        return;
    }
    const bool hC = sl & (D >> 2);
    double v = hC ? m1 : m0;
    const double u = hC ? m0 : m1;
    v = __dadd_rn(v, shfl_xor_f64(u, D >> 2));
#pragma unroll
    for (int d = D >> 3; d > 0; d >>= 1) v = __dadd_rn(v, shfl_xor_f64(v, d));
    v0 = v;
    v1 = 0.0;
    s1 = -1;
    s0 = ((sl & ((D >> 2) - 1)) || (hB && hC)) ? -1 : base + (hB ? 2 : (hC ? 1 : 0));
}

// lane -> slot of the 6-double record it writes after reduce_six (or -1)
__device__ __forceinline__ int six_slot(int lane) {
    if (lane & 3) return -1;
    const int g = lane >> 2;  // 0..7: (bit16, bit8, bit4)
    // bit16 selects {a0,a1,a2} / {a3,a4,a5}; inside: g&3 == 0 -> first, 1 -> second (bit4), 2 -> third (bit8), 3 -> none
    const int half = g >> 2, j = g & 3;
    const int idx = j == 0 ? 0 : (j == 1 ? 1 : (j == 2 ? 2 : -1));
    return idx < 0 ? -1 : half * 3 + idx;
}

template <bool GRAD>
__device__ __forceinline__ void write_record(const double (&a)[6], double* rec_n) {
    double2* o = reinterpret_cast<double2*>(rec_n);
    o[0] = make_double2(a[0], a[1]);
    if (GRAD) {
        o[1] = make_double2(a[2], a[3]);
        o[2] = make_double2(a[4], a[5]);
    }
}

template <bool GRAD>
__device__ __forceinline__ void acc4(uint32_t aF, uint32_t aGx, uint32_t aGy, uint32_t aR, double (&a)[6]) {
    const float4 f = lds128(aF), r = lds128(aR);
    float4 gx = make_float4(0.f, 0.f, 0.f, 0.f), gy = gx;
    if (GRAD) {
        gx = lds128(aGx);
        gy = lds128(aGy);
    }
    acc1<GRAD>(f.x, r.x, gx.x, gy.x, a);
    acc1<GRAD>(f.y, r.y, gx.y, gy.y, a);
    acc1<GRAD>(f.z, r.z, gx.z, gy.z, a);
    acc1<GRAD>(f.w, r.w, gx.w, gy.w, a);
}

// FULL whole-warp copies (512 B each) of each of the ROWS rows s[r] to dst + r * rowB
// points a warp projects at a time, given the points its CTA streams in this evaluation
__device__ __forceinline__ int batch_width(int span) { return span > 128 ? 32 : (span > 64 ? 16 : (span > 32 ? 8 : 4)); }

template <int ROWS, int FULL>
__device__ __forceinline__ void copy_rows_full(uint32_t dst, const char* const (&s)[4], uint32_t rowB) {
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
        const uint32_t d = dst + (uint32_t)r * rowB;
        const char* g = s[r];
        if (FULL > 0) cp_async16_off<0>(d, g);
        if (FULL > 1) cp_async16_off<512>(d, g);
        if (FULL > 2) cp_async16_off<1024>(d, g);
        if (FULL > 3) cp_async16_off<1536>(d, g);
        if (FULL > 4) cp_async16_off<2048>(d, g);
        if (FULL > 5) cp_async16_off<2560>(d, g);
        if (FULL > 6) cp_async16_off<3072>(d, g);
        if (FULL > 7) cp_async16_off<3584>(d, g);
    }
}

template <bool GRAD, int LOG2L>
__device__ void stream_staged(const EvalCtx& c, const double* R, const double* t, Shared& sh, RingPos& rp) {
    constexpr int ROWS = GRAD ? 4 : 2;
    constexpr int L = 1 << LOG2L, G = 32 / L;
    constexpr bool SHORT = LOG2L < 5;  // several points per item, at most 2 float4 of a row per lane
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const QueryDev& q = *c.q;
    const int N = c.n1;
    const int CH = c.CH, nCh = c.nCh, depth = c.depth, Cl = c.Cl;
    const uint32_t ring = smem_u32(c.ring) + (uint32_t)(warp * c.ring_per_warp);
    const uint32_t slot_bytes = (uint32_t)c.slot_bytes;
    const int32_t* bcell = &sh.bl_cell[warp][0][0];  // [2][32]
    const int32_t* bn = &sh.bl_n[warp][0][0];
    // batches of BW points owned by this warp: base = n0 + (warp + b * kWarps) * BW.  BW = 32 unless the CTA
    // has so few points (a small task of the scheduler) that narrower batches are needed to occupy every warp.
    const int BW = batch_width(N - c.n0);
    const int first = c.n0 + warp * BW;
    const int nb = first < N ? (N - first + kWarps * BW - 1) / (kWarps * BW) : 0;
    if (nb == 0) return;

    // row base pointers with this lane's 16-byte piece folded in
    const char* baseF = reinterpret_cast<const char*>(q.F + c.c0) + lane * 16;
    const char* baseGx = GRAD ? reinterpret_cast<const char*>(q.Gx + c.c0) + lane * 16 : nullptr;
    const char* baseGy = GRAD ? reinterpret_cast<const char*>(q.Gy + c.c0) + lane * 16 : nullptr;
    const char* baseR = reinterpret_cast<const char*>(q.ref + c.c0) + lane * 16;
    const uint32_t pitch = (uint32_t)q.C * 4u;  // bytes between rows of a map / of ref

    // projects batch b and leaves its supported points compacted in buffer `buf`; returns their count
    auto load_batch = [&](int b, int buf) -> int {
        const int n = first + b * kWarps * BW + lane;
        int cell = -1;
        if (lane < BW && n < N) cell = project_and_flag(c, R, t, n);
        const unsigned mask = __ballot_sync(0xffffffffu, cell >= 0);
        if (cell >= 0) {
            const int rank = __popc(mask & ((1u << lane) - 1u));
            sh.bl_cell[warp][buf][rank] = cell;
            sh.bl_n[warp][buf][rank] = n;
        }
        __syncwarp();
        return __popc(mask);
    };
    auto items_of = [&](int cnt) -> int { return SHORT ? (cnt + G - 1) >> (5 - LOG2L) : cnt * nCh; };

    // copies `nv` float4 of the ROWS rows of point k (buffer buf), starting at channel offset `off`
    // bytes, to shared address dst (rows packed back to back)
    auto copy_point = [&](int buf, int k, uint32_t off, int nv, uint32_t dst) {
        const uint32_t cell = (uint32_t)bcell[buf * 32 + k], n = (uint32_t)bn[buf * 32 + k];
        const char* s[4];
        s[0] = baseF + (size_t)cell * pitch + off;
        s[ROWS - 1] = baseR + (size_t)n * pitch + off;
        if (GRAD) {
            s[1] = baseGx + (size_t)cell * pitch + off;
            s[2] = baseGy + (size_t)cell * pitch + off;
        }
        dst += lane * 16;
        const int full = nv >> 5, tail = nv & 31;  // whole 512-byte warp copies, then a partial one
        if (SHORT) {  // nv <= 32: one 16-byte piece of every row per lane
            if (lane < nv) {
#pragma unroll
                for (int r = 0; r < ROWS; ++r) cp_async16_s(dst + (uint32_t)(r * nv) * 16u, s[r]);
            }
            return;
        }
        if (tail == 0 && full <= 8) {  // the usual case: immediate offsets, no loop
            const uint32_t rowB = (uint32_t)nv * 16u;
            switch (full) {
                case 1: copy_rows_full<ROWS, 1>(dst, s, rowB); break;
                case 2: copy_rows_full<ROWS, 2>(dst, s, rowB); break;
                case 3: copy_rows_full<ROWS, 3>(dst, s, rowB); break;
                case 4: copy_rows_full<ROWS, 4>(dst, s, rowB); break;
                case 5: copy_rows_full<ROWS, 5>(dst, s, rowB); break;
                case 6: copy_rows_full<ROWS, 6>(dst, s, rowB); break;
                case 7: copy_rows_full<ROWS, 7>(dst, s, rowB); break;
                default: copy_rows_full<ROWS, 8>(dst, s, rowB); break;
            }
            return;
        }
#pragma unroll
        for (int r = 0; r < ROWS; ++r) {
            uint32_t d = dst + (uint32_t)(r * nv) * 16u;
            const char* g = s[r];
#pragma unroll 4
            for (int i = 0; i < full; ++i, d += 512, g += 512) cp_async16_s(d, g);
            if (lane < tail) cp_async16_s(d, g);
        }
    };

    int ik = 0, ich = 0;  // issue side: point / chunk of the next item (long rows)
    auto issue = [&](int buf, int cnt, int it, int slot) {
        const uint32_t dst0 = ring + (uint32_t)slot * slot_bytes;
        if (SHORT) {
            const int k0 = it * G, gc = min(G, cnt - k0);
            for (int g = 0; g < gc; ++g) copy_point(buf, k0 + g, 0u, Cl >> 2, dst0 + (uint32_t)(g * ROWS * Cl) * 4u);
        } else {
            const int off = ich * CH;
            copy_point(buf, ik, (uint32_t)off * 4u, min(CH, Cl - off) >> 2, dst0);
            if (++ich == nCh) {
                ich = 0;
                ++ik;
            }
        }
        cp_async_commit();
    };

    double a[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    int ck = 0, cch = 0;  // consume side: point / chunk of the next item (long rows)
    auto consume = [&](int buf, int cnt, int it, int slot, int inflight) {
        cp_async_wait_pending(inflight - 1);  // the oldest of the `inflight` groups has landed
        __syncwarp();
        const uint32_t src0 = ring + (uint32_t)slot * slot_bytes;
        if (SHORT) {
            const int sg = lane >> LOG2L, sl = lane & (L - 1);
            const int k = it * G + sg, nvec = Cl >> 2;
            const bool act = k < cnt;
#pragma unroll
            for (int i = 0; i < 6; ++i) a[i] = 0.0;
            const uint32_t rowB = (uint32_t)Cl * 4u;
            const uint32_t p = src0 + (uint32_t)sg * ROWS * rowB + (uint32_t)sl * 16u;
            if (act && sl < nvec) acc4<GRAD>(p, p + rowB, p + 2 * rowB, p + (ROWS - 1) * rowB, a);
            if (act && sl + L < nvec) {
                const uint32_t p2 = p + L * 16;
                acc4<GRAD>(p2, p2 + rowB, p2 + 2 * rowB, p2 + (ROWS - 1) * rowB, a);
            }
            double* rec_n = c.rec + (size_t)bn[buf * 32 + (act ? k : 0)] * kRecDoubles;
            if (GRAD) {
                constexpr int LG = SHORT ? LOG2L : 4;  // (the long-row instantiation never gets here)
                double v0, v1;
                int s0, s1;
                reduce_group<LG>(a, sl, v0, v1, s0, s1);
                if (act && s0 >= 0) rec_n[s0] = v0;
                if (LG == 2 && act && s1 >= 0) rec_n[s1] = v1;
            } else {
                reduce_lanes<GRAD, LOG2L>(a);
                if (act && sl == 0) rec_n[0] = a[0];
            }
        } else {
            if (cch == 0) {
#pragma unroll
                for (int i = 0; i < 6; ++i) a[i] = 0.0;
            }
            const int nv = min(CH, Cl - cch * CH) >> 2;
            const uint32_t rowB = (uint32_t)nv * 16u;
            uint32_t p = src0 + (uint32_t)lane * 16u;
#pragma unroll 2
            for (int v = lane; v < nv; v += 32, p += 512) acc4<GRAD>(p, p + rowB, p + 2 * rowB, p + (ROWS - 1) * rowB, a);
            if (++cch == nCh) {
                cch = 0;
                double* rec_n = c.rec + (size_t)bn[buf * 32 + ck] * kRecDoubles;
                if (GRAD) {
                    const double v = reduce_six(a, lane);
                    const int slot = six_slot(lane);
                    if (slot >= 0) rec_n[slot] = v;
                } else {
                    reduce_lanes<GRAD, 5>(a);
                    if (lane == 0) rec_n[0] = a[0];
                }
                ++ck;
            }
        }
        __syncwarp();  // every lane is done with the slot: it may be refilled
    };

    int cnt0 = load_batch(0, 0), cnt1 = 0;
    int ib = 0, ii = 0, cb = 0, ci = 0, inflight = 0;
    int items_i = items_of(cnt0), items_c = items_i;
    for (;;) {
        // ---- issue side: keep `depth` items in flight, at most one batch ahead of the consume side
        for (;;) {
            if (ii < items_i) {
                if (inflight == depth) break;
                int slot = rp.pos + inflight;
                if (slot >= depth) slot -= depth;
                issue(ib & 1, (ib & 1) ? cnt1 : cnt0, ii, slot);
                ++ii;
                ++inflight;
            } else {
                if (ib + 1 >= nb || ib > cb) break;
                ++ib;
                ii = 0;
                ik = 0;
                ich = 0;
                const int cnew = load_batch(ib, ib & 1);
                if (ib & 1)
                    cnt1 = cnew;
                else
                    cnt0 = cnew;
                items_i = items_of(cnew);
            }
        }
        // ---- consume side
        if (ci < items_c) {
            consume(cb & 1, (cb & 1) ? cnt1 : cnt0, ci, rp.pos, inflight);
            if (++rp.pos == depth) rp.pos = 0;
            ++ci;
            --inflight;
        } else {
            if (cb == ib) break;  // the issue side could not advance: that was the last batch
            ++cb;
            ci = 0;
            ck = 0;
            cch = 0;
            items_c = items_of((cb & 1) ? cnt1 : cnt0);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Streaming phase: one record per supported point.  Every warp is its own producer and consumer.
// ---------------------------------------------------------------------------------------------
template <bool GRAD>
__device__ void stream_points(const EvalCtx& c, const double* R, const double* t, Shared& sh, RingPos& rp) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const QueryDev& q = *c.q;
    const int N = c.n1;

    if (!c.aligned) {
        // ---------------- generic path: every warp projects its own 32 points, loads directly
        const int BW = batch_width(N - c.n0);
        for (int base = c.n0 + warp * BW; base < N; base += kWarps * BW) {
            const int n = base + lane;
            int cell = -1;
            if (lane < BW && n < N) cell = project_and_flag(c, R, t, n);
            for (int j = 0; j < BW; ++j) {
                const int cj = __shfl_sync(0xffffffffu, cell, j);
                if (cj < 0) continue;
                const int nj = base + j;
                const float* srcF = q.F + (size_t)cj * q.C + c.c0;
                const float* srcGx = GRAD ? q.Gx + (size_t)cj * q.C + c.c0 : nullptr;
                const float* srcGy = GRAD ? q.Gy + (size_t)cj * q.C + c.c0 : nullptr;
                const float* srcR = q.ref + (size_t)nj * q.C + c.c0;
                double a[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
                for (int ch = lane; ch < c.Cl; ch += 32) {
                    const float f = __ldg(srcF + ch), r = __ldg(srcR + ch);
                    float gx = 0.f, gy = 0.f;
                    if (GRAD) {
                        gx = __ldg(srcGx + ch);
                        gy = __ldg(srcGy + ch);
                    }
                    acc1<GRAD>(f, r, gx, gy, a);
                }
                reduce_lanes<GRAD, 5>(a);
                if (lane == 0) write_record<GRAD>(a, c.rec + (size_t)nj * kRecDoubles);
            }
        }
        return;
    }

    switch (c.L) {
        case 4: stream_staged<GRAD, 2>(c, R, t, sh, rp); break;
        case 8: stream_staged<GRAD, 3>(c, R, t, sh, rp); break;
        case 16: stream_staged<GRAD, 4>(c, R, t, sh, rp); break;
        default: stream_staged<GRAD, 5>(c, R, t, sh, rp); break;
    }
}

// ---------------------------------------------------------------------------------------------
// Deterministic reductions.  The points of a query are cut into kRanges = 128 fixed ranges (the same
// for every launch geometry).  A range is summed by ONE warp in a fixed order (lanes stride the range,
// then a transposition through the warp's idle slice of the staging ring and a pairwise tree over the
// 32 lanes); range totals are added pairwise along the aligned binary tree over the range index: first
// over the consecutive ranges a warp owns, then over the warps of the CTA, then over the CTAs of the
// cluster or the tasks of the evaluation.  Every owner holds an aligned power-of-two run of ranges, so
// the result is bit-identical for every cluster size, task count, batch size and ring geometry.
// ---------------------------------------------------------------------------------------------
// (kLaneStride, kWarpScratchDoubles: fmpnp_kernels.cuh -- the host sizes the ring so that a warp's slice holds them)

// Combines the per-CTA values every CTA of the cluster left in its own Shared::cpart[par] /
// cpart_i[par]: all CTAs read all partials through distributed shared memory and add them in
// the same pairwise order, so every CTA ends up with bit-identical totals (and then takes the
// same LM decisions).  One hardware cluster barrier per reduction; the two halves of cpart
// alternate, so a half is only rewritten after the next barrier, when every CTA has read it.
template <bool MAX>
__device__ __forceinline__ void cluster_combine(Shared& sh, int csize, int32_t& par, int nv, double* out, int ni,
                                                int32_t* iout) {
    cg::cluster_group cl = cg::this_cluster();
    cl.sync();
    const int tid = threadIdx.x;
    if (tid < nv) {
        double v[kMaxCluster];
#pragma unroll
        for (int r = 0; r < kMaxCluster; ++r) v[r] = r < csize ? *cl.map_shared_rank(&sh.cpart[par][tid], r) : 0.0;
#pragma unroll
        for (int w = 1; w < kMaxCluster; w <<= 1) {
            if (w < csize) {
#pragma unroll
                for (int i = 0; i + w < kMaxCluster; i += 2 * w) v[i] = MAX ? fmax(v[i], v[i + w]) : __dadd_rn(v[i], v[i + w]);
            }
        }
        out[tid] = v[0];
    } else if (tid >= 32 && tid < 32 + ni) {
        const int k = tid - 32;
        int32_t s = 0;
        for (int r = 0; r < csize; ++r) {
            const int32_t x = *cl.map_shared_rank(&sh.cpart_i[par][k], r);
            s = MAX ? (s | x) : (s + x);
        }
        iout[k] = s;
    }
    par ^= 1;
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// Per-point phase: robust loss, ratio test, contraction with the projection Jacobian, sums.
// Leaves sh.ev[0..27], sh.ev_count (survivors), sh.ev_nsup (supported) and sh.ev_cost (mean rho).
// ---------------------------------------------------------------------------------------------
// max |rho| over the supported points [c.n0, c.n1) of this CTA and whether any is NaN (torch.max propagates NaN):
// the ingredients of ratio_threshold_feature_errors' limit (model.py:120-129).  Valid in thread 0 after the call.
__device__ __forceinline__ void block_ratio_max(const EvalCtx& c, Shared& sh, double& mm, int& nf) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    double m = -INFINITY;
    int nanf = 0;
    for (int n = c.n0 + tid; n < c.n1; n += kThreads) {
        if (!__ldcg(c.supp + n)) continue;  // .cg: records may have been written by another CTA (task scheduler)
        const double x = __dmul_rn(0.5, __ldcg(c.rec + (size_t)n * kRecDoubles));
        double rho = x, w = 1.0;
        if (c.ratio_mode == 1) robust_loss(c.loss_id, c.alpha, x, rho, w);
        const double a = fabs(rho);
        if (a != a) nanf = 1;
        m = fmax(m, a);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        m = fmax(m, shfl_xor_f64(m, off));
        nanf |= __shfl_xor_sync(0xffffffffu, nanf, off);
    }
    if (lane == 0) {
        sh.red[warp][0] = m;
        sh.red_i[warp][0] = nanf;
    }
    __syncthreads();
    mm = -INFINITY;
    nf = 0;
    if (tid == 0) {
        mm = sh.red[0][0];
        nf = sh.red_i[0][0];
        for (int w = 1; w < kWarps; ++w) {
            mm = fmax(mm, sh.red[w][0]);
            nf |= sh.red_i[w][0];
        }
    }
}

template <bool GRAD>
__device__ void reduce_points(const EvalCtx& c, const double* R, const double* t, Shared& sh, RingPos& rp) {
    const QueryDev& q = *c.q;
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const bool use_ratio = c.ratio_mode != 0;

    if (use_ratio && !c.limit_given) {
        double mm;
        int nf;
        block_ratio_max(c, sh, mm, nf);
        if (tid == 0) {
            if (c.csize > 1) {
                sh.cpart[rp.cpar][0] = mm;
                sh.cpart_i[rp.cpar][0] = nf;
            } else {
                sh.limit = nf ? NAN : __dmul_rn(mm, c.thr);  // torch.max propagates NaN
            }
        }
        if (c.csize > 1) {
            cluster_combine<true>(sh, c.csize, rp.cpar, 1, &sh.climit, 1, &sh.climit_nan);
            if (tid == 0) sh.limit = sh.climit_nan ? NAN : __dmul_rn(sh.climit, c.thr);
        }
        __syncthreads();
    }
    const double limit = use_ratio ? sh.limit : 0.0;

    const double fx = q.K[0], fy = q.K[4];
    const int per = c.range_len;              // points per fixed range
    const int nr = kRanges / c.csize;         // ranges this CTA owns: an aligned power-of-two run
    const int r_first = c.crank * nr;
    const int rpw = nr >= kWarps ? nr / kWarps : 1;  // consecutive ranges per warp ...
    const int wact = nr >= kWarps ? kWarps : nr;     // ... of the warps that own any
    double* const wscr = reinterpret_cast<double*>(c.ring + (size_t)warp * c.ring_per_warp);  // idle ring slice
    double* const stack = wscr + 28 * kLaneStride;
    int32_t cnt[2] = {0, 0};  // survivors, supported (this lane, all its ranges)
    if (warp < wact) {
        for (int rr = 0; rr < rpw; ++rr) {
            const int r0 = min(q.N, (r_first + warp * rpw + rr) * per), r1 = min(q.N, r0 + per);
            double acc[28];
#pragma unroll
            for (int i = 0; i < 28; ++i) acc[i] = 0.0;
            for (int n = r0 + lane; n < r1; n += 32) {
                const bool sup = __ldcg(c.supp + n) != 0;
                bool keep = false;
                if (sup) {
                    ++cnt[1];
                    const double2* r = reinterpret_cast<const double2*>(c.rec + (size_t)n * kRecDoubles);
                    const double2 rc0 = __ldcg(r);
                    const double x = __dmul_rn(0.5, rc0.x);  // cost = 0.5 * sum e^2 (model.py:338)
                    double rho = x, w = 1.0;
                    if (c.ratio_mode != 2) robust_loss(c.loss_id, c.alpha, x, rho, w);
                    keep = !use_ratio || (fabs(rho) < limit);
                    if (keep) {
                        ++cnt[0];
                        acc[27] = __dadd_rn(acc[27], rho);
                        if (GRAD) {
                            const double2 rc1 = __ldcg(r + 1), rc2 = __ldcg(r + 2);
                            const Proj pr = project_point(R, t, q.K, q.pts[3 * n], q.pts[3 * n + 1], q.pts[3 * n + 2],
                                                          q.im_w, q.im_h);
                            contract_point(fx, fy, pr.X, pr.Y, pr.Z, w, rc0.y, rc1.x, rc1.y, rc2.x, rc2.y, acc);
                        }
                    }
                }
                if (c.h_kept) c.h_kept[n] = keep ? 1 : 0;
                if (c.inl_out) c.inl_out[n] = keep ? 1 : 0;
            }
            // the 32 lane partials of each of the 28 sums -> lane i (transposition), pairwise tree over the lanes
            __syncwarp();  // the rows of the previous range have been read
#pragma unroll
            for (int i = 0; i < 28; ++i) wscr[i * kLaneStride + lane] = acc[i];
            __syncwarp();
            if (lane < 28) {
                const double* row = wscr + lane * kLaneStride;
                double x[16];
#pragma unroll
                for (int k = 0; k < 16; ++k) x[k] = __dadd_rn(row[2 * k], row[2 * k + 1]);
#pragma unroll
                for (int w = 1; w < 16; w <<= 1) {
#pragma unroll
                    for (int k = 0; k + w < 16; k += 2 * w) x[k] = __dadd_rn(x[k], x[k + w]);
                }
                double tot = x[0];
                int level = 0;
                for (int kk = rr; kk & 1; kk >>= 1) {  // binary-counter stack: pairwise tree over this warp's ranges
                    tot = __dadd_rn(stack[level * 28 + lane], tot);
                    ++level;
                }
                stack[level * 28 + lane] = tot;
                if (rr == rpw - 1) sh.wtot[warp][lane] = tot;  // rpw is a power of two: the total is on top
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) cnt[i] += __shfl_xor_sync(0xffffffffu, cnt[i], off);
    }
    if (lane == 0) {
        sh.red_i[warp][0] = cnt[0];
        sh.red_i[warp][1] = cnt[1];
    }
    __syncthreads();
    int32_t* iout = &sh.ev_count;  // ev_count, ev_nsup are adjacent
    if (tid < 28) {  // pairwise tree over the warps that own ranges
        double v[kWarps];
#pragma unroll
        for (int w = 0; w < kWarps; ++w) v[w] = w < wact ? sh.wtot[w][tid] : 0.0;
#pragma unroll
        for (int w = 1; w < kWarps; w <<= 1) {
            if (w < wact) {
#pragma unroll
                for (int k = 0; k + w < kWarps; k += 2 * w) v[k] = __dadd_rn(v[k], v[k + w]);
            }
        }
        if (c.part_out)
            c.part_out[tid] = v[0];
        else if (c.csize > 1)
            sh.cpart[rp.cpar][tid] = v[0];
        else
            sh.ev[tid] = v[0];
    } else if (tid >= 32 && tid < 34) {
        int32_t tot_i = 0;
        for (int w = 0; w < kWarps; ++w) tot_i += sh.red_i[w][tid - 32];
        if (c.part_out)
            reinterpret_cast<long long*>(c.part_out)[28 + (tid - 32)] = tot_i;
        else if (c.csize > 1)
            sh.cpart_i[rp.cpar][tid - 32] = tot_i;
        else
            iout[tid - 32] = tot_i;
    }
    if (c.part_out) {  // the CTA that completes the evaluation combines the tasks' partials (combine_parts)
        __syncthreads();
        return;
    }
    if (c.csize > 1)
        cluster_combine<false>(sh, c.csize, rp.cpar, 28, sh.ev, 2, iout);
    else
        __syncthreads();
    if (tid == 0) sh.ev_cost = sh.ev[27] / (double)sh.ev_count;  // mean; 0/0 = NaN like torch's empty mean
    __syncthreads();
}

template <bool GRAD>
__device__ __forceinline__ void evaluate(EvalCtx& c, const double* R, const double* t, Shared& sh, RingPos& rp) {
    set_geometry(c, GRAD ? 4 : 2);
    rp.pos = 0;  // the ring is empty between evaluations; its depth may change with the slice
    stream_points<GRAD>(c, R, t, sh, rp);
    __syncthreads();
    reduce_points<GRAD>(c, R, t, sh, rp);
    if (threadIdx.x == 0) {
        sh.n_evals += 1;
        sh.gathered += (long long)sh.ev_nsup * c.Cl * (GRAD ? 4 : 2);
    }
}

__device__ __forceinline__ bool slice_aligned(const QueryDev& q, int c0, int Cl, bool grad) {
    uintptr_t bits = reinterpret_cast<uintptr_t>(q.F) | reinterpret_cast<uintptr_t>(q.ref);
    if (grad) bits |= reinterpret_cast<uintptr_t>(q.Gx) | reinterpret_cast<uintptr_t>(q.Gy);
    return (bits & 15u) == 0 && (q.C & 3) == 0 && (c0 & 3) == 0 && (Cl & 3) == 0;
}

__device__ __forceinline__ int tri_index(int i, int j) {  // upper triangle, row-major, i <= j
    return i * 6 - (i * (i - 1)) / 2 + (j - i);
}

// optimizer_step (model.py:37-72) + pose update (model.py:416-426) on warp 0.
// Reads sh.lin / sh.lambda / sh.lr / sh.R / sh.t; writes sh.Rn / sh.tn, sh.flag = 1 on a NaN step.
__device__ void solve_and_update(Shared& sh, int solver, int lane) {
    auto build = [&]() {
        for (int e = lane; e < 36; e += 32) {
            const int i = e / 6, j = e % 6;
            double v = sh.lin[i <= j ? tri_index(i, j) : tri_index(j, i)];
            if (i == j && sh.lambda != 0.0)  // `if lambda_:` D = (diag(H) + 1e-9); H = H + D * lambda_
                v = __dadd_rn(v, __dmul_rn(__dadd_rn(v, 1e-9), sh.lambda));
            sh.A[e] = v;
        }
        if (lane < 6) sh.b[lane] = sh.lin[21 + lane];
        __syncwarp();
    };
    build();
    bool ok = false;
    if (solver == FMPNP_SOLVER_CHOLESKY) ok = warp_cholesky_solve(sh.A, sh.b, lane);
    if (!ok) {
        build();
        warp_lu_solve(sh.A, sh.b, lane);
    }
    if (lane == 0) {
        double delta[6];
        bool has_nan = false;
        for (int i = 0; i < 6; ++i) {
            delta[i] = __dmul_rn(-sh.lr, sh.b[i]);  // delta = -lr * x   (model.py:62)
            if (delta[i] != delta[i]) has_nan = true;
        }
        if (has_nan) {
            sh.flag = 1;  // model.py:411-413 (intended `break`)
        } else {
            double dr[9];
            so3exp(delta + 3, dr);                               // dw = delta[3:6]
            compose_pose(dr, delta, sh.R, sh.t, sh.Rn, sh.tn);   // dt = delta[0:3]
            sh.flag = 0;
        }
    }
}

__device__ __forceinline__ void hist_record(const LaunchParams& p, int qi, int e, const double* R, const double* t,
                                            double cost) {
    if (!p.hist.enabled || e >= p.hist.cap || (p.cluster_size > 1 && cg::this_cluster().block_rank() != 0)) return;
    const size_t o = (size_t)qi * p.hist.cap + e;
    for (int i = 0; i < 9; ++i) p.hist.Rs[o * 9 + i] = R[i];
    for (int i = 0; i < 3; ++i) p.hist.ts[o * 3 + i] = t[i];
    p.hist.costs[o] = cost;
}

__device__ __forceinline__ void hist_slices(const LaunchParams& p, int qi, int e, EvalCtx& c) {
    if (p.hist.enabled && e < p.hist.cap) {
        const size_t o = ((size_t)qi * p.hist.cap + e) * p.hist.n_stride;
        c.h_px = p.hist.points2d + 2 * o;
        c.h_mask = p.hist.mask + o;
        c.h_kept = p.hist.kept + o;
    } else {
        c.h_px = nullptr;
        c.h_mask = nullptr;
        c.h_kept = nullptr;
    }
}

template <int MODE>
__device__ __forceinline__ void persistent_body(const LaunchParams& p) {
    extern __shared__ __align__(128) unsigned char ring[];
    __shared__ Shared sh;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    // A query is refined by a thread-block cluster of `csize` CTAs (1 when the batch alone fills the
    // GPU): each CTA streams a contiguous share of the points, the partial normal equations are
    // combined through distributed shared memory, and every CTA then takes the same LM decisions.
    const int csize = p.cluster_size, crank = csize > 1 ? (int)cg::this_cluster().block_rank() : 0;
    if (csize > 1) cg::this_cluster().sync();  // every CTA of the cluster is resident before any remote access
    RingPos rp;
    rp.pos = 0;
    rp.cpar = 0;
    double* rec = p.rec + (size_t)blockIdx.x * p.n_max * kRecDoubles;
    uint8_t* supp = p.supp + (size_t)blockIdx.x * p.n_max;

    for (;;) {
        __syncthreads();
        if (csize > 1) {
            cg::cluster_group cl = cg::this_cluster();
            if (crank == 0 && tid == 0) {
                const int32_t nq = (int32_t)atomicAdd(p.work_counter, 1u);
                for (int r = 0; r < csize; ++r) *cl.map_shared_rank(&sh.query, r) = nq;
            }
            cl.sync();
        } else {
            if (tid == 0) sh.query = (int32_t)atomicAdd(p.work_counter, 1u);
            __syncthreads();
        }
        const int qi = sh.query;
        if (qi >= p.B) break;
        const QueryDev& q = p.queries[qi];

        EvalCtx c;
        c.q = &q;
        c.rec = rec;
        c.supp = supp;
        c.ring = ring;
        c.ring_per_warp = p.ring_per_warp;
        c.min_depth = p.min_depth;
        c.csize = csize;
        c.crank = crank;
        c.part_out = nullptr;
        c.limit_given = false;
        {
            const int per = max(1, (q.N + kRanges - 1) / kRanges);
            c.range_len = per;
            c.n0 = min(q.N, crank * (kRanges / csize) * per);
            c.n1 = min(q.N, (crank + 1) * (kRanges / csize) * per);
        }
        c.h_px = nullptr;
        c.h_mask = nullptr;
        c.h_kept = nullptr;
        c.inl_out = nullptr;

        if (tid == 0) {
            for (int i = 0; i < 9; ++i) sh.R[i] = q.R0[i];
            for (int i = 0; i < 3; ++i) sh.t[i] = q.t0[i];
            sh.n_evals = 0;
            sh.gathered = 0;
        }
        __syncthreads();

        if (MODE == kModeComputeCost) {  // sparseFeaturePnP.compute_cost (model.py:216-243)
            c.c0 = 0;
            c.Cl = q.C;
            c.aligned = slice_aligned(q, 0, q.C, false);
            c.loss_id = FMPNP_LOSS_SQUARED;
            c.alpha = 0.0;
            c.ratio_mode = p.use_ratio ? 2 : 0;
            c.thr = p.ratio_thr;
            evaluate<false>(c, sh.R, sh.t, sh, rp);
            __syncthreads();
            if (tid == 0 && crank == 0) {
                p.cost_out[qi] = sh.ev_nsup > 0 ? sh.ev_cost : NAN;
                if (p.nsup_out) p.nsup_out[qi] = sh.ev_nsup;
            }
            continue;
        }
        if (MODE == kModeFindInliers) {  // find_inliers (model.py:131-152)
            c.c0 = 0;
            c.Cl = q.C;
            c.aligned = slice_aligned(q, 0, q.C, false);
            c.loss_id = p.loss_id;
            c.alpha = p.alpha;
            c.ratio_mode = 1;
            c.thr = p.ratio_thr;
            c.inl_out = p.inliers_out + (size_t)qi * p.inliers_stride;
            evaluate<false>(c, sh.R, sh.t, sh, rp);
            continue;
        }

        if (MODE != kModeRefine) continue;
        // ------------------------------------------------------------------ refine
        if (tid == 0) {
            sh.status = FMPNP_STATUS_OK;
            sh.has_best = 0;
            sh.best_cost = NAN;
            sh.best_n = -1;
            sh.hist_len = 0;
            sh.initial_set = 0;
            sh.initial_cost = NAN;
            for (int i = 0; i < FMPNP_MAX_LEVELS; ++i) sh.level_status[i] = 0;
        }
        __syncthreads();
        bool run_levels = true;
        if (p.n_levels > 0) {
            // multilevel_optimization: initial_cost_ = compute_cost(...) on ALL channels (model.py:182)
            c.c0 = 0;
            c.Cl = q.C;
            c.aligned = slice_aligned(q, 0, q.C, false);
            c.loss_id = FMPNP_LOSS_SQUARED;
            c.alpha = 0.0;
            c.ratio_mode = p.use_ratio ? 2 : 0;
            c.thr = p.ratio_thr;
            evaluate<false>(c, sh.R, sh.t, sh, rp);
            __syncthreads();
            if (sh.ev_nsup == 0) {  // compute_cost is None -> return the initial pose (model.py:183-187)
                run_levels = false;
                if (tid == 0) sh.status = FMPNP_STATUS_NO_INITIAL_COST;
            } else if (tid == 0) {
                sh.initial_cost = sh.ev_cost;
                sh.initial_set = 1;
            }
            __syncthreads();
        }
        const int nlev = p.n_levels > 0 ? p.n_levels : 1;
        for (int lvl = 0; run_levels && lvl < nlev; ++lvl) {
            c.c0 = p.n_levels > 0 ? p.levels[lvl][0] : 0;
            c.Cl = p.n_levels > 0 ? p.levels[lvl][1] - p.levels[lvl][0] : q.C;
            c.aligned = slice_aligned(q, c.c0, c.Cl, true);
            c.loss_id = p.loss_id;
            c.alpha = p.alpha;
            c.ratio_mode = p.use_ratio ? 1 : 0;
            c.thr = p.ratio_thr;

            // ---- i == 0: cost and linearisation at the incoming pose (model.py:303-359)
            hist_slices(p, qi, sh.hist_len, c);
            evaluate<true>(c, sh.R, sh.t, sh, rp);
            __syncthreads();
            if (sh.ev_nsup == 0) {  // model.py:316-320 -> forward returns the CURRENT pose
                if (tid == 0) {
                    sh.status = FMPNP_STATUS_EMPTY_SUPPORT;
                    sh.level_status[lvl] = FMPNP_STATUS_EMPTY_SUPPORT;
                }
                __syncthreads();
                continue;
            }
            if (tid == 0) {
                sh.prev_cost = sh.ev_cost;
                sh.best_cost = sh.ev_cost;
                sh.best_n = sh.ev_count;
                sh.has_best = 1;
                for (int i = 0; i < 9; ++i) sh.Rb[i] = sh.R[i];
                for (int i = 0; i < 3; ++i) sh.tb[i] = sh.t[i];
                if (!sh.initial_set) {  // `if self.initial_cost_ is None` (model.py:355-356)
                    sh.initial_cost = sh.ev_cost;
                    sh.initial_set = 1;
                }
                for (int i = 0; i < 27; ++i) sh.lin[i] = sh.ev[i];
                sh.lambda = p.lambda0;  // model.py:281
                sh.lr = 1.0;
                sh.status = FMPNP_STATUS_OK;
                sh.flag = 0;
                hist_record(p, qi, sh.hist_len, sh.R, sh.t, sh.ev_cost);
                sh.hist_len += 1;
            }
            __syncthreads();

            for (int it = 0; it < p.n_iters; ++it) {
                if (warp == 0) solve_and_update(sh, p.solver, lane);
                __syncthreads();
                if (sh.flag == 1) {
                    if (tid == 0) sh.status = FMPNP_STATUS_NAN_STEP;
                    break;
                }
                // ---- trial pose: cost (model.py:428-462) and, if accepted, the next linearisation
                hist_slices(p, qi, sh.hist_len, c);
                evaluate<true>(c, sh.Rn, sh.tn, sh, rp);
                __syncthreads();
                if (sh.ev_nsup == 0) {  // model.py:441-445 -> CURRENT pose
                    if (tid == 0) {
                        sh.status = FMPNP_STATUS_EMPTY_SUPPORT;
                        sh.flag = 2;
                    }
                    break;
                }
                if (tid == 0) {
                    const double new_cost = sh.ev_cost;
                    hist_record(p, qi, sh.hist_len, sh.Rn, sh.tn, new_cost);  // model.py:464-465
                    sh.hist_len += 1;
                    const bool worse = new_cost > sh.prev_cost;
                    {  // model.py:469-470
                        const double l = __dmul_rn(sh.lambda, worse ? 10.0 : 0.1);
                        sh.lambda = l < 1e-6 ? 1e-6 : (l > 1e4 ? 1e4 : l);
                    }
                    if (worse) {  // model.py:472-476: reject, keep (R, t) and its linearisation
                        const double l = __dmul_rn(0.1, sh.lr);
                        sh.lr = l < 1e-3 ? 1e-3 : (l > 1.0 ? 1.0 : l);
                    } else {  // model.py:477-486: accept
                        sh.lr = 1.0;
                        if (new_cost < sh.best_cost) {
                            for (int i = 0; i < 9; ++i) sh.Rb[i] = sh.Rn[i];
                            for (int i = 0; i < 3; ++i) sh.tb[i] = sh.tn[i];
                            sh.best_n = sh.ev_count;
                            sh.best_cost = new_cost;
                        }
                        sh.prev_cost = new_cost;
                        for (int i = 0; i < 9; ++i) sh.R[i] = sh.Rn[i];
                        for (int i = 0; i < 3; ++i) sh.t[i] = sh.tn[i];
                        for (int i = 0; i < 27; ++i) sh.lin[i] = sh.ev[i];
                    }
                }
                __syncthreads();
            }
            __syncthreads();
            if (tid == 0) {
                if (sh.flag != 2) {  // model.py:491-494: best pose (else: current pose, unchanged)
                    for (int i = 0; i < 9; ++i) sh.R[i] = sh.Rb[i];
                    for (int i = 0; i < 3; ++i) sh.t[i] = sh.tb[i];
                }
                sh.level_status[lvl] = sh.status;
                sh.flag = 0;
            }
            __syncthreads();
        }
        if (tid == 0 && crank == 0) {
            fmpnp_result& r = p.results[qi];
            for (int i = 0; i < 9; ++i) r.R[i] = sh.R[i];
            for (int i = 0; i < 3; ++i) r.t[i] = sh.t[i];
            r.best_cost = sh.best_cost;
            r.initial_cost = sh.initial_cost;
            r.best_num_inliers = sh.best_n;
            r.status = sh.status;
            r.n_evaluations = sh.n_evals;
            for (int i = 0; i < FMPNP_MAX_LEVELS; ++i) r.level_status[i] = sh.level_status[i];
            r.gathered_rows = sh.gathered;
            if (p.hist.enabled) p.hist.len[qi] = sh.hist_len;
        }
    }
}

// Three entry points over the same persistent body, so profilers see them by name.
__global__ void __launch_bounds__(kThreads, 2) refine_kernel(const LaunchParams p) { persistent_body<kModeRefine>(p); }
__global__ void __launch_bounds__(kThreads, 2) compute_cost_kernel(const LaunchParams p) {
    persistent_body<kModeComputeCost>(p);
}
__global__ void __launch_bounds__(kThreads, 2) find_inliers_kernel(const LaunchParams p) {
    persistent_body<kModeFindInliers>(p);
}

// ---------------------------------------------------------------------------------------------
// Task scheduler (LaunchParams::scheduler == 1).
// One evaluation of a query = `units` tasks (1..128), each an aligned run of the 128 fixed point ranges.  Tasks
// sit in one global FIFO (ticket queue: a consumer takes a ticket and waits for that position to be filled).
// A task streams its points and sums its own ranges (28 + 2 partials in global memory); the CTA that completes
// the last task of an evaluation adds the partials along the same fixed pairwise tree (bit-identical to the
// one-CTA-per-query path), runs the solve / accept / reject step on the query's state (kept in global memory
// between evaluations) and queues the tasks of the next evaluation.  With a ratio test an evaluation takes two
// rounds of tasks (max |rho| first, sums second).  Any number of queries keeps every CTA busy: no waves, no
// tail, no idle slots; a single query is spread over up to 128 CTAs.
// ---------------------------------------------------------------------------------------------
constexpr unsigned long long kQueueEmpty = ~0ull;
constexpr int kPartStride = 32;  // doubles per task partial: 28 sums, 2 counts (as int64), max |rho|, NaN flag (int64)
constexpr unsigned long long kWaitLimitNs = 20ull * 1000 * 1000 * 1000;  // then the kernel gives up instead of hanging

__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    return *reinterpret_cast<const volatile unsigned long long*>(p);
}
__device__ __forceinline__ unsigned int ld_volatile_u32(const unsigned int* p) {
    return *reinterpret_cast<const volatile unsigned int*>(p);
}

// thread 0 only: queues the `units` tasks of query q's next evaluation (its state is already in global memory)
__device__ __forceinline__ void enqueue_evaluation(const SchedDev& s, int q) {
    __threadfence();  // state (and, for the finisher, everything before) visible before the tasks are
    const unsigned int base = atomicAdd(s.tail, (unsigned int)s.units);
    for (int u = 0; u < s.units; ++u) {
        const unsigned int pos = base + (unsigned int)u;
        const unsigned long long v = ((unsigned long long)pos << 32) | (unsigned int)(q * kRanges + u + 1);
        *reinterpret_cast<volatile unsigned long long*>(&s.queue[pos & (s.cap - 1)]) = v;
    }
}

// thread 0 only: next task id (q * kRanges + u), or -1 when every query is finished (or on a protocol error)
__device__ __forceinline__ int dequeue_task(const SchedDev& s, int B) {
    const unsigned int ticket = atomicAdd(s.head, 1u);
    const unsigned long long* slot = &s.queue[ticket & (s.cap - 1)];
    unsigned long long t0 = 0;
    for (uint32_t spin = 0;; ++spin) {
        const unsigned long long v = ld_volatile_u64(slot);
        if (v != kQueueEmpty && (unsigned int)(v >> 32) == ticket) {
            __threadfence();
            return (int)(unsigned int)(v & 0xffffffffu) - 1;
        }
        if (ld_volatile_u32(s.done) >= (unsigned int)B || ld_volatile_u32(s.error) != 0u) return -1;
        if ((spin & 1023u) == 1023u) {
            const unsigned long long now = global_timer_ns();
            if (t0 == 0) t0 = now;
            if (now - t0 > kWaitLimitNs) {
                atomicExch(s.error, 1u);
                return -1;
            }
        }
        __nanosleep(200);
    }
}

// thread 0: query state global <-> shared
__device__ __forceinline__ void load_state(Shared& sh, const volatile QueryState& g) {  // volatile: L1 may hold a stale copy
    for (int i = 0; i < 9; ++i) {
        sh.R[i] = g.R[i];
        sh.Rn[i] = g.Rn[i];
        sh.Rb[i] = g.Rb[i];
    }
    for (int i = 0; i < 3; ++i) {
        sh.t[i] = g.t[i];
        sh.tn[i] = g.tn[i];
        sh.tb[i] = g.tb[i];
    }
    for (int i = 0; i < 27; ++i) sh.lin[i] = g.lin[i];
    sh.prev_cost = g.prev_cost;
    sh.best_cost = g.best_cost;
    sh.lambda = g.lambda;
    sh.lr = g.lr;
    sh.initial_cost = g.initial_cost;
    sh.gathered = g.gathered;
    sh.best_n = g.best_n;
    sh.has_best = g.has_best;
    sh.status = g.status;
    sh.flag = g.flag;
    sh.hist_len = g.hist_len;
    sh.n_evals = g.n_evals;
    sh.initial_set = g.initial_set;
    for (int i = 0; i < FMPNP_MAX_LEVELS; ++i) sh.level_status[i] = g.level_status[i];
}

__device__ __forceinline__ void store_state(const Shared& sh, QueryState& g) {
    for (int i = 0; i < 9; ++i) {
        g.R[i] = sh.R[i];
        g.Rn[i] = sh.Rn[i];
        g.Rb[i] = sh.Rb[i];
    }
    for (int i = 0; i < 3; ++i) {
        g.t[i] = sh.t[i];
        g.tn[i] = sh.tn[i];
        g.tb[i] = sh.tb[i];
    }
    for (int i = 0; i < 27; ++i) g.lin[i] = sh.lin[i];
    g.prev_cost = sh.prev_cost;
    g.best_cost = sh.best_cost;
    g.lambda = sh.lambda;
    g.lr = sh.lr;
    g.initial_cost = sh.initial_cost;
    g.gathered = sh.gathered;
    g.best_n = sh.best_n;
    g.has_best = sh.has_best;
    g.status = sh.status;
    g.flag = sh.flag;
    g.hist_len = sh.hist_len;
    g.n_evals = sh.n_evals;
    g.initial_set = sh.initial_set;
    for (int i = 0; i < FMPNP_MAX_LEVELS; ++i) g.level_status[i] = sh.level_status[i];
}

// describes the next evaluation of a query in its global state (thread 0)
__device__ __forceinline__ void describe_evaluation(QueryState& g, const double* R, const double* t, int phase, int grad,
                                                    int c0, int Cl, int ratio_mode, int loss, int hist_entry, int units) {
    for (int i = 0; i < 9; ++i) g.evR[i] = R[i];
    for (int i = 0; i < 3; ++i) g.evt[i] = t[i];
    g.phase = phase;
    g.ev_grad = grad;
    g.ev_c0 = c0;
    g.ev_Cl = Cl;
    g.ev_ratio_mode = ratio_mode;
    g.ev_loss = loss;
    g.ev_hist = hist_entry;
    g.ev_round = 0;
    g.pending = units;
}

// Adds the partial sums the `units` (a power of two, up to 128) tasks of an evaluation left in global memory
// along the same aligned pairwise tree: groups of up to 16 consecutive tasks by 8 x 28 threads, then the groups.
// Leaves sh.ev[0..27], ev_count, ev_nsup, ev_cost.
__device__ __forceinline__ void combine_parts(Shared& sh, const double* part, int units) {
    const int tid = threadIdx.x, grp = tid >> 5, i = tid & 31;
    const int in_group = units < 16 ? units : 16, groups = units < 16 ? 1 : units >> 4;
    if (i < 28 && grp < groups) {
        const double* src = part + (size_t)(grp * 16) * kPartStride + i;
        double v[16];
#pragma unroll
        for (int r = 0; r < 16; ++r) v[r] = r < in_group ? __ldcg(src + r * kPartStride) : 0.0;
#pragma unroll
        for (int w = 1; w < 16; w <<= 1) {
            if (w < in_group) {
#pragma unroll
                for (int k = 0; k + w < 16; k += 2 * w) v[k] = __dadd_rn(v[k], v[k + w]);
            }
        }
        sh.wtot[grp][i] = v[0];
    }
    __syncthreads();
    if (tid < 28) {
        double v[kWarps];
#pragma unroll
        for (int g = 0; g < kWarps; ++g) v[g] = g < groups ? sh.wtot[g][tid] : 0.0;
#pragma unroll
        for (int w = 1; w < kWarps; w <<= 1) {
            if (w < groups) {
#pragma unroll
                for (int k = 0; k + w < kWarps; k += 2 * w) v[k] = __dadd_rn(v[k], v[k + w]);
            }
        }
        sh.ev[tid] = v[0];
    } else if (tid >= 32 && tid < 34) {
        long long sum = 0;
        for (int r = 0; r < units; ++r) sum += __ldcg(reinterpret_cast<const long long*>(part + (size_t)r * kPartStride) + 28 + (tid - 32));
        (&sh.ev_count)[tid - 32] = (int32_t)sum;
    }
    __syncthreads();
    if (tid == 0) sh.ev_cost = sh.ev[27] / (double)sh.ev_count;
    __syncthreads();
}

template <int MODE>
__device__ __forceinline__ void task_body(const LaunchParams& p) {
    extern __shared__ __align__(128) unsigned char ring[];
    __shared__ Shared sh;
    __shared__ double evpose[12];
    __shared__ int32_t s_task, s_last;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const SchedDev& sc = p.sched;
    RingPos rp;
    rp.pos = 0;
    rp.cpar = 0;

    // ---- queries are started (initial state + the tasks of their first evaluation) by whichever CTA comes by:
    // one per pass through the loop below until none is left, so the launch makes progress with any number of
    // resident CTAs (nothing waits for a CTA that has not been scheduled yet).
    auto start_query = [&](int q) {
        const QueryDev& qd = p.queries[q];
        QueryState& g = sc.states[q];
        for (int i = 0; i < 9; ++i) g.R[i] = g.Rn[i] = g.Rb[i] = qd.R0[i];
        for (int i = 0; i < 3; ++i) g.t[i] = g.tn[i] = g.tb[i] = qd.t0[i];
        for (int i = 0; i < 27; ++i) g.lin[i] = 0.0;
        g.prev_cost = g.best_cost = NAN;
        g.lambda = p.lambda0;
        g.lr = 1.0;
        g.initial_cost = NAN;
        g.gathered = 0;
        g.best_n = -1;
        g.has_best = 0;
        g.status = FMPNP_STATUS_OK;
        g.flag = 0;
        g.hist_len = 0;
        g.n_evals = 0;
        g.initial_set = 0;
        for (int i = 0; i < FMPNP_MAX_LEVELS; ++i) g.level_status[i] = 0;
        g.level = 0;
        g.it = 0;
        if (MODE == kModeRefine) p.results[q].status = FMPNP_STATUS_INCOMPLETE;
        if (MODE == kModeComputeCost)
            describe_evaluation(g, qd.R0, qd.t0, kPhaseCost, 0, 0, qd.C, p.use_ratio ? 2 : 0, FMPNP_LOSS_SQUARED, -1, sc.units);
        else if (MODE == kModeFindInliers)
            describe_evaluation(g, qd.R0, qd.t0, kPhaseCost, 0, 0, qd.C, 1, p.loss_id, -1, sc.units);
        else if (p.n_levels > 0)  // multilevel: compute_cost on ALL channels first (model.py:182)
            describe_evaluation(g, qd.R0, qd.t0, kPhaseCost, 0, 0, qd.C, p.use_ratio ? 2 : 0, FMPNP_LOSS_SQUARED, -1, sc.units);
        else
            describe_evaluation(g, qd.R0, qd.t0, kPhaseFirst, 1, 0, qd.C, p.use_ratio ? 1 : 0, p.loss_id, 0, sc.units);
        enqueue_evaluation(sc, q);
    };

    for (;;) {
        __syncthreads();
        if (tid == 0) {
            if (ld_volatile_u32(sc.started) < (unsigned int)p.B) {
                const unsigned int q = atomicAdd(sc.started, 1u);
                if (q < (unsigned int)p.B) start_query((int)q);
            }
            s_task = dequeue_task(sc, p.B);
        }
        __syncthreads();
        const int task = s_task;
        if (task < 0) break;
        const int qi = task / kRanges, u = task - qi * kRanges;
        const QueryDev& q = p.queries[qi];
        QueryState& g = sc.states[qi];
        const volatile QueryState& gv = g;
        if (tid < 9) evpose[tid] = gv.evR[tid];
        if (tid >= 9 && tid < 12) evpose[tid] = gv.evt[tid - 9];

        EvalCtx c;
        c.q = &q;
        c.rec = p.rec + (size_t)qi * p.n_max * kRecDoubles;
        c.supp = p.supp + (size_t)qi * p.n_max;
        c.ring = ring;
        c.ring_per_warp = p.ring_per_warp;
        c.min_depth = p.min_depth;
        c.csize = 1;
        c.crank = 0;
        c.part_out = nullptr;
        c.limit_given = false;
        c.c0 = gv.ev_c0;
        c.Cl = gv.ev_Cl;
        c.loss_id = gv.ev_loss;
        c.alpha = p.alpha;
        c.ratio_mode = gv.ev_ratio_mode;
        c.thr = p.ratio_thr;
        const bool grad = gv.ev_grad != 0;
        const int hist_entry = gv.ev_hist;
        c.aligned = slice_aligned(q, c.c0, c.Cl, grad);
        c.inl_out = MODE == kModeFindInliers ? p.inliers_out + (size_t)qi * p.inliers_stride : nullptr;
        c.h_px = nullptr;
        c.h_mask = nullptr;
        c.h_kept = nullptr;
        if (hist_entry >= 0) hist_slices(p, qi, hist_entry, c);
        {
            const int per = max(1, (q.N + kRanges - 1) / kRanges);
            const int rpu = kRanges / sc.units;  // fixed ranges per task
            c.range_len = per;
            c.n0 = min(q.N, u * rpu * per);
            c.n1 = min(q.N, (u + 1) * rpu * per);
        }
        __syncthreads();

        // ---- the task.  Round 0 streams this run of points (one record per supported point).  Then the
        // per-point phase of the task's own ranges: without a ratio test it needs nothing from the other tasks
        // and follows at once (28 + 2 partial sums); with one, the limit is a maximum over ALL points, so
        // round 0 only leaves the task's max |rho|, the CTA that completes the round publishes the limit and
        // queues the tasks again (round 1: sums only, the records come from L2).
        const int round = gv.ev_round;
        const bool multi = sc.units > 1, ratio = c.ratio_mode != 0;
        double* const parts = sc.part + (size_t)qi * sc.units * kPartStride;
        if (round == 0) {
            rp.pos = 0;
            if (grad) {
                set_geometry(c, 4);
                stream_points<true>(c, evpose, evpose + 9, sh, rp);
            } else {
                set_geometry(c, 2);
                stream_points<false>(c, evpose, evpose + 9, sh, rp);
            }
            __syncthreads();
        }
        if (multi && ratio && round == 0) {
            double mm;
            int nf;
            block_ratio_max(c, sh, mm, nf);
            if (tid == 0) {
                parts[u * kPartStride + 30] = mm;
                reinterpret_cast<long long*>(parts)[u * kPartStride + 31] = nf;
            }
        } else if (multi) {
            if (ratio) {
                if (tid == 0) sh.limit = gv.ev_limit;
                __syncthreads();
            }
            c.csize = sc.units;
            c.crank = u;
            c.part_out = parts + u * kPartStride;
            c.limit_given = ratio;
            if (grad)
                reduce_points<true>(c, evpose, evpose + 9, sh, rp);
            else
                reduce_points<false>(c, evpose, evpose + 9, sh, rp);
            c.csize = 1;
            c.crank = 0;
            c.part_out = nullptr;
            c.limit_given = false;
        }
        if (tid == 0) {
            __threadfence();  // records / partials before the count
            s_last = atomicSub(&g.pending, 1) == 1;
            if (s_last) __threadfence();
        }
        __syncthreads();
        if (!s_last) continue;

        if (multi && ratio && round == 0) {  // every task has streamed: limit = max |rho| * threshold, then round 1
            if (tid == 0) {
                double mm = -INFINITY;
                long long nf = 0;
                for (int r = 0; r < sc.units; ++r) {
                    mm = fmax(mm, __ldcg(parts + r * kPartStride + 30));
                    nf |= __ldcg(reinterpret_cast<const long long*>(parts) + r * kPartStride + 31);
                }
                g.ev_limit = nf ? NAN : __dmul_rn(mm, c.thr);  // torch.max propagates NaN
                g.ev_round = 1;
                g.pending = sc.units;
                enqueue_evaluation(sc, qi);
            }
            continue;
        }

        // ---- this CTA completed the evaluation: total the query, advance its state machine
        c.n0 = 0;
        c.n1 = q.N;
        if (tid == 0) load_state(sh, gv);
        __syncthreads();
        if (multi)
            combine_parts(sh, parts, sc.units);
        else if (grad)
            reduce_points<true>(c, evpose, evpose + 9, sh, rp);
        else
            reduce_points<false>(c, evpose, evpose + 9, sh, rp);
        if (tid == 0) {
            sh.n_evals += 1;
            sh.gathered += (long long)sh.ev_nsup * c.Cl * (grad ? 4 : 2);
        }
        __syncthreads();

        bool finished = false, have_eval = false;
        int lvl = gv.level, it = gv.it;
        const int phase = gv.phase;
        enum { kStartLevel, kTryStep, kFinishLevel } next = kStartLevel;

        if (MODE == kModeComputeCost) {  // sparseFeaturePnP.compute_cost (model.py:216-243)
            if (tid == 0) {
                p.cost_out[qi] = sh.ev_nsup > 0 ? sh.ev_cost : NAN;
                if (p.nsup_out) p.nsup_out[qi] = sh.ev_nsup;
            }
            finished = true;
        } else if (MODE == kModeFindInliers) {  // find_inliers (model.py:131-152): the mask was written by reduce_points
            finished = true;
        } else if (phase == kPhaseCost) {
            if (sh.ev_nsup == 0) {  // compute_cost is None -> return the initial pose (model.py:183-187)
                if (tid == 0) sh.status = FMPNP_STATUS_NO_INITIAL_COST;
                finished = true;
            } else {
                if (tid == 0) {
                    sh.initial_cost = sh.ev_cost;
                    sh.initial_set = 1;
                }
                lvl = 0;
                next = kStartLevel;
            }
        } else if (phase == kPhaseFirst) {  // i == 0 of a level (model.py:303-359)
            if (sh.ev_nsup == 0) {          // model.py:316-320 -> forward returns the CURRENT pose
                if (tid == 0) {
                    sh.status = FMPNP_STATUS_EMPTY_SUPPORT;
                    sh.level_status[lvl] = FMPNP_STATUS_EMPTY_SUPPORT;
                }
                ++lvl;
                next = kStartLevel;
                finished = lvl >= (p.n_levels > 0 ? p.n_levels : 1);
            } else {
                if (tid == 0) {
                    sh.prev_cost = sh.ev_cost;
                    sh.best_cost = sh.ev_cost;
                    sh.best_n = sh.ev_count;
                    sh.has_best = 1;
                    for (int i = 0; i < 9; ++i) sh.Rb[i] = sh.R[i];
                    for (int i = 0; i < 3; ++i) sh.tb[i] = sh.t[i];
                    if (!sh.initial_set) {  // `if self.initial_cost_ is None` (model.py:355-356)
                        sh.initial_cost = sh.ev_cost;
                        sh.initial_set = 1;
                    }
                    for (int i = 0; i < 27; ++i) sh.lin[i] = sh.ev[i];
                    sh.lambda = p.lambda0;  // model.py:281
                    sh.lr = 1.0;
                    sh.status = FMPNP_STATUS_OK;
                    sh.flag = 0;
                    hist_record(p, qi, sh.hist_len, sh.R, sh.t, sh.ev_cost);
                    sh.hist_len += 1;
                }
                it = 0;
                next = kTryStep;
            }
        } else {  // trial pose evaluated (model.py:428-486)
            if (sh.ev_nsup == 0) {  // model.py:441-445 -> CURRENT pose
                if (tid == 0) {
                    sh.status = FMPNP_STATUS_EMPTY_SUPPORT;
                    sh.flag = 2;
                }
                next = kFinishLevel;
            } else {
                if (tid == 0) {
                    const double new_cost = sh.ev_cost;
                    hist_record(p, qi, sh.hist_len, sh.Rn, sh.tn, new_cost);  // model.py:464-465
                    sh.hist_len += 1;
                    const bool worse = new_cost > sh.prev_cost;
                    {  // model.py:469-470
                        const double l = __dmul_rn(sh.lambda, worse ? 10.0 : 0.1);
                        sh.lambda = l < 1e-6 ? 1e-6 : (l > 1e4 ? 1e4 : l);
                    }
                    if (worse) {  // model.py:472-476: reject, keep (R, t) and its linearisation
                        const double l = __dmul_rn(0.1, sh.lr);
                        sh.lr = l < 1e-3 ? 1e-3 : (l > 1.0 ? 1.0 : l);
                    } else {  // model.py:477-486: accept
                        sh.lr = 1.0;
                        if (new_cost < sh.best_cost) {
                            for (int i = 0; i < 9; ++i) sh.Rb[i] = sh.Rn[i];
                            for (int i = 0; i < 3; ++i) sh.tb[i] = sh.tn[i];
                            sh.best_n = sh.ev_count;
                            sh.best_cost = new_cost;
                        }
                        sh.prev_cost = new_cost;
                        for (int i = 0; i < 9; ++i) sh.R[i] = sh.Rn[i];
                        for (int i = 0; i < 3; ++i) sh.t[i] = sh.tn[i];
                        for (int i = 0; i < 27; ++i) sh.lin[i] = sh.ev[i];
                    }
                }
                ++it;
                next = kTryStep;
            }
        }

        const int nlev = p.n_levels > 0 ? p.n_levels : 1;
        while (!finished && !have_eval) {
            __syncthreads();
            if (next == kStartLevel) {
                if (tid == 0) {
                    const int c0 = p.n_levels > 0 ? p.levels[lvl][0] : 0;
                    const int Cl = p.n_levels > 0 ? p.levels[lvl][1] - p.levels[lvl][0] : q.C;
                    describe_evaluation(g, sh.R, sh.t, kPhaseFirst, 1, c0, Cl, p.use_ratio ? 1 : 0, p.loss_id, sh.hist_len,
                                        sc.units);
                }
                have_eval = true;
            } else if (next == kTryStep) {
                if (it >= p.n_iters) {
                    next = kFinishLevel;
                    continue;
                }
                if (warp == 0) solve_and_update(sh, p.solver, lane);
                __syncthreads();
                if (sh.flag == 1) {  // NaN step: leave the loop, best pose (model.py:411-413)
                    __syncthreads();
                    if (tid == 0) sh.status = FMPNP_STATUS_NAN_STEP;
                    next = kFinishLevel;
                    continue;
                }
                if (tid == 0)
                    describe_evaluation(g, sh.Rn, sh.tn, kPhaseTrial, 1, gv.ev_c0, gv.ev_Cl, p.use_ratio ? 1 : 0, p.loss_id,
                                        sh.hist_len, sc.units);
                have_eval = true;
            } else {  // kFinishLevel (model.py:491-494): best pose unless the loop was left on an empty support
                if (tid == 0) {
                    if (sh.flag != 2) {
                        for (int i = 0; i < 9; ++i) sh.R[i] = sh.Rb[i];
                        for (int i = 0; i < 3; ++i) sh.t[i] = sh.tb[i];
                    }
                    sh.level_status[lvl] = sh.status;
                    sh.flag = 0;
                }
                ++lvl;
                next = kStartLevel;
                finished = lvl >= nlev;
            }
        }
        __syncthreads();
        if (tid == 0) {
            if (finished) {
                if (MODE == kModeRefine) {
                    fmpnp_result& r = p.results[qi];
                    for (int i = 0; i < 9; ++i) r.R[i] = sh.R[i];
                    for (int i = 0; i < 3; ++i) r.t[i] = sh.t[i];
                    r.best_cost = sh.best_cost;
                    r.initial_cost = sh.initial_cost;
                    r.best_num_inliers = sh.best_n;
                    r.status = sh.status;
                    r.n_evaluations = sh.n_evals;
                    for (int i = 0; i < FMPNP_MAX_LEVELS; ++i) r.level_status[i] = sh.level_status[i];
                    r.gathered_rows = sh.gathered;
                    if (p.hist.enabled) p.hist.len[qi] = sh.hist_len;
                }
                __threadfence();
                atomicAdd(sc.done, 1u);
            } else {
                store_state(sh, g);
                g.level = lvl;
                g.it = it;
                enqueue_evaluation(sc, qi);
            }
        }
    }
}

__global__ void __launch_bounds__(kThreads, 2) refine_tasks_kernel(const LaunchParams p) { task_body<kModeRefine>(p); }
__global__ void __launch_bounds__(kThreads, 2) compute_cost_tasks_kernel(const LaunchParams p) {
    task_body<kModeComputeCost>(p);
}
__global__ void __launch_bounds__(kThreads, 2) find_inliers_tasks_kernel(const LaunchParams p) {
    task_body<kModeFindInliers>(p);
}

// ---------------------------------------------------------------------------------------------
// CHW (float32 / float64) -> HWC float32 pack: a (C) x (Hf*Wf) transpose through shared memory.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void pack_chw_to_hwc_kernel(const T* __restrict__ src, float* __restrict__ dst, int C, int HW) {
    __shared__ float tile[32][33];
    const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, p = p0 + threadIdx.x;
        if (c < C && p < HW) tile[i][threadIdx.x] = (float)src[(size_t)c * HW + p];
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int p = p0 + i, c = c0 + threadIdx.x;
        if (c < C && p < HW) dst[(size_t)p * C + c] = tile[threadIdx.x][i];
    }
}

cudaError_t launch_pack_chw_to_hwc(const void* src, int src_is_f64, float* dst, int C, int Hf, int Wf,
                                   cudaStream_t stream) {
    const int HW = Hf * Wf;
    dim3 block(32, 8), grid((HW + 31) / 32, (C + 31) / 32);
    if (src_is_f64)
        pack_chw_to_hwc_kernel<double><<<grid, block, 0, stream>>>(static_cast<const double*>(src), dst, C, HW);
    else
        pack_chw_to_hwc_kernel<float><<<grid, block, 0, stream>>>(static_cast<const float*>(src), dst, C, HW);
    return cudaGetLastError();
}

size_t refine_static_smem_bytes() { return sizeof(Shared) + 128; }

cudaError_t refine_set_max_dynamic_smem(int bytes) {
    cudaError_t e = cudaFuncSetAttribute(refine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(compute_cost_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(find_inliers_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(refine_tasks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(compute_cost_tasks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(find_inliers_tasks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    return e;
}

// clusters of 16 CTAs (single-query latency mode) are beyond the portable size of 8
cudaError_t refine_allow_nonportable_clusters(int allow) {
    cudaError_t e = cudaFuncSetAttribute(refine_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, allow);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(compute_cost_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, allow);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(find_inliers_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, allow);
    return e;
}

static void (*kernel_of(int mode, int scheduler))(LaunchParams) {
    if (scheduler == 1) {
        switch (mode) {
            case kModeComputeCost: return compute_cost_tasks_kernel;
            case kModeFindInliers: return find_inliers_tasks_kernel;
            default: return refine_tasks_kernel;
        }
    }
    switch (mode) {
        case kModeComputeCost: return compute_cost_kernel;
        case kModeFindInliers: return find_inliers_kernel;
        default: return refine_kernel;
    }
}

// grid = clusters x cluster_size CTAs; cluster_size > 1 launches thread-block clusters
cudaError_t launch_refine(const LaunchParams& p, int grid, int ring_bytes, cudaStream_t stream) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = (size_t)ring_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)(p.cluster_size > 1 ? p.cluster_size : 1);
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = p.cluster_size > 1 ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel_of(p.mode, p.scheduler), p);
}

// how many clusters of `cluster_size` CTAs (each with ring_bytes of dynamic shared memory) can be
// resident at once; 0 if such a cluster cannot be scheduled at all
int max_active_clusters(int cluster_size, int ring_bytes) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)cluster_size);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = (size_t)ring_bytes;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)cluster_size;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, refine_kernel, &cfg) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

}  // namespace fmpnp
