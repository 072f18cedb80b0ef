// fmpnp_stream.cuh -- shared-memory state of a CTA and the streaming phase of a pose evaluation: every warp projects its points, copies their rows into its slice of the staging ring with cp.async and accumulates one record per point.
// Part of the refinement kernels (fmpnp_kernels.cu); not a public interface.
#pragma once
#include <cstdint>

#include "fmpnp_device.cuh"
#include "fmpnp_kernels.cuh"

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
    long long gathered, reused;
    int32_t red_i[kWarps][4];
    int32_t cpart_i[2][4];
    int32_t climit_nan;
    int32_t ev_count, ev_nsup, ev_ngath;  // survivors, supported, supported and gathered (not reused); adjacent
    int32_t best_n, has_best, status, flag, hist_len, query, n_evals, initial_set;
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

// Number of fixed reduction ranges of a query: a function of its point count only (never of the launch), a power
// of two in [16, 128] such that a range holds about kRangePoints = 128 points -- four rows of a warp, one transposition
// and lane tree per range -- when the query has enough of them (measured: 32 / 64 / 128 points per range give cfg2
// 4.56 / 4.40 / 4.32 ms, cfg3 4.42 / 4.28 / 4.22 ms, one cfg3 query alone 1.20 / 0.86 / 0.75 ms; profiles/r02/).
__device__ __forceinline__ int ranges_of(int N) {
    int r = kMinRanges;
    while (r < kRanges && N >= 2 * kRangePoints * r) r <<= 1;
    return r;
}

struct EvalCtx {
    const QueryDev* q;
    int32_t c0, Cl;        // channel slice [c0, c0 + Cl)
    bool aligned;          // staged (16-byte) path allowed
    int32_t loss_id;
    double alpha;
    int32_t ratio_mode;    // 0 none, 1 on rho (forward / find_inliers), 2 on raw SSD (compute_cost)
    double thr;
    double* rec;           // per-point records of this CTA
    uint8_t* supp;         // per-point support flags of this CTA: 0 unsupported, 1 gathered, 2 record reused
    const uint8_t* active; // points to leave out (0), or null: the caller's mask, or the fused find_inliers pass's
    uint32_t reuse_gen;    // generation of the record tags this evaluation may reuse (0 = gather every row)
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
    int32_t nranges, range_len;        // fixed reduction ranges of this query (ranges_of(N)), points per range
    int32_t csize, crank;              // thread-block cluster working on this query (or: tasks of this evaluation)
    double* part_out;                  // task scheduler: this task's partial sums go here (global), no combine
    bool limit_given;                  // task scheduler: sh.limit already holds the ratio-test limit
    uint64_t row_pol, rec_pol;         // L2 policies: gathered rows (read once), staged records / coordinates (read again)
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
    const bool act = (c.active == nullptr) || (__ldcg(c.active + n) != 0);  // .cg: may be another CTA's inlier mask
    // the record's tag does not depend on the projection: its load travels with the point's (one L2 round trip, not two)
    long long* const tag = reinterpret_cast<long long*>(c.rec + (size_t)n * kRecDoubles + kRecTagSlot);
    long long have = 0;
    if (c.reuse_gen != 0u) have = __ldcg(tag);  // .cg: another CTA may have written it (task scheduler)
    const double x = q.pts[3 * n + 0], y = q.pts[3 * n + 1], z = q.pts[3 * n + 2];
    const Proj pr = project_point(R, t, q.K, x, y, z, q.im_w, q.im_h);
    const bool sup = act && pr.supported;
    if (c.h_px) {
        c.h_px[2 * n + 0] = pr.px;
        c.h_px[2 * n + 1] = pr.py;
        c.h_mask[n] = sup ? 1 : 0;
    }
    int32_t cell = sup ? cell_index(pr.px, pr.py, q.Hf, q.Wf, q.im_w, q.im_h) : -1;
    uint8_t flag = sup ? 1 : 0;
    if (c.reuse_gen != 0u && sup) {
        // nearest-cell lookup (model.py:74-97): same cell, same slice => same rows => the record the point already
        // has is the one this evaluation would compute.
        const long long want = (long long)(((unsigned long long)c.reuse_gen << 32) | (unsigned int)cell);
        if (have == want) {
            flag = 2;
            cell = -1;  // nothing to gather
        } else {
            __stcg(tag, want);  // its rows are gathered by this evaluation (the record follows before anyone reads it)
        }
    }
    c.supp[n] = flag;
    return cell;
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
__device__ __forceinline__ void copy_rows_full(uint32_t dst, const char* const (&s)[4], uint32_t rowB, uint64_t pol) {
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
        const uint32_t d = dst + (uint32_t)r * rowB;
        const char* g = s[r];
        if (FULL > 0) cp_async16_pol_off<0>(d, g, pol);
        if (FULL > 1) cp_async16_pol_off<512>(d, g, pol);
        if (FULL > 2) cp_async16_pol_off<1024>(d, g, pol);
        if (FULL > 3) cp_async16_pol_off<1536>(d, g, pol);
        if (FULL > 4) cp_async16_pol_off<2048>(d, g, pol);
        if (FULL > 5) cp_async16_pol_off<2560>(d, g, pol);
        if (FULL > 6) cp_async16_pol_off<3072>(d, g, pol);
        if (FULL > 7) cp_async16_pol_off<3584>(d, g, pol);
    }
}

// list_cnt >= 0: the warp's supported points are already in its buffer 0 (Shared::bl_cell / bl_n, list_cnt entries: the
// points a light evaluation found without a valid record, fmpnp_light.cuh) -- nothing is projected, that one batch is
// streamed.
template <bool GRAD, int LOG2L>
__device__ void stream_staged(const EvalCtx& c, const double* R, const double* t, Shared& sh, RingPos& rp, int list_cnt = -1) {
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
    const bool list = list_cnt >= 0;
    const int BW = batch_width(N - c.n0);
    const int first = c.n0 + warp * BW;
    const int nb = list ? (list_cnt > 0 ? 1 : 0) : (first < N ? (N - first + kWarps * BW - 1) / (kWarps * BW) : 0);
    if (nb == 0) return;

    // row base pointers with this lane's 16-byte piece folded in
    const char* baseF = reinterpret_cast<const char*>(q.F + c.c0) + lane * 16;
    const char* baseGx = GRAD ? reinterpret_cast<const char*>(q.Gx + c.c0) + lane * 16 : nullptr;
    const char* baseGy = GRAD ? reinterpret_cast<const char*>(q.Gy + c.c0) + lane * 16 : nullptr;
    const char* baseR = reinterpret_cast<const char*>(q.ref + c.c0) + lane * 16;
    const uint32_t pitch = (uint32_t)q.C * 4u;  // bytes between rows of a map / of ref
    const uint64_t pol = c.row_pol;             // L2 policy of the row copies (read once: evict_first)

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
        if (SHORT) {  // off == 0
            s[0] = row_address(baseF, cell, pitch);
            s[ROWS - 1] = row_address(baseR, n, pitch);
            if (GRAD) {
                s[1] = row_address(baseGx, cell, pitch);
                s[2] = row_address(baseGy, cell, pitch);
            }
        } else {
            s[0] = baseF + (size_t)cell * pitch + off;
            s[ROWS - 1] = baseR + (size_t)n * pitch + off;
            if (GRAD) {
                s[1] = baseGx + (size_t)cell * pitch + off;
                s[2] = baseGy + (size_t)cell * pitch + off;
            }
        }
        dst += lane * 16;
        const int full = nv >> 5, tail = nv & 31;  // whole 512-byte warp copies, then a partial one
        if (SHORT) {  // nv <= 32: one 16-byte piece of every row per lane
            if (lane < nv) {
#pragma unroll
                for (int r = 0; r < ROWS; ++r) cp_async16_pol(dst + (uint32_t)(r * nv) * 16u, s[r], pol);
            }
            return;
        }
        if (tail == 0 && full <= 8) {  // the usual case: immediate offsets, no loop
            const uint32_t rowB = (uint32_t)nv * 16u;
            switch (full) {
                case 1: copy_rows_full<ROWS, 1>(dst, s, rowB, pol); break;
                case 2: copy_rows_full<ROWS, 2>(dst, s, rowB, pol); break;
                case 3: copy_rows_full<ROWS, 3>(dst, s, rowB, pol); break;
                case 4: copy_rows_full<ROWS, 4>(dst, s, rowB, pol); break;
                case 5: copy_rows_full<ROWS, 5>(dst, s, rowB, pol); break;
                case 6: copy_rows_full<ROWS, 6>(dst, s, rowB, pol); break;
                case 7: copy_rows_full<ROWS, 7>(dst, s, rowB, pol); break;
                default: copy_rows_full<ROWS, 8>(dst, s, rowB, pol); break;
            }
            return;
        }
#pragma unroll
        for (int r = 0; r < ROWS; ++r) {
            uint32_t d = dst + (uint32_t)(r * nv) * 16u;
            const char* g = s[r];
#pragma unroll 4
            for (int i = 0; i < full; ++i, d += 512, g += 512) cp_async16_pol(d, g, pol);
            if (lane < tail) cp_async16_pol(d, g, pol);
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
            if (GRAD && (LOG2L == 3 || LOG2L == 4)) {
                // Sum over the L lanes of a point through the slot that has just been consumed: every lane stores
                // its six partials, then one lane per (point, value) adds the L of them in the order of the
                // butterfly (pairs at distance L/2, L/4, ...).  45 instructions instead of 80 for the shuffles,
                // selects and register moves of the butterfly.
                constexpr int STR = L + 2;  // doubles per (point, value) row: 16-byte aligned, spreads the banks
                __syncwarp();               // every lane has read its pieces of the slot
                const uint32_t wbase = src0 + (uint32_t)((sg * 6) * STR + sl) * 8u;
#pragma unroll
                for (int j = 0; j < 6; ++j) sts64(wbase + (uint32_t)(j * STR) * 8u, a[j]);
                __syncwarp();
                if (lane < 6 * G) {
                    const int g = lane / 6, j = lane - g * 6, kk = it * G + g;
                    const uint32_t rbase = src0 + (uint32_t)(lane * STR) * 8u;
                    double x[L];
#pragma unroll
                    for (int i = 0; i < L / 2; ++i) {
                        const double2 v = lds128d(rbase + (uint32_t)i * 16u);
                        x[2 * i] = v.x;
                        x[2 * i + 1] = v.y;
                    }
#pragma unroll
                    for (int w = L / 2; w >= 1; w >>= 1) {
#pragma unroll
                        for (int i = 0; i < w; ++i) x[i] = __dadd_rn(x[i], x[i + w]);
                    }
                    if (kk < cnt) c.rec[(size_t)bn[buf * 32 + kk] * kRecDoubles + j] = x[0];
                }
                __syncwarp();  // every lane is done with the slot: it may be refilled
                return;
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

    int cnt0 = list ? list_cnt : load_batch(0, 0), cnt1 = 0;
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

// generic path (channel ranges that are not 16-byte aligned): the whole warp loads the rows of one point directly
template <bool GRAD>
__device__ __forceinline__ void gather_point_generic(const EvalCtx& c, int cell, int n, int lane) {
    const QueryDev& q = *c.q;
    const float* srcF = q.F + (size_t)cell * q.C + c.c0;
    const float* srcGx = GRAD ? q.Gx + (size_t)cell * q.C + c.c0 : nullptr;
    const float* srcGy = GRAD ? q.Gy + (size_t)cell * q.C + c.c0 : nullptr;
    const float* srcR = q.ref + (size_t)n * q.C + c.c0;
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
    if (lane == 0) write_record<GRAD>(a, c.rec + (size_t)n * kRecDoubles);
}

// ---------------------------------------------------------------------------------------------
// Streaming phase: one record per supported point.  Every warp is its own producer and consumer.
// ---------------------------------------------------------------------------------------------
template <bool GRAD>
__device__ void stream_points(const EvalCtx& c, const double* R, const double* t, Shared& sh, RingPos& rp) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
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
                gather_point_generic<GRAD>(c, cj, base + j, lane);
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

// Streams the `cnt` (<= 32) points this warp has listed in its buffer 0 (cell, index): the rows of the points a light
// evaluation found without a valid record.  Called by whole warps, independently of each other.
__device__ void stream_list(const EvalCtx& c, Shared& sh, RingPos& rp, int cnt) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (cnt <= 0) return;
    if (!c.aligned) {
        for (int j = 0; j < cnt; ++j) gather_point_generic<true>(c, sh.bl_cell[warp][0][j], sh.bl_n[warp][0][j], lane);
        return;
    }
    switch (c.L) {
        case 4: stream_staged<true, 2>(c, nullptr, nullptr, sh, rp, cnt); break;
        case 8: stream_staged<true, 3>(c, nullptr, nullptr, sh, rp, cnt); break;
        case 16: stream_staged<true, 4>(c, nullptr, nullptr, sh, rp, cnt); break;
        default: stream_staged<true, 5>(c, nullptr, nullptr, sh, rp, cnt); break;
    }
}

}  // namespace fmpnp
