// fmpnp_light.cuh -- a pose evaluation that expects to find (nearly) every record in place: the evaluations after the
// first one of a level when records are reused (fmpnp_options.gather_every_row = 0).
// Part of the refinement kernels (fmpnp_kernels.cu); not a public interface.
//
// Same results as evaluate<true>() in fmpnp_lm.cuh, bit for bit -- the same projection, the same tag rule, the same
// records, the same per-point phase (reduce_points) -- but organised for latency instead of bandwidth:
//   * a warp works on the points of the reduction ranges it owns, in both phases, so nothing but the ratio-test limit
//     crosses warps: one block barrier per evaluation (none without a ratio test) instead of three;
//   * the projection runs over two points per lane at a time (two independent pairs of float64 division chains per
//     thread), and what it reads of them -- coordinates, tag, first record entry -- is staged into the warp's ring
//     slice with cp.async one group AHEAD: the L2 round trip is behind the arithmetic, no register holds data in flight;
//   * the ratio-test maximum is taken from the records on the way (the first entry of a record shares the tag's
//     cache line), not in a pass of its own;
//   * the few points that left their cell are listed per warp and streamed by that warp through its slice of the ring
//     (stream_list) before it goes on -- no other warp waits for them.
#pragma once
#include "fmpnp_reduce.cuh"

namespace fmpnp {

// |rho| of a supported point for the ratio test (model.py:120-129), from the first entry of its record
__device__ __forceinline__ double ratio_abs_rho(const EvalCtx& c, double rec0) {
    const double x = __dmul_rn(0.5, rec0);
    double rho = x, w = 1.0;
    if (c.ratio_mode == 1) robust_loss(c.loss_id, c.alpha, x, rho, w);
    return fabs(rho);
}

// Projection + tag check of the points of this warp's ranges; streams the rows of those without a valid record.
// Leaves supp[] / tags / history exactly as project_and_flag() does.  vmax / vnan: max |rho| over the warp's supported
// points and whether any is NaN (valid in every lane only after the caller's reduction over lanes).
__device__ void light_project(const EvalCtx& c, const double* R, const double* t, Shared& sh, RingPos& rp, double& vmax,
                              int& vnan) {
    constexpr int G = 2;  // points per lane worked on together (and as many being loaded)
    const QueryDev& q = *c.q;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int per = c.range_len, nr = c.nranges / c.csize, r_first = c.crank * nr;
    const int rpw = nr >= kWarps ? nr / kWarps : 1, wact = nr >= kWarps ? kWarps : nr;
    vmax = -INFINITY;
    vnan = 0;
    if (warp >= wact) return;
    const int w0 = min(q.N, (r_first + warp * rpw) * per), w1 = min(q.N, (r_first + (warp + 1) * rpw) * per);
    if (w0 >= w1) return;
    const bool ratio = c.ratio_mode != 0;
    const double* const pts = q.pts;
    const uint8_t* const active = c.active;
    const int32_t im_w = q.im_w, im_h = q.im_h, Hf = q.Hf, Wf = q.Wf;
    const int32_t sx = cell_shift(im_w, Wf), sy = cell_shift(im_h, Hf);
    const unsigned long long gen = (unsigned long long)c.reuse_gen << 32;
    double K[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) K[i] = q.K[i];
    int lcnt = 0;  // entries of the warp's list (buffer 0 of bl_cell / bl_n)

    bool flushed = false;  // a flush used the ring: what was staged there for the next group is gone
    auto flush = [&]() {
        if (lcnt == 0) return;
        __syncwarp();
        stream_list(c, sh, rp, lcnt);
        __syncwarp();  // the records are written (by whichever lanes reduced them)
        if (ratio && lane < lcnt) {
            const double a = ratio_abs_rho(c, __ldcg(c.rec + (size_t)sh.bl_n[warp][0][lane] * kRecDoubles));
            if (a != a) vnan = 1;
            vmax = fmax(vmax, a);
        }
        __syncwarp();  // the list may be overwritten
        lcnt = 0;
        flushed = true;
    };

    // What the projection reads of a point -- coordinates (24 B), the record's first entry and its tag (16 B around
    // each) -- is copied into the warp's ring slice with cp.async one group of G rows AHEAD of
    // the group being projected; every lane copies and reads its own points only.  The mask byte travels in a register.
    constexpr int kRowBytes = 32 * (24 + 32), kGroupBytes = G * kRowBytes;
    const bool stage = c.ring_per_warp >= 2 * kGroupBytes;
    const uint32_t sbuf = smem_u32(c.ring) + (uint32_t)(warp * c.ring_per_warp);
    auto issue_group = [&](int base, int b) {
#pragma unroll
        for (int k = 0; k < G; ++k) {
            const int n = base + 32 * k + lane;
            if (n < w1) {
                // row layout: 32 x [c, Sxx], 32 x [tag, pad], 32 x [x, y, z].  The record pieces are written during the
                // launch (by other CTAs too): 16-byte .cg copies, never through L1; the coordinates are input
                const uint32_t d = sbuf + (uint32_t)(b * kGroupBytes + k * kRowBytes);
                const char* r = reinterpret_cast<const char*>(c.rec + (size_t)n * kRecDoubles);
                cp_async16_pol(d + lane * 16, r, c.rec_pol);
                cp_async16_pol(d + 512 + lane * 16, r + 8 * kRecTagSlot, c.rec_pol);
                const char* pt = reinterpret_cast<const char*>(pts + 3 * (size_t)n);
                const uint32_t e = d + 1024 + lane * 24;
                cp_async8_s(e, pt);
                cp_async8_s(e + 8, pt + 8);
                cp_async8_s(e + 16, pt + 16);
            }
        }
        cp_async_commit();
    };
    auto mask_of = [&](int base, int k) -> bool {
        const int n = base + 32 * k + lane;
        return (active == nullptr) || (n < w1 && __ldcg(active + n) != 0);  // .cg: may be another CTA's inlier mask
    };
    bool act_nx[G];
    if (stage) {
        issue_group(w0, 0);
#pragma unroll
        for (int k = 0; k < G; ++k) act_nx[k] = mask_of(w0, k);
    }
    int b = 0;
    for (int base = w0; base < w1; base += 32 * G, b ^= 1) {
        double x[G], y[G], z[G], r0[G];
        long long have[G];
        bool act[G];
        const bool more = base + 32 * G < w1;
        if (stage) {
#pragma unroll
            for (int k = 0; k < G; ++k) act[k] = act_nx[k];
            if (more) {
                issue_group(base + 32 * G, b ^ 1);
#pragma unroll
                for (int k = 0; k < G; ++k) act_nx[k] = mask_of(base + 32 * G, k);
                cp_async_wait_pending(1);
            } else {
                cp_async_wait_pending(0);
            }
#pragma unroll
            for (int k = 0; k < G; ++k) {
                const uint32_t d = sbuf + (uint32_t)(b * kGroupBytes + k * kRowBytes);
                const bool valid = base + 32 * k + lane < w1;
                x[k] = y[k] = r0[k] = 0.0;
                z[k] = 1.0;
                have[k] = 0ll;
                if (valid) {
                    const uint32_t e = d + 1024 + lane * 24;
                    x[k] = lds64d(e);
                    y[k] = lds64d(e + 8);
                    z[k] = lds64d(e + 16);
                    r0[k] = lds128d(d + lane * 16).x;
                    have[k] = __double_as_longlong(lds128d(d + 512 + lane * 16).x);
                }
            }
        } else {
#pragma unroll
            for (int k = 0; k < G; ++k) {  // every load of the group first
                const int n = base + 32 * k + lane;
                const int m = n < w1 ? n : w1 - 1;  // lanes past the end load a valid address and discard
                act[k] = (active == nullptr) || (__ldcg(active + m) != 0);
                const double* rec = c.rec + (size_t)m * kRecDoubles;
                have[k] = __ldcg(reinterpret_cast<const long long*>(rec + kRecTagSlot));
                r0[k] = __ldcg(rec);
                x[k] = pts[3 * m + 0];
                y[k] = pts[3 * m + 1];
                z[k] = pts[3 * m + 2];
            }
        }
        flushed = false;
#pragma unroll
        for (int k = 0; k < G; ++k) {
            const int n = base + 32 * k + lane;
            const bool valid = n < w1;
            const Proj pr = project_point(R, t, K, x[k], y[k], z[k], im_w, im_h);
            const bool sup = valid && act[k] && pr.supported;
            if (c.h_px && valid) {
                c.h_px[2 * n + 0] = pr.px;
                c.h_px[2 * n + 1] = pr.py;
                c.h_mask[n] = sup ? 1 : 0;
            }
            int32_t cell = -1;
            bool hit = false;
            if (sup) {
                cell = cell_index_shifted(pr.px, pr.py, Hf, Wf, im_w, im_h, sx, sy);
                // nearest-cell lookup (model.py:74-97): same cell, same slice => the record the point has is the one
                // this evaluation would compute
                const long long want = (long long)(gen | (unsigned int)cell);
                hit = have[k] == want;
                if (!hit) __stcg(reinterpret_cast<long long*>(c.rec + (size_t)n * kRecDoubles + kRecTagSlot), want);
            }
            if (valid) c.supp[n] = sup ? (hit ? 2 : 1) : 0;
            if (ratio && hit) {
                const double a = ratio_abs_rho(c, r0[k]);
                if (a != a) vnan = 1;
                vmax = fmax(vmax, a);
            }
            const bool miss = sup && !hit;
            const unsigned mask = __ballot_sync(0xffffffffu, miss);
            if (mask != 0u) {  // uniform
                const int add = __popc(mask);
                if (lcnt + add > 32) flush();
                if (miss) {
                    const int rank = lcnt + __popc(mask & ((1u << lane) - 1u));
                    sh.bl_cell[warp][0][rank] = cell;
                    sh.bl_n[warp][0][rank] = n;
                }
                lcnt += add;
            }
        }
        if (stage && flushed && more) {  // the flush overwrote the next group's staged rows: copy them again
            cp_async_wait_pending(0);
            issue_group(base + 32 * G, b ^ 1);
        }
    }
    flush();
}

// One evaluation at pose (R, t) with gradients: leaves sh.ev[0..27], sh.ev_count, ev_nsup, ev_ngath, ev_cost like
// evaluate<true>().  The CTA owns ranges [crank * nranges / csize, ...) of the query and c.reuse_gen != 0; csize == 1
// unless the caller combines partials itself (c.part_out).
__device__ void light_evaluate(EvalCtx& c, const double* R, const double* t, Shared& sh, RingPos& rp) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    set_geometry(c, 4);
    rp.pos = 0;
    double m;
    int nanf;
    light_project(c, R, t, sh, rp, m, nanf);
    const bool ratio = c.ratio_mode != 0;
    if (ratio) {  // limit = max |rho| * threshold over ALL supported points (model.py:120-129)
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
        if (tid == 0) {
            double mm = sh.red[0][0];
            int nf = sh.red_i[0][0];
            for (int w = 1; w < kWarps; ++w) {
                mm = fmax(mm, sh.red[w][0]);
                nf |= sh.red_i[w][0];
            }
            sh.limit = nf ? NAN : __dmul_rn(mm, c.thr);  // torch.max propagates NaN
        }
        __syncthreads();
    } else {
        __syncwarp();  // a warp's per-point phase reads what its own lanes wrote
    }
    const bool given = c.limit_given;
    c.limit_given = ratio;
    reduce_points<true>(c, R, t, sh, rp);
    c.limit_given = given;
}

}  // namespace fmpnp
