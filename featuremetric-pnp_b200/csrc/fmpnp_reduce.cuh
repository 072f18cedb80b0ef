// fmpnp_reduce.cuh -- the per-point phase of a pose evaluation (robust loss, ratio test, contraction with the projection Jacobian) and the fixed-order reductions of the normal equations.
// Part of the refinement kernels (fmpnp_kernels.cu); not a public interface.
#pragma once
#include <cooperative_groups.h>

#include "fmpnp_stream.cuh"

namespace fmpnp {

namespace cg = cooperative_groups;

// ---------------------------------------------------------------------------------------------
// Deterministic reductions.  The points of a query are cut into ranges_of(N) = 16...128 fixed ranges (a function
// of the query only, the same for every launch geometry).  A range is summed by ONE warp in a fixed order (lanes stride the range,
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
    const int nr = c.nranges / c.csize;       // ranges this CTA owns: an aligned power-of-two run
    const int r_first = c.crank * nr;
    const int rpw = nr >= kWarps ? nr / kWarps : 1;  // consecutive ranges per warp ...
    const int wact = nr >= kWarps ? kWarps : nr;     // ... of the warps that own any
    double* const wscr = reinterpret_cast<double*>(c.ring + (size_t)warp * c.ring_per_warp);  // idle ring slice
    double* const stack = wscr + kHalfSums * kLaneStride;
    int32_t cnt[3] = {0, 0, 0};  // survivors, supported, supported and gathered by this evaluation (this lane, all its ranges)
    // Staging (with gradients, when the warp's ring slice has room beside the reduction scratch): what the per-point
    // phase reads of a point -- record (48 B) and coordinates (24 B); the addresses depend on n only -- is copied into
    // the idle part of the slice with cp.async one row of 32 points AHEAD, so the L2 round trip is behind the arithmetic
    // and the transposition of the current row instead of in front of them, and no register holds data in flight.
    // Every lane copies and reads its own point only: its own wait_group is all the synchronisation this needs.
    constexpr int kRowRec = 32 * 48, kRowBytes = kRowRec + 32 * 24;
    const bool stage = GRAD && c.ring_per_warp >= kMinRingPerWarp + 2 * kRowBytes;
    const uint32_t sbuf = smem_u32(wscr) + (uint32_t)kMinRingPerWarp;
    const int rows_pr = (per + 31) >> 5;  // rows of 32 points per range
    const int rbase = r_first + warp * rpw;
    auto point_of = [&](int rr, int j) -> int {
        const int r0 = min(q.N, (rbase + rr) * per), r1 = min(q.N, r0 + per);
        const int n = r0 + 32 * j + lane;
        return n < r1 ? n : -1;
    };
    auto issue_row = [&](int n, int buf) {  // one commit group per row (empty for a lane without a point in it)
        if (n >= 0) {
            const char* r = reinterpret_cast<const char*>(c.rec + (size_t)n * kRecDoubles);
            const uint32_t d = sbuf + (uint32_t)(buf * kRowBytes + lane * 48);
            cp_async16_pol(d, r, c.rec_pol);
            cp_async16_pol(d + 16, r + 16, c.rec_pol);
            cp_async16_pol(d + 32, r + 32, c.rec_pol);
            const char* pt = reinterpret_cast<const char*>(q.pts + 3 * (size_t)n);
            const uint32_t e = sbuf + (uint32_t)(buf * kRowBytes + kRowRec + lane * 24);
            cp_async8_s(e, pt);
            cp_async8_s(e + 8, pt + 8);
            cp_async8_s(e + 16, pt + 16);
        }
        cp_async_commit();
    };
    if (warp < wact) {
        int buf = 0;
        uint8_t flag_nx = 0;
        if (stage) {
            const int n0 = point_of(0, 0);
            issue_row(n0, 0);
            if (n0 >= 0) flag_nx = __ldcg(c.supp + n0);
        }
        for (int rr = 0; rr < rpw; ++rr) {
            const int r0 = min(q.N, (rbase + rr) * per), r1 = min(q.N, r0 + per);
            double acc[28];
#pragma unroll
            for (int i = 0; i < 28; ++i) acc[i] = 0.0;
            for (int j = 0; j < rows_pr; ++j, buf ^= 1) {
                const int n = r0 + 32 * j + lane;
                const bool has = n < r1;
                uint8_t sflag = flag_nx;
                if (stage) {
                    int nrr = rr, nj = j + 1;
                    if (nj == rows_pr) {
                        nj = 0;
                        ++nrr;
                    }
                    if (nrr < rpw) {  // uniform
                        const int nn = point_of(nrr, nj);
                        issue_row(nn, buf ^ 1);
                        if (nn >= 0) flag_nx = __ldcg(c.supp + nn);
                        cp_async_wait_pending(1);
                    } else {
                        cp_async_wait_pending(0);
                    }
                }
                if (has) {
                    // flag, record and point: from the staged row, or loaded side by side (one L2 round trip per point;
                    // .cg: flags and records may have been written by another CTA); what an unsupported point's record
                    // holds is never used
                    double2 rc0, rc1 = make_double2(0.0, 0.0), rc2 = rc1;
                    double ptx = 0.0, pty = 0.0, ptz = 0.0;
                    if (stage) {
                        const uint32_t d = sbuf + (uint32_t)(buf * kRowBytes + lane * 48);
                        rc0 = lds128d(d);
                        rc1 = lds128d(d + 16);
                        rc2 = lds128d(d + 32);
                        const uint32_t e = sbuf + (uint32_t)(buf * kRowBytes + kRowRec + lane * 24);
                        ptx = lds64d(e);
                        pty = lds64d(e + 8);
                        ptz = lds64d(e + 16);
                    } else {
                        sflag = __ldcg(c.supp + n);
                        const double2* r = reinterpret_cast<const double2*>(c.rec + (size_t)n * kRecDoubles);
                        rc0 = __ldcg(r);
                        if (GRAD) {
                            rc1 = __ldcg(r + 1);
                            rc2 = __ldcg(r + 2);
                            ptx = q.pts[3 * n];
                            pty = q.pts[3 * n + 1];
                            ptz = q.pts[3 * n + 2];
                        }
                    }
                    const bool sup = sflag != 0;
                    bool keep = false;
                    if (sup) {
                        ++cnt[1];
                        cnt[2] += sflag == 1;
                        const double x = __dmul_rn(0.5, rc0.x);  // cost = 0.5 * sum e^2 (model.py:338)
                        double rho = x, w = 1.0;
                        if (c.ratio_mode != 2) robust_loss(c.loss_id, c.alpha, x, rho, w);
                        keep = !use_ratio || (fabs(rho) < limit);
                        if (keep) {
                            ++cnt[0];
                            acc[27] = __dadd_rn(acc[27], rho);
                            if (GRAD) {
                                double X, Y, Z;
                                transform_point(R, t, ptx, pty, ptz, X, Y, Z);
                                contract_point(fx, fy, X, Y, Z, w, rc0.y, rc1.x, rc1.y, rc2.x, rc2.y, acc);
                            }
                        }
                    }
                    if (c.h_kept) c.h_kept[n] = keep ? 1 : 0;
                    if (c.inl_out) c.inl_out[n] = keep ? 1 : 0;
                }
            }
            // the 32 lane partials of each of the 28 sums -> lane i (transposition, two rounds of 14 sums through the
            // same scratch), pairwise tree over the lanes
            double tot = 0.0;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                __syncwarp();  // the rows of the previous round / range have been read
#pragma unroll
                for (int i = 0; i < kHalfSums; ++i) wscr[i * kLaneStride + lane] = acc[h * kHalfSums + i];
                __syncwarp();
                if (lane >= h * kHalfSums && lane < (h + 1) * kHalfSums) {
                    const double2* row = reinterpret_cast<const double2*>(wscr + (lane - h * kHalfSums) * kLaneStride);
                    double x[16];
#pragma unroll
                    for (int k = 0; k < 16; ++k) {
                        const double2 v = row[k];
                        x[k] = __dadd_rn(v.x, v.y);
                    }
#pragma unroll
                    for (int w = 1; w < 16; w <<= 1) {
#pragma unroll
                        for (int k = 0; k + w < 16; k += 2 * w) x[k] = __dadd_rn(x[k], x[k + w]);
                    }
                    tot = x[0];
                }
            }
            if (lane < 28) {
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
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) cnt[i] += __shfl_xor_sync(0xffffffffu, cnt[i], off);
    }
    if (lane == 0) {
        sh.red_i[warp][0] = cnt[0];
        sh.red_i[warp][1] = cnt[1];
        sh.red_i[warp][2] = cnt[2];
    }
    __syncthreads();
    int32_t* iout = &sh.ev_count;  // ev_count, ev_nsup, ev_ngath are adjacent
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
    } else if (tid >= 32 && tid < 35) {
        int32_t tot_i = 0;
        for (int w = 0; w < kWarps; ++w) tot_i += sh.red_i[w][tid - 32];
        if (c.part_out) {  // two int64 slots: survivors; supported (low half) + gathered (high half), see combine_parts
            if (tid == 33) {
                int32_t ng = 0;
                for (int w = 0; w < kWarps; ++w) ng += sh.red_i[w][2];
                reinterpret_cast<long long*>(c.part_out)[29] = (long long)tot_i + ((long long)ng << 32);
            } else if (tid == 32) {
                reinterpret_cast<long long*>(c.part_out)[28] = tot_i;
            }
        } else if (c.csize > 1)
            sh.cpart_i[rp.cpar][tid - 32] = tot_i;
        else
            iout[tid - 32] = tot_i;
    }
    if (c.part_out) {  // the CTA that completes the evaluation combines the tasks' partials (combine_parts)
        __syncthreads();
        return;
    }
    if (c.csize > 1)
        cluster_combine<false>(sh, c.csize, rp.cpar, 28, sh.ev, 3, iout);
    else
        __syncthreads();
    if (tid == 0) sh.ev_cost = sh.ev[27] / (double)sh.ev_count;  // mean; 0/0 = NaN like torch's empty mean
    __syncthreads();
}

}  // namespace fmpnp
