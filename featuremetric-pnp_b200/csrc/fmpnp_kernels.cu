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
#include <cstddef>
#include <cstdint>
#include <mutex>

#include <cooperative_groups.h>

#include "fmpnp_device.cuh"
#include "fmpnp_kernels.cuh"
#include "fmpnp_lm.cuh"      // evaluate, solve_and_update, lm_* (includes fmpnp_reduce.cuh, fmpnp_stream.cuh)
#include "fmpnp_light.cuh"   // light_evaluate

namespace cg = cooperative_groups;

namespace fmpnp {

// evaluate<true>(), or its light form when nearly every record is expected to be in place (fmpnp_light.cuh)
__device__ __forceinline__ void evaluate_grad(EvalCtx& c, const double* R, const double* t, Shared& sh, RingPos& rp, bool light) {
    if (light) {
        light_evaluate(c, R, t, sh, rp);
        if (threadIdx.x == 0) {
            sh.n_evals += 1;
            sh.gathered += (long long)sh.ev_ngath * c.Cl * 4;
            sh.reused += (long long)(sh.ev_nsup - sh.ev_ngath) * c.Cl * 4;
        }
    } else {
        evaluate<true>(c, R, t, sh, rp);
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
    uint8_t* inl = p.inl ? p.inl + (size_t)blockIdx.x * p.n_max : nullptr;

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
        c.row_pol = p.l2_policy ? l2_policy_evict_first() : l2_policy_normal();
        c.rec_pol = p.l2_policy ? l2_policy_evict_last() : l2_policy_normal();
        c.reuse_gen = 0u;
        c.active = q.active;
        {
            c.nranges = ranges_of(q.N);
            const int per = max(1, (q.N + c.nranges - 1) / c.nranges);
            c.range_len = per;
            c.n0 = min(q.N, crank * (c.nranges / csize) * per);
            c.n1 = min(q.N, (crank + 1) * (c.nranges / csize) * per);
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
            sh.reused = 0;
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
        if (p.inlier_on) {
            // fused find_inliers (model.py:131-152) at (R0, t0): the points it keeps are the ones refined
            // (optimize_feature_pnp.py:42-46).  Without a pyramid the pass streams all four rows and leaves
            // tagged records, so the i == 0 evaluation below gathers nothing.
            const bool share = p.n_levels == 0 && p.reuse;
            c.c0 = 0;
            c.Cl = q.C;
            c.aligned = slice_aligned(q, 0, q.C, share);
            c.loss_id = p.loss_id;
            c.alpha = p.alpha;
            c.ratio_mode = 1;
            c.thr = p.inlier_thr;
            c.inl_out = inl;
            c.reuse_gen = share ? p.reuse_base + (uint32_t)qi : 0u;  // level 0 of this query (one level)
            if (share)
                evaluate<true, false, false>(c, sh.R, sh.t, sh, rp);
            else
                evaluate<false, false, false>(c, sh.R, sh.t, sh, rp);
            __syncthreads();
            c.inl_out = nullptr;
            c.reuse_gen = 0u;
            c.active = inl;
        }
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
            c.reuse_gen = p.reuse ? p.reuse_base + (uint32_t)(qi * nlev + lvl) : 0u;  // records of this level only

            // ---- i == 0: cost and linearisation at the incoming pose (model.py:303-359)
            hist_slices(p, qi, sh.hist_len, c);
            const bool can_light = p.light != 0 && csize == 1 && c.reuse_gen != 0u;
            evaluate_grad(c, sh.R, sh.t, sh, rp, can_light && p.inlier_on && p.n_levels == 0);  // the inlier pass left every record
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
                lm_first_evaluation(sh, p, qi);
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
                // (light only while the previous evaluation found nearly every record in place)
                evaluate_grad(c, sh.Rn, sh.tn, sh, rp,
                              can_light && (it == 0 || (long long)sh.ev_ngath * p.light_miss_den <= (long long)sh.ev_nsup));
                __syncthreads();
                if (sh.ev_nsup == 0) {  // model.py:441-445 -> CURRENT pose
                    if (tid == 0) {
                        sh.status = FMPNP_STATUS_EMPTY_SUPPORT;
                        sh.flag = 2;
                    }
                    break;
                }
                if (tid == 0) {
                    lm_trial_evaluation(sh, p, qi);
                }
                __syncthreads();
            }
            __syncthreads();
            if (tid == 0) {
                lm_finish_level(sh, lvl);
            }
            __syncthreads();
        }
        if (tid == 0 && crank == 0) {
            write_result(sh, p, qi);
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
__device__ __forceinline__ void enqueue_evaluation(const SchedDev& s, int q, int units) {
    __threadfence();  // state (and, for the finisher, everything before) visible before the tasks are
    const unsigned int base = atomicAdd(s.tail, (unsigned int)units);
    for (int u = 0; u < units; ++u) {
        const unsigned int pos = base + (unsigned int)u;
        const unsigned long long v = ((unsigned long long)pos << 32) | (unsigned int)(q * kRanges + u + 1);
        volatile unsigned long long* slot = reinterpret_cast<volatile unsigned long long*>(&s.queue[pos & (s.cap - 1)]);
        // A slot is written only while it is empty: its consumer of the previous lap (ticket pos - cap) hands it back
        // after reading it.  The FIFO holds twice the tasks that can be queued at once, so this never waits in
        // practice; it is what makes overwriting an entry that has not been taken yet impossible, not just unlikely.
        unsigned long long t0 = 0;
        for (uint32_t spin = 0; *slot != kQueueEmpty; ++spin) {
            if (ld_volatile_u32(s.error) != 0u) return;
            if ((spin & 1023u) == 1023u) {
                const unsigned long long now = global_timer_ns();
                if (t0 == 0) t0 = now;
                if (s.wait_limit_ns != 0 && now - t0 > s.wait_limit_ns) {  // give up instead of hanging
                    atomicExch(s.error, 1u);
                    return;
                }
            }
            __nanosleep(100);
        }
        *slot = v;
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
            *reinterpret_cast<volatile unsigned long long*>(const_cast<unsigned long long*>(slot)) = kQueueEmpty;  // hand the slot back
            return (int)(unsigned int)(v & 0xffffffffu) - 1;
        }
        if (ld_volatile_u32(s.done) >= (unsigned int)B || ld_volatile_u32(s.error) != 0u) return -1;
        if ((spin & 1023u) == 1023u) {
            const unsigned long long now = global_timer_ns();
            if (t0 == 0) t0 = now;
            if (s.wait_limit_ns != 0 && now - t0 > s.wait_limit_ns) {  // give up instead of hanging
                atomicExch(s.error, 1u);
                return -1;
            }
        }
        __nanosleep(200);
    }
}

// warp 0: query state global <-> shared.  The doubles R .. initial_cost are laid out in the same order in QueryState
// and in Shared's copy (R, t, Rn, tn, Rb, tb, lin, then the scalars), so the lanes copy them as one array each.
constexpr int kStatePoseDoubles = 36 + 27;  // R, t, Rn, tn, Rb, tb, lin
__device__ __forceinline__ void load_state(Shared& sh, const volatile QueryState& g, int lane) {  // volatile: L1 may hold a stale copy
    static_assert(offsetof(QueryState, lin) == 36 * sizeof(double) && offsetof(Shared, lin) == 36 * sizeof(double), "layout");
    const volatile double* src = g.R;
    double* dst = sh.R;
    for (int i = lane; i < kStatePoseDoubles; i += 32) dst[i] = src[i];
    if (lane == 0) {
        sh.prev_cost = g.prev_cost;
        sh.best_cost = g.best_cost;
        sh.lambda = g.lambda;
        sh.lr = g.lr;
        sh.initial_cost = g.initial_cost;
        sh.gathered = g.gathered;
        sh.reused = g.reused;
    } else if (lane == 1) {
        sh.best_n = g.best_n;
        sh.has_best = g.has_best;
        sh.status = g.status;
        sh.flag = g.flag;
        sh.hist_len = g.hist_len;
        sh.n_evals = g.n_evals;
        sh.initial_set = g.initial_set;
    } else if (lane >= 2 && lane < 2 + FMPNP_MAX_LEVELS) {
        sh.level_status[lane - 2] = g.level_status[lane - 2];
    }
}

__device__ __forceinline__ void store_state(const Shared& sh, QueryState& g, int lane) {
    const double* src = sh.R;
    double* dst = g.R;
    for (int i = lane; i < kStatePoseDoubles; i += 32) dst[i] = src[i];
    if (lane == 0) {
        g.prev_cost = sh.prev_cost;
        g.best_cost = sh.best_cost;
        g.lambda = sh.lambda;
        g.lr = sh.lr;
        g.initial_cost = sh.initial_cost;
        g.gathered = sh.gathered;
        g.reused = sh.reused;
    } else if (lane == 1) {
        g.best_n = sh.best_n;
        g.has_best = sh.has_best;
        g.status = sh.status;
        g.flag = sh.flag;
        g.hist_len = sh.hist_len;
        g.n_evals = sh.n_evals;
        g.initial_set = sh.initial_set;
    } else if (lane >= 2 && lane < 2 + FMPNP_MAX_LEVELS) {
        g.level_status[lane - 2] = sh.level_status[lane - 2];
    }
    __threadfence();  // every lane's part before the queue entry that announces the state (enqueue_evaluation, lane 0)
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
        (&sh.ev_count)[tid - 32] = (int32_t)(sum & 0xffffffffll);          // survivors / supported
        if (tid == 33) sh.ev_ngath = (int32_t)(sum >> 32);                 // gathered rides in the high half of slot 29
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
        const int uq = min(sc.units, ranges_of(qd.N));  // tasks per evaluation of THIS query
        for (int i = 0; i < 9; ++i) g.R[i] = g.Rn[i] = g.Rb[i] = qd.R0[i];
        for (int i = 0; i < 3; ++i) g.t[i] = g.tn[i] = g.tb[i] = qd.t0[i];
        for (int i = 0; i < 27; ++i) g.lin[i] = 0.0;
        g.prev_cost = g.best_cost = NAN;
        g.lambda = p.lambda0;
        g.lr = 1.0;
        g.initial_cost = NAN;
        g.gathered = 0;
        g.reused = 0;
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
        if (MODE == kModeRefine && p.inlier_on)  // fused find_inliers first (streams all four rows when its records are reused)
            describe_evaluation(g, qd.R0, qd.t0, kPhaseInliers, (p.n_levels == 0 && p.reuse) ? 1 : 0, 0, qd.C, 1, p.loss_id, -1, uq);
        else if (MODE == kModeComputeCost)
            describe_evaluation(g, qd.R0, qd.t0, kPhaseCost, 0, 0, qd.C, p.use_ratio ? 2 : 0, FMPNP_LOSS_SQUARED, -1, uq);
        else if (MODE == kModeFindInliers)
            describe_evaluation(g, qd.R0, qd.t0, kPhaseCost, 0, 0, qd.C, 1, p.loss_id, -1, uq);
        else if (p.n_levels > 0)  // multilevel: compute_cost on ALL channels first (model.py:182)
            describe_evaluation(g, qd.R0, qd.t0, kPhaseCost, 0, 0, qd.C, p.use_ratio ? 2 : 0, FMPNP_LOSS_SQUARED, -1, uq);
        else
            describe_evaluation(g, qd.R0, qd.t0, kPhaseFirst, 1, 0, qd.C, p.use_ratio ? 1 : 0, p.loss_id, 0, uq);
        enqueue_evaluation(sc, q, uq);
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
        const int uq = min(sc.units, ranges_of(q.N));  // tasks per evaluation of this query
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
        c.row_pol = p.l2_policy ? l2_policy_evict_first() : l2_policy_normal();
        c.rec_pol = p.l2_policy ? l2_policy_evict_last() : l2_policy_normal();
        c.c0 = gv.ev_c0;
        c.Cl = gv.ev_Cl;
        c.loss_id = gv.ev_loss;
        c.alpha = p.alpha;
        c.ratio_mode = gv.ev_ratio_mode;
        c.inl_out = MODE == kModeFindInliers ? p.inliers_out + (size_t)qi * p.inliers_stride : nullptr;
        c.active = q.active;
        c.thr = p.ratio_thr;
        if (MODE == kModeRefine && p.inlier_on) {  // fused find_inliers: its pass writes the mask every later evaluation reads
            uint8_t* const inl = p.inl + (size_t)qi * p.n_max;
            if (gv.phase == kPhaseInliers) {
                c.inl_out = inl;
                c.thr = p.inlier_thr;
            } else {
                c.active = inl;
            }
        }
        const bool grad = gv.ev_grad != 0;             // rows streamed (the per-point phase of the inlier pass builds no normal equations)
        const int hist_entry = gv.ev_hist;
        c.reuse_gen = (MODE == kModeRefine && p.reuse && grad)
                          ? p.reuse_base + (uint32_t)(qi * (p.n_levels > 0 ? p.n_levels : 1) + gv.level)
                          : 0u;
        c.aligned = slice_aligned(q, c.c0, c.Cl, grad);
        c.h_px = nullptr;
        c.h_mask = nullptr;
        c.h_kept = nullptr;
        if (hist_entry >= 0) hist_slices(p, qi, hist_entry, c);
        {
            c.nranges = ranges_of(q.N);
            const int per = max(1, (q.N + c.nranges - 1) / c.nranges);
            const int rpu = c.nranges / uq;  // fixed ranges per task
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
        const bool multi = uq > 1, ratio = c.ratio_mode != 0;
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
            c.csize = uq;
            c.crank = u;
            c.part_out = parts + u * kPartStride;
            c.limit_given = ratio;
            if (grad && !(MODE == kModeRefine && c.inl_out != nullptr))
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
                for (int r = 0; r < uq; ++r) {
                    mm = fmax(mm, __ldcg(parts + r * kPartStride + 30));
                    nf |= __ldcg(reinterpret_cast<const long long*>(parts) + r * kPartStride + 31);
                }
                g.ev_limit = nf ? NAN : __dmul_rn(mm, c.thr);  // torch.max propagates NaN
                g.ev_round = 1;
                g.pending = uq;
                enqueue_evaluation(sc, qi, uq);
            }
            continue;
        }

        // ---- this CTA completed the evaluation: total the query, advance its state machine
        c.n0 = 0;
        c.n1 = q.N;
        if (warp == 0) load_state(sh, gv, lane);
        __syncthreads();
        if (multi)
            combine_parts(sh, parts, uq);
        else if (grad && !(MODE == kModeRefine && c.inl_out != nullptr))
            reduce_points<true>(c, evpose, evpose + 9, sh, rp);
        else
            reduce_points<false>(c, evpose, evpose + 9, sh, rp);

        int lvl = gv.level, it = gv.it, phase = gv.phase;
        bool ev_grad = grad;
        const int nlev = p.n_levels > 0 ? p.n_levels : 1;
        // Owner loop: the evaluations that follow are run right here, without the task queue, for as long as they are
        // expected to be light (LaunchParams::light); otherwise the next evaluation is described in the query's global
        // state and queued, and the loop is left after one pass.
        for (;;) {
            if (tid == 0) {
                if (!(MODE == kModeRefine && phase == kPhaseInliers)) sh.n_evals += 1;  // the inlier pass is not an LM evaluation
                sh.gathered += (long long)sh.ev_ngath * c.Cl * (ev_grad ? 4 : 2);
                sh.reused += (long long)(sh.ev_nsup - sh.ev_ngath) * c.Cl * (ev_grad ? 4 : 2);
            }
            __syncthreads();

            bool finished = false, have_eval = false;
            enum { kStartLevel, kTryStep, kFinishLevel, kCostEval } next = kStartLevel;
            // the evaluation that comes next (uniform across the CTA): phase, rows, slice, ratio mode, loss, pose
            int nd_phase = kPhaseCost, nd_grad = 0, nd_c0 = 0, nd_Cl = q.C, nd_ratio = 0, nd_loss = p.loss_id;
            bool nd_hist = false, nd_trial_pose = false;

            if (MODE == kModeComputeCost) {  // sparseFeaturePnP.compute_cost (model.py:216-243)
                if (tid == 0) {
                    p.cost_out[qi] = sh.ev_nsup > 0 ? sh.ev_cost : NAN;
                    if (p.nsup_out) p.nsup_out[qi] = sh.ev_nsup;
                }
                finished = true;
            } else if (MODE == kModeFindInliers) {  // find_inliers (model.py:131-152): the mask was written by reduce_points
                finished = true;
            } else if (phase == kPhaseInliers) {  // the mask is in `inl`: on to compute_cost (pyramid) or level 0
                lvl = 0;
                next = p.n_levels > 0 ? kCostEval : kStartLevel;
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
                    finished = lvl >= nlev;
                } else {
                    if (tid == 0) {
                        lm_first_evaluation(sh, p, qi);
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
                        lm_trial_evaluation(sh, p, qi);
                    }
                    ++it;
                    next = kTryStep;
                }
            }

            while (!finished && !have_eval) {
                __syncthreads();
                if (next == kCostEval) {  // multilevel: compute_cost on ALL channels (model.py:182), on the inliers
                    nd_phase = kPhaseCost;
                    nd_grad = 0;
                    nd_c0 = 0;
                    nd_Cl = q.C;
                    nd_ratio = p.use_ratio ? 2 : 0;
                    nd_loss = FMPNP_LOSS_SQUARED;
                    nd_hist = false;
                    nd_trial_pose = false;
                    have_eval = true;
                } else if (next == kStartLevel) {
                    nd_phase = kPhaseFirst;
                    nd_grad = 1;
                    nd_c0 = p.n_levels > 0 ? p.levels[lvl][0] : 0;
                    nd_Cl = p.n_levels > 0 ? p.levels[lvl][1] - p.levels[lvl][0] : q.C;
                    nd_ratio = p.use_ratio ? 1 : 0;
                    nd_loss = p.loss_id;
                    nd_hist = true;
                    nd_trial_pose = false;
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
                    nd_phase = kPhaseTrial;
                    nd_grad = 1;
                    nd_c0 = c.c0;  // the level's slice: a trial evaluation follows an evaluation of the same level
                    nd_Cl = c.Cl;
                    nd_ratio = p.use_ratio ? 1 : 0;
                    nd_loss = p.loss_id;
                    nd_hist = true;
                    nd_trial_pose = true;
                    have_eval = true;
                } else {  // kFinishLevel (model.py:491-494): best pose unless the loop was left on an empty support
                    if (tid == 0) {
                        lm_finish_level(sh, lvl);
                    }
                    ++lvl;
                    next = kStartLevel;
                    finished = lvl >= nlev;
                }
            }
            __syncthreads();
            if (finished) {
                if (tid == 0) {
                    if (MODE == kModeRefine) {
                        write_result(sh, p, qi);
                    }
                    __threadfence();
                    atomicAdd(sc.done, 1u);
                }
                break;
            }
            // light: a trial evaluation, or the first evaluation of the only level right after an inlier pass that left
            // its records (every tag matches) -- records are reused and nearly all of them are in place
            // ... as far as the evaluation that has just been completed can tell: one that re-used records and still had
            // to gather more than 1 / light_miss_den of its points says the query is moving across cells, and then the
            // next evaluation is better cut into tasks again (its rows streamed by several CTAs)
            const bool few_misses = phase != kPhaseTrial || (long long)sh.ev_ngath * p.light_miss_den <= (long long)sh.ev_nsup;
            const bool light = MODE == kModeRefine && p.light != 0 && p.reuse != 0 &&
                               ((nd_phase == kPhaseTrial && few_misses) ||
                                (nd_phase == kPhaseFirst && phase == kPhaseInliers && p.n_levels == 0));
            if (!light) {
                if (tid == 0) {
                    describe_evaluation(g, nd_trial_pose ? sh.Rn : sh.R, nd_trial_pose ? sh.tn : sh.t, nd_phase, nd_grad, nd_c0,
                                        nd_Cl, nd_ratio, nd_loss, nd_hist ? sh.hist_len : -1, uq);
                    g.level = lvl;
                    g.it = it;
                }
                if (warp == 0) {
                    store_state(sh, g, lane);
                    __syncwarp();
                    if (lane == 0) enqueue_evaluation(sc, qi, uq);
                }
                break;
            }
            // ---- the next evaluation, here and now: all points of the query, this CTA alone
            c.c0 = nd_c0;
            c.Cl = nd_Cl;
            c.loss_id = nd_loss;
            c.ratio_mode = nd_ratio;
            c.thr = p.ratio_thr;
            c.inl_out = nullptr;
            c.active = p.inlier_on ? p.inl + (size_t)qi * p.n_max : q.active;
            c.reuse_gen = p.reuse_base + (uint32_t)(qi * nlev + lvl);
            c.aligned = slice_aligned(q, c.c0, c.Cl, true);
            c.csize = 1;
            c.crank = 0;
            c.part_out = nullptr;
            c.limit_given = false;
            hist_slices(p, qi, sh.hist_len, c);
            light_evaluate(c, nd_trial_pose ? sh.Rn : sh.R, nd_trial_pose ? sh.tn : sh.t, sh, rp);
            phase = nd_phase;
            ev_grad = true;
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


// ---------------------------------------------------------------------------------------------
// compute_cost without any allocation (fmpnp_compute_cost_batch_noalloc; no counterpart in the reference).
// The per-CTA scratch of compute_cost_kernel -- records, support flags, the query counter -- lives in a
// file-scope device array of a fixed, modest size instead of the context's cudaMalloc'ed workspace, and the
// problem descriptors travel as a kernel parameter instead of through the context's staging buffers.  Every GPU
// that loads the library holds the array until the process exits; every call on a device shares it (calls that
// overlap on the GPU race), and it keeps the last call's contents: the only part that has to start at zero is
// the query counter, which the launcher clears on the caller's stream (flags are written before they are read,
// records are read only where the flag says so).
// ---------------------------------------------------------------------------------------------
struct StaticScratch {
    double rec[kStaticMaxQueries][kStaticMaxPoints][kRecDoubles];  // 2 MB
    uint8_t supp[kStaticMaxQueries][kStaticMaxPoints];             // 32 KB
    unsigned int counter;
};
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

__global__ void __launch_bounds__(kThreads, 2)
    compute_cost_static_kernel(const LaunchParams p, const __grid_constant__ StaticQueries pack) {
    LaunchParams lp = p;
    lp.queries = pack.q;  // the descriptors are read from the parameter space: nothing was staged in global memory
    persistent_body<kModeComputeCost>(lp);
}

// Set-up, once per device (with that device current): the device address of the scratch array -- in host code
// its name is only a stand-in -- and the dynamic shared memory the kernel may ask for.  Allocates nothing.
cudaError_t static_scratch_setup(int device, int ring_bytes, void** scratch) {
    constexpr int kMaxDevices = 64;
    static void* ptrs[kMaxDevices] = {nullptr};
    static std::mutex mu;
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (ptrs[device] == nullptr) {
        void* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_static_scratch);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(compute_cost_static_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ring_bytes);
        if (e != cudaSuccess) return e;
        ptrs[device] = ptr;
    }
    *scratch = ptrs[device];
    return cudaSuccess;
}

// p.rec / p.supp / p.work_counter are filled in here from `scratch` (the pointer static_scratch_setup returned)
cudaError_t launch_compute_cost_static(LaunchParams p, const StaticQueries& pack, void* scratch, int grid, int ring_bytes,
                                       cudaStream_t stream) {
    StaticScratch* s = static_cast<StaticScratch*>(scratch);
    p.rec = &s->rec[0][0][0];
    p.supp = &s->supp[0][0];
    p.inl = nullptr;
    p.n_max = kStaticMaxPoints;
    p.work_counter = &s->counter;
    cudaError_t e = cudaMemsetAsync(&s->counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    compute_cost_static_kernel<<<grid, kThreads, (size_t)ring_bytes, stream>>>(p, pack);
    return cudaGetLastError();
}

}  // namespace fmpnp
