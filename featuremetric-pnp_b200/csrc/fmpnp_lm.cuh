// fmpnp_lm.cuh -- one pose evaluation, the damped solve + pose update, history records and the Levenberg-Marquardt bookkeeping shared by the two launch routes.
// Part of the refinement kernels (fmpnp_kernels.cu); not a public interface.
#pragma once
#include "fmpnp_reduce.cuh"

namespace fmpnp {

// GRAD: rows streamed (F, Gx, Gy, ref or F, ref); RGRAD: whether the per-point phase builds the normal equations
template <bool GRAD, bool RGRAD = GRAD, bool COUNT = true>
__device__ __forceinline__ void evaluate(EvalCtx& c, const double* R, const double* t, Shared& sh, RingPos& rp) {
    set_geometry(c, GRAD ? 4 : 2);
    rp.pos = 0;  // the ring is empty between evaluations; its depth may change with the slice
    stream_points<GRAD>(c, R, t, sh, rp);
    __syncthreads();
    reduce_points<RGRAD>(c, R, t, sh, rp);
    if (threadIdx.x == 0) {
        if (COUNT) sh.n_evals += 1;
        sh.gathered += (long long)sh.ev_ngath * c.Cl * (GRAD ? 4 : 2);
        sh.reused += (long long)(sh.ev_nsup - sh.ev_ngath) * c.Cl * (GRAD ? 4 : 2);
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

// ---------------------------------------------------------------------------------------------
// Levenberg-Marquardt bookkeeping on the state in shared memory (thread 0 only).  Shared by the two
// launch routes, so the reference's accept / reject semantics exist once.
// ---------------------------------------------------------------------------------------------
// i == 0 of a level: the evaluation at the incoming pose becomes the current linearisation (model.py:347-359)
__device__ __forceinline__ void lm_first_evaluation(Shared& sh, const LaunchParams& p, int qi) {
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

// the trial pose (Rn, tn) has been evaluated: record it, adapt the damping, accept or reject (model.py:464-486)
__device__ __forceinline__ void lm_trial_evaluation(Shared& sh, const LaunchParams& p, int qi) {
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

// end of a level (model.py:491-494): the best pose, unless the loop was left on an empty support (current pose)
__device__ __forceinline__ void lm_finish_level(Shared& sh, int lvl) {
    if (sh.flag != 2) {
        for (int i = 0; i < 9; ++i) sh.R[i] = sh.Rb[i];
        for (int i = 0; i < 3; ++i) sh.t[i] = sh.tb[i];
    }
    sh.level_status[lvl] = sh.status;
    sh.flag = 0;
}

__device__ __forceinline__ void write_result(const Shared& sh, const LaunchParams& p, int qi) {
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
    r.reused_rows = sh.reused;
    if (p.hist.enabled) p.hist.len[qi] = sh.hist_len;
}

}  // namespace fmpnp
