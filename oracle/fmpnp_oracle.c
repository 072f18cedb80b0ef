/*
 * fmpnp_oracle.c -- plain-C, float64, CHW restatement of the reference optimizer.
 *
 * TEST INFRASTRUCTURE ONLY (see fmpnp_oracle.h).  It deliberately keeps the REFERENCE's
 * structure -- project twice per iteration, gather F / Gx / Gy / F(trial) separately from CHW
 * maps, materialise the per-point C x 6 Jacobian, LU solve -- so that it shares no design
 * decision with the CUDA product it checks.
 *
 * Each function cites the reference lines it restates (paths relative to /root/reference).
 * Compile with -ffp-contract=off: the only fused multiply-adds are the explicit fma() calls
 * that reproduce what torch.mm does on the reference's operand layouts (see project()).
 */
#include "fmpnp_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

#define EPS_F32 1.1920928955078125e-07 /* torch.finfo(torch.float32).eps */

/* ------------------------------------------------------------------------------------------
 * double -> int32 as torch's `.type(torch.IntTensor)` does on x86-64 (cvttsd2si): values
 * outside the int32 range and NaN become INT32_MIN ("integer indefinite").
 * ---------------------------------------------------------------------------------------- */
static int32_t cast_i32_x86(double r) {
    if (!(r > -2147483649.0 && r < 2147483648.0)) return INT32_MIN;
    return (int32_t)r;
}

/* model.py:303-308 (again :428-433, :134-135, :217-218):
 *   points_3d = torch.mm(R, pts3D.T).T + t
 *   points_2d = torch.round(from_homogeneous(torch.mm(K, points_3d.T).T)).type(IntTensor) - 1
 * Rounding order of the two torch.mm calls, measured in the build container (torch 2.11 CPU,
 * MKL): the first (rhs = transposed view of a contiguous N x 3) accumulates k = 0,1,2 with
 * separately rounded products and NO fma; the second (rhs contiguous 3 x N) is the fma chain
 * fma(K[i][2], Z, fma(K[i][1], Y, K[i][0] * X)).  from_homogeneous: helpers/utils.py:122-129.
 * torch.round = round-half-to-even = rint() in the default rounding mode. The "- 1" is int32
 * arithmetic (wraps). */
void oracle_project(const double *pts3D, int32_t N, const double R[9], const double t[3], const double K[9],
                    double *P, int32_t *px) {
    for (int32_t n = 0; n < N; ++n) {
        const double x = pts3D[3 * n + 0], y = pts3D[3 * n + 1], z = pts3D[3 * n + 2];
        double Pc[3];
        for (int i = 0; i < 3; ++i) {
            double a = R[3 * i + 0] * x;
            double b = R[3 * i + 1] * y;
            double c = R[3 * i + 2] * z;
            Pc[i] = ((a + b) + c) + t[i];
        }
        double q[3];
        for (int i = 0; i < 3; ++i) q[i] = fma(K[3 * i + 2], Pc[2], fma(K[3 * i + 1], Pc[1], K[3 * i + 0] * Pc[0]));
        const double u = q[0] / q[2], v = q[1] / q[2];
        P[3 * n + 0] = Pc[0];
        P[3 * n + 1] = Pc[1];
        P[3 * n + 2] = Pc[2];
        px[2 * n + 0] = (int32_t)((uint32_t)cast_i32_x86(rint(u)) - 1u);
        px[2 * n + 1] = (int32_t)((uint32_t)cast_i32_x86(rint(v)) - 1u);
    }
}

/* model.py:99-117 points_within_image, pad = 0; x against width, y against height */
static int within_image(const int32_t *px, int32_t width, int32_t height) {
    return px[0] >= 0 && px[1] >= 0 && px[0] < width && px[1] < height;
}

/* model.py:74-97 indexing_, called with flip(points) so column 0 is y:
 *   row = floor(double(y) * Hf / height) -> int16 ; col = floor(double(x) * Wf / width) -> int16 */
static void cell_of(int32_t x, int32_t y, int32_t Hf, int32_t Wf, int32_t width, int32_t height, int32_t *row,
                    int32_t *col) {
    *row = (int16_t)floor(((double)y * (double)Hf) / (double)height);
    *col = (int16_t)floor(((double)x * (double)Wf) / (double)width);
}

/* helpers/utils.py:15-78: returns rho (loss) and w (first derivative) */
static void loss1(int32_t id, double alpha, double x, double *rho, double *w) {
    switch (id) {
    case ORACLE_LOSS_SQUARED: /* :15-17 */
        *rho = x;
        *w = 1.0;
        return;
    case ORACLE_LOSS_HUBER: { /* :19-29, rho = 1 */
        const int m = x <= 1.0;
        const double sx = sqrt(x);
        const double isx = fmax(EPS_F32, 1.0 / sx);
        *rho = m ? x : 2.0 * sx - 1.0;
        *w = m ? 1.0 : isx;
        return;
    }
    case ORACLE_LOSS_CAUCHY: /* :31-34 -> barron(alpha = 0) */
        alpha = 0.0;
        break;
    case ORACLE_LOSS_GEMAN_MCCLURE: /* :36-38 -> barron(alpha = -2) */
        alpha = -2.0;
        break;
    default:
        break;
    }
    /* barron_loss :40-78 */
    if (alpha == 0.0) {
        *rho = 2.0 * log1p(fmin(0.5 * x, 33e37));
        *w = 2.0 / (x + 2.0);
    } else if (alpha == 2.0) {
        *rho = x;
        *w = 1.0;
    } else {
        const double beta_safe = fmax(EPS_F32, fabs(alpha - 2.0));
        const double alpha_safe = (alpha >= 0.0 ? 1.0 : -1.0) * fmax(EPS_F32, fabs(alpha));
        *rho = 2.0 * (beta_safe / alpha_safe) * (pow(x / beta_safe + 1.0, 0.5 * alpha) - 1.0);
        *w = pow(x / beta_safe + 1.0, 0.5 * alpha - 1.0);
    }
}

void oracle_loss(int32_t loss_id, double alpha, const double *x, int32_t n, double *rho, double *w) {
    for (int32_t i = 0; i < n; ++i) loss1(loss_id, alpha, x[i], &rho[i], &w[i]);
}

/* model.py:37-72 optimizer_step. torch.lu / lu_solve = LAPACK getrf / getrs: partial pivoting
 * on the largest |entry| of the column (first occurrence), multipliers formed with the
 * reciprocal of the pivot, then unit-lower forward and upper backward substitution.
 * Returns 0, or 1 if a zero pivot was met (LAPACK "info > 0"; torch carries on and the
 * division produces inf/nan exactly as here). */
int oracle_optimizer_step(const double g[6], const double Hin[36], double lambda_, double lr, double delta[6]) {
    double A[36], b[6];
    int singular = 0;
    memcpy(A, Hin, sizeof(A));
    if (lambda_ != 0.0) { /* `if lambda_:` :46-48 */
        for (int i = 0; i < 6; ++i) A[7 * i] = A[7 * i] + (Hin[7 * i] + 1e-9) * lambda_;
    }
    for (int i = 0; i < 6; ++i) b[i] = g[i];
    for (int j = 0; j < 6; ++j) {
        int p = j;
        double best = fabs(A[6 * j + j]);
        for (int i = j + 1; i < 6; ++i)
            if (fabs(A[6 * i + j]) > best) {
                best = fabs(A[6 * i + j]);
                p = i;
            }
        if (p != j) {
            for (int k = 0; k < 6; ++k) {
                double tmp = A[6 * j + k];
                A[6 * j + k] = A[6 * p + k];
                A[6 * p + k] = tmp;
            }
            double tb = b[j];
            b[j] = b[p];
            b[p] = tb;
        }
        const double piv = A[6 * j + j];
        if (piv == 0.0) {
            singular = 1;
            continue;
        }
        const double rp = 1.0 / piv;
        for (int i = j + 1; i < 6; ++i) {
            const double l = A[6 * i + j] * rp;
            A[6 * i + j] = l;
            for (int k = j + 1; k < 6; ++k) A[6 * i + k] = A[6 * i + k] - l * A[6 * j + k];
        }
    }
    /* L y = P b */
    for (int i = 1; i < 6; ++i) {
        double s = b[i];
        for (int k = 0; k < i; ++k) s = s - A[6 * i + k] * b[k];
        b[i] = s;
    }
    /* U x = y */
    for (int i = 5; i >= 0; --i) {
        double s = b[i];
        for (int k = i + 1; k < 6; ++k) s = s - A[6 * i + k] * b[k];
        b[i] = s / A[6 * i + i];
    }
    for (int i = 0; i < 6; ++i) delta[i] = -lr * b[i]; /* :62 */
    return singular;
}

/* helpers/utils.py:174-183 */
static void skew(const double v[3], double M[9]) {
    M[0] = 0.0;
    M[1] = -v[2];
    M[2] = v[1];
    M[3] = v[2];
    M[4] = 0.0;
    M[5] = -v[0];
    M[6] = -v[1];
    M[7] = v[0];
    M[8] = 0.0;
}

static void mat3_mul(const double A[9], const double B[9], double C[9]) {
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            C[3 * i + j] = (A[3 * i + 0] * B[0 + j] + A[3 * i + 1] * B[3 + j]) + A[3 * i + 2] * B[6 + j];
}

/* helpers/utils.py:209-221 so3exp_map */
void oracle_so3exp(const double w[3], double dr[9]) {
    const double theta = sqrt((w[0] * w[0] + w[1] * w[1]) + w[2] * w[2]);
    double wn[3] = {w[0] / theta, w[1] / theta, w[2] / theta};
    double W[9], WW[9];
    skew(wn, W);
    mat3_mul(W, W, WW);
    const double s = sin(theta), c1 = 1.0 - cos(theta);
    for (int i = 0; i < 9; ++i) {
        double res = W[i] * s + WW[i] * c1;
        if (theta < 1e-12) res = 0.0; /* torch.where(theta < 1e-12, 0, res): also swallows the 0/0 NaNs */
        dr[i] = ((i % 4) == 0 ? 1.0 : 0.0) + res;
    }
}

/* ------------------------------------------------------------------------------------------
 * Residual evaluation at one pose: the block the reference repeats at :303-339, :428-462,
 * :217-241 and :134-147.
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    double *P;     /* N x 3 */
    int32_t *px;   /* N x 2 */
    uint8_t *mask; /* N: supported */
    uint8_t *kept; /* N: supported and survived ratio test (== mask when no ratio test) */
    int32_t *idx;  /* survivors' point indices, length Nk */
    double *err;   /* Nk x C, survivors only */
    double *rho;   /* Nk */
    double *w;     /* Nk */
    int32_t Ns, Nk;
} eval_buf;

static int eval_alloc(eval_buf *e, int32_t N, int32_t C) {
    memset(e, 0, sizeof(*e));
    e->P = (double *)malloc(sizeof(double) * 3 * (size_t)(N ? N : 1));
    e->px = (int32_t *)malloc(sizeof(int32_t) * 2 * (size_t)(N ? N : 1));
    e->mask = (uint8_t *)malloc((size_t)(N ? N : 1));
    e->kept = (uint8_t *)malloc((size_t)(N ? N : 1));
    e->idx = (int32_t *)malloc(sizeof(int32_t) * (size_t)(N ? N : 1));
    e->err = (double *)malloc(sizeof(double) * (size_t)(N ? N : 1) * (size_t)(C ? C : 1));
    e->rho = (double *)malloc(sizeof(double) * (size_t)(N ? N : 1));
    e->w = (double *)malloc(sizeof(double) * (size_t)(N ? N : 1));
    return !(e->P && e->px && e->mask && e->kept && e->idx && e->err && e->rho && e->w);
}

static void eval_free(eval_buf *e) {
    free(e->P);
    free(e->px);
    free(e->mask);
    free(e->kept);
    free(e->idx);
    free(e->err);
    free(e->rho);
    free(e->w);
}

/* gather one map at the survivors' cells: indexing_ (model.py:74-97) */
static void gather_map(const oracle_problem *p, const double *map, const eval_buf *e, double *out /*Nk x C*/) {
    const size_t HW = (size_t)p->Hf * (size_t)p->Wf;
    for (int32_t k = 0; k < e->Nk; ++k) {
        const int32_t n = e->idx[k];
        int32_t row, col;
        cell_of(e->px[2 * n], e->px[2 * n + 1], p->Hf, p->Wf, p->im_width, p->im_height, &row, &col);
        const double *src = map + (size_t)row * (size_t)p->Wf + (size_t)col;
        double *dst = out + (size_t)k * (size_t)p->C;
        for (int32_t c = 0; c < p->C; ++c) dst[c] = src[(size_t)c * HW];
    }
}

/* ratio_mode: 0 = none, 1 = ratio test on loss_fn(cost).rho (forward, find_inliers),
 *             2 = ratio test on the raw SSD cost (compute_cost :231-233).
 * After the call: e->err holds the survivors' residual rows, e->rho / e->w = loss_fn(0.5*sum e^2)
 * of the survivors (for ratio_mode 2 and for cost-only users rho = raw SSD when loss_id == squared).
 * Returns Ns (supported count); 0 means "no supported point" and nothing else is valid. */
static int32_t eval_pose(const oracle_problem *p, int32_t loss_id, double alpha, int ratio_mode, double thr,
                         const double R[9], const double t[3], eval_buf *e) {
    const int32_t N = p->N, C = p->C;
    oracle_project(p->pts3D, N, R, t, p->K, e->P, e->px);
    int32_t Ns = 0;
    for (int32_t n = 0; n < N; ++n) {
        e->mask[n] = (uint8_t)within_image(&e->px[2 * n], p->im_width, p->im_height); /* :311 */
        e->kept[n] = 0;
        if (e->mask[n]) e->idx[Ns++] = n; /* :314-315 */
    }
    e->Ns = Ns;
    e->Nk = Ns;
    if (Ns == 0) return 0;
    /* error = indexing_(F, flip(px_s)) - feature_ref[mask]   (:322) */
    gather_map(p, p->F, e, e->err);
    for (int32_t k = 0; k < Ns; ++k) {
        const double *r = p->feature_ref + (size_t)e->idx[k] * (size_t)p->ref_ld;
        double *d = e->err + (size_t)k * (size_t)C;
        for (int32_t c = 0; c < C; ++c) d[c] = d[c] - r[c];
    }
    if (ratio_mode) { /* :324-334 / :231-236 / :146-150 ; ratio_threshold_feature_errors :120-129 */
        double mx = -INFINITY;
        int any_nan = 0;
        for (int32_t k = 0; k < Ns; ++k) {
            const double *d = e->err + (size_t)k * (size_t)C;
            double s = 0.0;
            for (int32_t c = 0; c < C; ++c) s += d[c] * d[c];
            const double cost = 0.5 * s;
            double rho, w;
            if (ratio_mode == 1)
                loss1(loss_id, alpha, cost, &rho, &w);
            else
                rho = cost;
            e->rho[k] = rho;
            const double a = fabs(rho);
            if (a != a) any_nan = 1;
            if (a > mx) mx = a;
        }
        const double limit = (any_nan ? NAN : mx) * thr; /* torch.max propagates NaN */
        int32_t Nk = 0;
        for (int32_t k = 0; k < Ns; ++k) {
            if (fabs(e->rho[k]) < limit) {
                if (Nk != k) memmove(e->err + (size_t)Nk * C, e->err + (size_t)k * C, sizeof(double) * (size_t)C);
                e->idx[Nk] = e->idx[k];
                ++Nk;
            }
        }
        e->Nk = Nk;
    }
    for (int32_t k = 0; k < e->Nk; ++k) { /* :338-339 */
        const double *d = e->err + (size_t)k * (size_t)C;
        double s = 0.0;
        for (int32_t c = 0; c < C; ++c) s += d[c] * d[c];
        const double cost = 0.5 * s;
        if (ratio_mode == 2) {
            e->rho[k] = cost;
            e->w[k] = 1.0;
        } else {
            loss1(loss_id, alpha, cost, &e->rho[k], &e->w[k]);
        }
        e->kept[e->idx[k]] = 1;
    }
    return Ns;
}

static double mean_of(const double *x, int32_t n) {
    double s = 0.0;
    for (int32_t i = 0; i < n; ++i) s += x[i];
    return s / (double)n; /* n == 0 -> nan, like torch's mean of an empty tensor */
}

static void track_push(oracle_track *trk, int32_t N, const double R[9], const double t[3], double cost,
                       const eval_buf *e, int use_ratio) {
    if (!trk || trk->len >= trk->cap) return;
    const int32_t i = trk->len++;
    if (trk->Rs) memcpy(trk->Rs + 9 * (size_t)i, R, 9 * sizeof(double));
    if (trk->ts) memcpy(trk->ts + 3 * (size_t)i, t, 3 * sizeof(double));
    if (trk->costs) trk->costs[i] = cost;
    if (trk->points2d) memcpy(trk->points2d + 2 * (size_t)N * i, e->px, 2 * (size_t)N * sizeof(int32_t));
    if (trk->mask) memcpy(trk->mask + (size_t)N * i, e->mask, (size_t)N);
    if (trk->thr_mask) memcpy(trk->thr_mask + (size_t)N * i, use_ratio ? e->kept : e->mask, (size_t)N);
}

int oracle_forward(const oracle_problem *p, const oracle_opts *o, const double R0[9], const double t0[3],
                   oracle_result *out, oracle_track *trk) {
    const int32_t N = p->N, C = p->C;
    eval_buf cur, tri;
    if (eval_alloc(&cur, N, C) || eval_alloc(&tri, N, C)) return -1;
    double *gx = (double *)malloc(sizeof(double) * (size_t)(N ? N : 1) * (size_t)(C ? C : 1));
    double *gy = (double *)malloc(sizeof(double) * (size_t)(N ? N : 1) * (size_t)(C ? C : 1));
    double *J = (double *)malloc(sizeof(double) * (size_t)(C ? C : 1) * 6);
    if (!gx || !gy || !J) return -1;

    double R[9], t[3], Rb[9], tb[3];
    memcpy(R, R0, sizeof(R));
    memcpy(t, t0, sizeof(t));
    memcpy(Rb, R0, sizeof(R));
    memcpy(tb, t0, sizeof(t));
    double lambda_ = o->lambda0, lr = 1.0; /* :281-283 */
    double prev_cost = NAN;
    const int ratio_mode = o->use_ratio_test ? 1 : 0;
    memset(out, 0, sizeof(*out));
    out->status = ORACLE_OK;
    out->first_cost = NAN;
    out->best_cost = NAN;
    const double fx = p->K[0], fy = p->K[4];

    for (int32_t i = 0; i < o->n_iters; ++i) { /* :300 */
        const int32_t Ns = eval_pose(p, o->loss_id, o->barron_alpha, ratio_mode, o->ratio_threshold, R, t, &cur);
        if (Ns == 0) { /* :316-320 -> CURRENT pose */
            out->status = ORACLE_EMPTY_SUPPORT;
            memcpy(out->R, R, sizeof(R));
            memcpy(out->t, t, sizeof(t));
            goto done;
        }
        if (i == 0) { /* :347-359 */
            prev_cost = mean_of(cur.rho, cur.Nk);
            out->best_cost = prev_cost;
            out->first_cost = prev_cost;
            out->best_num_inliers = cur.Nk;
            out->has_best = 1;
            memcpy(Rb, R, sizeof(R));
            memcpy(tb, t, sizeof(t));
            track_push(trk, N, R, t, prev_cost, &cur, ratio_mode);
        }
        /* :369-405 Jacobians, gradient, Gauss-Newton Hessian */
        gather_map(p, p->Gx, &cur, gx); /* :386 */
        gather_map(p, p->Gy, &cur, gy); /* :387 */
        double Grad[6] = {0, 0, 0, 0, 0, 0}, Hess[36];
        memset(Hess, 0, sizeof(Hess));
        for (int32_t k = 0; k < cur.Nk; ++k) {
            const double *Pn = cur.P + 3 * (size_t)cur.idx[k];
            const double X = Pn[0], Y = Pn[1], Z = Pn[2];
            /* J_p_T = [I3 | -skew(P)]   :369-370 */
            double S[9], JpT[18];
            skew(Pn, S);
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) {
                    JpT[6 * r + c] = (r == c) ? 1.0 : 0.0;
                    JpT[6 * r + 3 + c] = -S[3 * r + c];
                }
            /* J_px_p  :377-382 */
            double Jpx[6];
            Jpx[0] = (fx * 1.0) / Z;
            Jpx[1] = 0.0 / Z;
            Jpx[2] = ((-fx * X) / Z) / Z;
            Jpx[3] = 0.0 / Z;
            Jpx[4] = (fy * 1.0) / Z;
            Jpx[5] = ((-fy * Y) / Z) / Z;
            const double *e = cur.err + (size_t)k * C;
            const double *gxk = gx + (size_t)k * C, *gyk = gy + (size_t)k * C;
            double gk[6] = {0, 0, 0, 0, 0, 0}, Hk[36];
            memset(Hk, 0, sizeof(Hk));
            for (int32_t c = 0; c < C; ++c) { /* J_e_T = (J_f_px @ J_px_p) @ J_p_T   :390-394 */
                double a3[3];
                for (int m = 0; m < 3; ++m) a3[m] = gxk[c] * Jpx[m] + gyk[c] * Jpx[3 + m];
                double *Jc = J + 6 * (size_t)c;
                for (int m = 0; m < 6; ++m) Jc[m] = (a3[0] * JpT[m] + a3[1] * JpT[6 + m]) + a3[2] * JpT[12 + m];
                for (int m = 0; m < 6; ++m) gk[m] += Jc[m] * e[c]; /* einsum 'bij,bi->bj' :397 */
                for (int m = 0; m < 6; ++m)
                    for (int l = 0; l < 6; ++l) Hk[6 * m + l] += Jc[m] * Jc[l]; /* einsum 'ijk,ijl->ikl' :403 */
            }
            const double wk = cur.w[k];
            for (int m = 0; m < 6; ++m) Grad[m] += wk * gk[m];    /* :398-399 */
            for (int m = 0; m < 36; ++m) Hess[m] += wk * Hk[m];   /* :404-405 */
        }
        double delta[6];
        oracle_optimizer_step(Grad, Hess, lambda_, lr, delta); /* :408 */
        int has_nan = 0;
        for (int m = 0; m < 6; ++m)
            if (delta[m] != delta[m]) has_nan = 1;
        if (has_nan) { /* :411-413 intended `break` (the reference would raise NameError on `logging`) */
            out->status = ORACLE_NAN_STEP;
            break;
        }
        double dr[9], Rn[9], tn[3];
        oracle_so3exp(delta + 3, dr); /* :416-419 */
        mat3_mul(dr, R, Rn);          /* :425 */
        for (int r = 0; r < 3; ++r)   /* :426 */
            tn[r] = ((dr[3 * r] * t[0] + dr[3 * r + 1] * t[1]) + dr[3 * r + 2] * t[2]) + delta[r];
        const int32_t Ns2 = eval_pose(p, o->loss_id, o->barron_alpha, ratio_mode, o->ratio_threshold, Rn, tn, &tri);
        if (Ns2 == 0) { /* :441-445 -> CURRENT pose */
            out->status = ORACLE_EMPTY_SUPPORT;
            memcpy(out->R, R, sizeof(R));
            memcpy(out->t, t, sizeof(t));
            goto done;
        }
        const double new_cost = mean_of(tri.rho, tri.Nk); /* :460-462 */
        out->iters_run = i + 1;
        track_push(trk, N, Rn, tn, new_cost, &tri, ratio_mode); /* :464-465 */
        { /* :469-470  np.clip(lambda_ * (10 if new > prev else 1/10), 1e-6, 1e4) */
            double l = lambda_ * (new_cost > prev_cost ? 10.0 : (1.0 / 10.0));
            lambda_ = l < 1e-6 ? 1e-6 : (l > 1e4 ? 1e4 : l);
        }
        if (new_cost > prev_cost) { /* :472-476 */
            double l = 0.1 * lr;
            lr = l < 1e-3 ? 1e-3 : (l > 1.0 ? 1.0 : l);
            continue;
        } else { /* :477-483 */
            lr = 1.0;
            if (new_cost < out->best_cost) {
                memcpy(Rb, Rn, sizeof(R));
                memcpy(tb, tn, sizeof(t));
                out->best_num_inliers = tri.Nk;
                out->best_cost = new_cost;
            }
        }
        prev_cost = new_cost; /* :484-486 */
        memcpy(R, Rn, sizeof(R));
        memcpy(t, tn, sizeof(t));
    }
    memcpy(out->R, Rb, sizeof(R)); /* :491-494 */
    memcpy(out->t, tb, sizeof(t));
done:
    eval_free(&cur);
    eval_free(&tri);
    free(gx);
    free(gy);
    free(J);
    return out->status;
}

int oracle_compute_cost(const oracle_problem *p, const oracle_opts *o, const double R[9], const double t[3],
                        double *cost) {
    eval_buf e;
    if (eval_alloc(&e, p->N, p->C)) return -1;
    const int32_t Ns = eval_pose(p, ORACLE_LOSS_SQUARED, 0.0, o->use_ratio_test ? 2 : 0, o->ratio_threshold, R, t, &e);
    int rc = 0;
    if (Ns == 0) {
        rc = 1; /* None  :224-225 */
        *cost = NAN;
    } else {
        *cost = mean_of(e.rho, e.Nk); /* :242-243 */
    }
    eval_free(&e);
    return rc;
}

int oracle_find_inliers(const oracle_problem *p, int32_t loss_id, double barron_alpha, double threshold,
                        const double R[9], const double t[3], uint8_t *inliers) {
    eval_buf e;
    if (eval_alloc(&e, p->N, p->C)) return -1;
    eval_pose(p, loss_id, barron_alpha, 1, threshold, R, t, &e); /* :134-150 */
    memcpy(inliers, e.kept, (size_t)p->N);                       /* mask_supported[mask_supported] = threshold_mask */
    eval_free(&e);
    return 0;
}

int oracle_multilevel(const oracle_problem *p, const oracle_opts *o, const int32_t *levels, int32_t n_levels,
                      const double R0[9], const double t0[3], oracle_result *out, double *initial_cost,
                      oracle_result *lvl, oracle_track *trk) {
    double R[9], t[3];
    memcpy(R, R0, sizeof(R));
    memcpy(t, t0, sizeof(t));
    memset(out, 0, sizeof(*out));
    if (oracle_compute_cost(p, o, R, t, initial_cost)) { /* :182-187 */
        memcpy(out->R, R, sizeof(R));
        memcpy(out->t, t, sizeof(t));
        out->status = ORACLE_EMPTY_SUPPORT;
        out->best_cost = NAN;
        out->first_cost = NAN;
        return 1;
    }
    const size_t HW = (size_t)p->Hf * (size_t)p->Wf;
    for (int32_t l = 0; l < n_levels; ++l) { /* :193-210 with target_size = kernel_size = None */
        const int32_t s = levels[2 * l], e = levels[2 * l + 1];
        oracle_problem q = *p;
        q.C = e - s;
        q.F = p->F + (size_t)s * HW;
        q.Gx = p->Gx + (size_t)s * HW;
        q.Gy = p->Gy + (size_t)s * HW;
        q.feature_ref = p->feature_ref + s;
        oracle_result r;
        oracle_forward(&q, o, R, t, &r, trk);
        if (lvl) lvl[l] = r;
        memcpy(R, r.R, sizeof(R));
        memcpy(t, r.t, sizeof(t));
        *out = r;
    }
    return 0;
}

typedef struct {
    const oracle_problem *ps;
    int32_t B;
    const oracle_opts *o;
    const int32_t *levels;
    int32_t n_levels;
    const double *R0s, *t0s;
    oracle_result *outs;
    int32_t next; /* shared work counter */
} batch_job;

static void *batch_worker(void *arg) {
    batch_job *j = (batch_job *)arg;
    for (;;) {
        const int32_t b = __atomic_fetch_add(&j->next, 1, __ATOMIC_RELAXED);
        if (b >= j->B) break;
        if (j->levels && j->n_levels > 0) {
            double ic;
            oracle_multilevel(&j->ps[b], j->o, j->levels, j->n_levels, j->R0s + 9 * (size_t)b, j->t0s + 3 * (size_t)b,
                              &j->outs[b], &ic, NULL, NULL);
        } else {
            oracle_forward(&j->ps[b], j->o, j->R0s + 9 * (size_t)b, j->t0s + 3 * (size_t)b, &j->outs[b], NULL);
        }
    }
    return NULL;
}

int oracle_forward_batch(const oracle_problem *ps, int32_t B, const oracle_opts *o, const int32_t *levels,
                         int32_t n_levels, const double *R0s, const double *t0s, oracle_result *outs,
                         int32_t n_threads) {
    batch_job job = {ps, B, o, levels, n_levels, R0s, t0s, outs, 0};
    if (n_threads < 1) n_threads = 1;
    if (n_threads > B) n_threads = B;
    if (n_threads <= 1) {
        batch_worker(&job);
        return 0;
    }
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)n_threads);
    if (!th) return -1;
    int32_t started = 0;
    for (int32_t i = 0; i < n_threads; ++i)
        if (pthread_create(&th[i], NULL, batch_worker, &job) == 0) ++started;
    if (started == 0) batch_worker(&job);
    for (int32_t i = 0; i < started; ++i) pthread_join(th[i], NULL);
    free(th);
    return 0;
}
