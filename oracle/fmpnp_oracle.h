/*
 * fmpnp_oracle.h -- CPU restatement of the feature-metric PnP refinement of
 * aunagar/FeatureMetric-PnP (`featurePnP/model.py`, `featurePnP/helpers/utils.py`).
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under `featuremetric-pnp_b200/` may include, link or
 * call this.  It is used by `tests/`, by `__graft_entry__.smoke()` as the checker, and by
 * `bench.py` for the `cpu_baseline` leg / `--impl reference` arm.
 *
 * Layout and dtype follow the REFERENCE (not the product): feature maps are C x Hf x Wf
 * (CHW) float64, descriptors N x C float64, everything on the host.
 *
 * Parity pin: this restatement is checked (tests/test_oracle_golden.py) against
 *   (1) the reference notebook's stored known answer for toyexample_6
 *       (cost 27497.41105769231 -> 276.125, 51 entries), and
 *   (2) fixtures produced by running the unmodified reference in the build container
 *       (oracle/gen_golden.py -> tests/golden/ npz files).
 */
#ifndef FMPNP_ORACLE_H
#define FMPNP_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* loss ids: the gin-registered names of helpers/utils.py:15-38 plus barron_loss(alpha) :40-78 */
enum {
    ORACLE_LOSS_SQUARED = 0,
    ORACLE_LOSS_HUBER = 1,
    ORACLE_LOSS_CAUCHY = 2,
    ORACLE_LOSS_GEMAN_MCCLURE = 3,
    ORACLE_LOSS_BARRON = 4
};

/* forward() exit paths (model.py:316-320, 441-445, 411-413, 491-494) */
enum {
    ORACLE_OK = 0,            /* loop ran to the end: best pose returned                      */
    ORACLE_EMPTY_SUPPORT = 1, /* no point projected inside the image: CURRENT pose returned   */
    ORACLE_NAN_STEP = 2       /* NaN in delta: loop left early (intended `break`), best pose  */
};

typedef struct {
    int32_t n_iters;
    int32_t loss_id;
    double barron_alpha;    /* only for ORACLE_LOSS_BARRON */
    double lambda0;         /* sparseFeaturePnP.lambda_ */
    int32_t use_ratio_test; /* ratio_threshold is not None */
    double ratio_threshold;
} oracle_opts;

/* One problem, reference layout. `ref_ld` is the row stride of feature_ref in elements
 * (lets a channel slice [start,end) be passed without a copy, as model.py:209 does with a view). */
typedef struct {
    const double *pts3D;       /* N x 3 */
    const double *feature_ref; /* N x C (row stride ref_ld) */
    const double *F;           /* C x Hf x Wf */
    const double *Gx;          /* C x Hf x Wf (may be NULL for cost-only calls) */
    const double *Gy;
    int32_t N, C, Hf, Wf, ref_ld;
    double K[9];
    int32_t im_width, im_height;
} oracle_problem;

/* Optional history = the reference's `track_` dict (model.py:170-176). Arrays are caller
 * allocated for `cap` entries; `len` entries are written.  mask / thr_mask are stored over ALL
 * N points (thr_mask[n] = 1 iff point n is supported AND survived the ratio test); the
 * reference's `threshold_mask` (length Ns) is thr_mask[mask]. */
typedef struct {
    int32_t cap, len;
    double *Rs;        /* cap x 9  */
    double *ts;        /* cap x 3  */
    double *costs;     /* cap      */
    int32_t *points2d; /* cap x N x 2 (x, y) */
    uint8_t *mask;     /* cap x N  */
    uint8_t *thr_mask; /* cap x N  */
} oracle_track;

typedef struct {
    double R[9], t[3];
    double first_cost; /* mean rho at i == 0 (what forward() would store in initial_cost_ if unset) */
    double best_cost;  /* best_cost_ */
    int32_t best_num_inliers;
    int32_t status;
    int32_t iters_run; /* number of trial steps evaluated */
    int32_t has_best;  /* 0 if the loop body never reached the i == 0 block */
} oracle_result;

/* sparseFeaturePnP.forward (model.py:246-494) */
int oracle_forward(const oracle_problem *p, const oracle_opts *o, const double R0[9], const double t0[3],
                   oracle_result *out, oracle_track *trk);

/* sparseFeaturePnP.compute_cost (model.py:216-243); returns 0 and *cost, or 1 if no supported point (None) */
int oracle_compute_cost(const oracle_problem *p, const oracle_opts *o, const double R[9], const double t[3],
                        double *cost);

/* find_inliers (model.py:131-152), mode "ratio_max"; writes N flags */
int oracle_find_inliers(const oracle_problem *p, int32_t loss_id, double barron_alpha, double threshold,
                        const double R[9], const double t[3], uint8_t *inliers);

/* sparseFeaturePnP.multilevel_optimization (model.py:179-213) for channel-slice pyramids
 * (target_size = kernel_size = None, the only form any shipped config uses).
 * levels: n_levels x 2 ints (start, end). initial_cost: the compute_cost value (SSD mean);
 * returns 1 if compute_cost was None (init pose returned), else 0.  `lvl` (n_levels entries,
 * may be NULL) receives the per-level forward() results. */
int oracle_multilevel(const oracle_problem *p, const oracle_opts *o, const int32_t *levels, int32_t n_levels,
                      const double R0[9], const double t0[3], oracle_result *out, double *initial_cost,
                      oracle_result *lvl, oracle_track *trk);

/* Batch driver used only by bench.py (cpu_baseline / --impl reference): runs `B` independent
 * problems with `n_threads` POSIX threads (one query per thread at a time). levels may be NULL
 * (single forward). */
int oracle_forward_batch(const oracle_problem *ps, int32_t B, const oracle_opts *o, const int32_t *levels,
                         int32_t n_levels, const double *R0s, const double *t0s, oracle_result *outs,
                         int32_t n_threads);

/* building blocks exposed for stage-by-stage tests */
void oracle_project(const double *pts3D, int32_t N, const double R[9], const double t[3], const double K[9],
                    double *P /*N x 3*/, int32_t *px /*N x 2*/);
void oracle_loss(int32_t loss_id, double alpha, const double *x, int32_t n, double *rho, double *w);
int oracle_optimizer_step(const double g[6], const double H[36], double lambda_, double lr, double delta[6]);
void oracle_so3exp(const double w[3], double dr[9]);

#ifdef __cplusplus
}
#endif
#endif
