"""GPU: seeded random problem shapes against the oracle, through both launch routes.

Every case draws N, C, map size, loss, damping, ratio threshold and (sometimes) a channel pyramid, builds a
ragged batch of six such queries (B > 4: the task scheduler, with whatever number of tasks per evaluation the
library picks), and checks each query against the oracle: pixel indices and masks bit-exact, costs and poses to
the tight float64 bar.  The first query of the batch is then refined alone by a thread-block cluster of 8 CTAs and
must give the same record bit for bit."""
import numpy as np
import pytest
import torch

from test_gpu_edge_and_scale import assert_reuse_is_invisible, compare, fm, run_oracle, same_records, synth, to_problem

pytestmark = pytest.mark.gpu
LOSSES = ["squared_loss", "huber_loss", "cauchy_loss", "geman_mcclure_loss"]
K = ((50.0, 0.0, 63.5), (0.0, 50.0, 62.3), (0.0, 0.0, 1.0))


def draw_case(seed):
    r = np.random.RandomState(1000 + seed)
    C = int(r.choice([4, 8, 20, 36, 64, 100, 128, 132, 256, 300, 384, 512, 640, 1024, 7, 1]))
    n_hi = 600 if C >= 512 else 2500
    cfg = dict(C=C, loss=LOSSES[int(r.randint(4))], lambda_=float(r.choice([0.0, 0.01, 1.0])),
               ratio=None if r.rand() < 0.5 else float(r.choice([0.5, 0.8, 0.95])), n_iters=int(r.randint(2, 7)),
               Ns=[int(x) for x in r.randint(1, n_hi, size=6)], map=int(r.choice([8, 16, 24])), levels=None)
    if cfg["lambda_"] == 0.0:  # undamped Gauss-Newton needs a well-posed system: fewer than ~6 points make H
        cfg["Ns"] = [max(n, 12) for n in cfg["Ns"]]  # singular and the step rounding noise (no parity to speak of)
    if C >= 8 and r.rand() < 0.4:  # coarse-to-fine channel slices (not necessarily multiples of 4)
        a, b = sorted(int(x) for x in r.choice(np.arange(1, C), size=2, replace=False))
        cfg["levels"] = [(b, C), (a, b), (0, a)]
    return cfg


@pytest.mark.parametrize("seed", range(64))
def test_random_ragged_batches(seed):
    cfg = draw_case(seed)
    scs = [synth().make_scene(7000 + 10 * seed + i, N=n, C=cfg["C"], Hf=cfg["map"], Wf=cfg["map"], im_width=128,
                              im_height=128, K=K, grad_scale=1.0 / 64, outlier_frac=0.1) for i, n in enumerate(cfg["Ns"])]
    probs = [to_problem(sc) for sc in scs]
    kw = dict(loss=cfg["loss"], lambda_=cfg["lambda_"], ratio_threshold=cfg["ratio"], levels=cfg["levels"], track=True)
    out = fm().refine_batch(probs, cfg["n_iters"], **kw)
    g = fm().default_context().last_geometry
    assert g["tasks_per_evaluation"] >= 1 and g["cluster_size"] == 1, g  # six queries: the task scheduler
    for i, sc in enumerate(scs):
        o = run_oracle(sc, cfg["n_iters"], cfg["loss"], cfg["lambda_"], cfg["ratio"], cfg["levels"])
        compare("seed %d query %d %r" % (seed, i, cfg), out["results"][i], out["history"], o, b=i,
                use_ratio=cfg["ratio"] is not None)
    assert_reuse_is_invisible(out, fm().refine_batch(probs, cfg["n_iters"], gather_every_row=True, **kw), len(probs))
    one = fm().refine_batch(probs[:1], cfg["n_iters"], cluster_size=8, **kw)
    assert fm().default_context().last_geometry["tasks_per_evaluation"] == 0
    assert same_records(one["results"][0], out["results"][0]), "cluster route differs from scheduler route: %r" % (cfg,)
    n = int(one["history"]["len"][0])
    assert np.array_equal(one["history"]["costs"][0, :n], out["history"]["costs"][0, :n], equal_nan=True)
    # find_inliers and compute_cost on the same ragged batch
    if cfg["levels"] is None:
        thr = cfg["ratio"] or 0.8
        inl = fm().find_inliers_batch(probs, thr, loss=cfg["loss"]).cpu().numpy()
        costs, nsup = fm().compute_cost_batch(probs, ratio_threshold=cfg["ratio"])
        import oracle
        for i, sc in enumerate(scs):
            inp = sc.reference_inputs()
            a = (inp["pts3D"].numpy(), inp["R_init"].numpy(), inp["t_init"].numpy(), inp["feature_map_query"].numpy(),
                 inp["feature_ref"].numpy(), inp["K"].numpy(), sc.im_width, sc.im_height)
            exp = oracle.find_inliers(*a, thr, loss=cfg["loss"])
            assert np.array_equal(inl[i, :cfg["Ns"][i]].astype(bool), exp), "find_inliers %d %r" % (i, cfg)
            c = oracle.compute_cost(inp["pts3D"].numpy(), inp["feature_ref"].numpy(), inp["feature_map_query"].numpy(),
                                    inp["K"].numpy(), sc.im_width, sc.im_height, inp["R_init"].numpy(),
                                    inp["t_init"].numpy(), ratio_threshold=cfg["ratio"])
            if c is None:
                assert int(nsup[i]) == 0 and torch.isnan(costs[i])
            else:
                got = float(costs[i])
                assert (np.isnan(c) and np.isnan(got)) or abs(got - c) <= 1e-9 * abs(c), "compute_cost %d %r" % (i, cfg)
