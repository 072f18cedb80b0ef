"""GPU: the BASELINE.json configurations at their FULL per-query size (N = 4096 / 8192, C = 1024, 128 x 128 maps,
20 iterations), whole trajectory against the oracle -- the shape bench.py times.

    cfg2  Gauss-Newton, squared loss, ratio_threshold 0.8, find_inliers(0.8) fused into the launch
          (model.py:131-152 -> :246-494 on the kept points, optimize_feature_pnp.py:42-46)
    cfg3  Geman-McClure LM, lambda_ = 0.01                                     (model.py:246-494)
    cfg4  3-level channel pyramid [(512,1024),(128,512),(0,128)], N = 8192     (model.py:179-213)

Every history entry is compared: projected pixel indices, support masks and ratio-test masks bit-exact, cost
<= 1e-4 relative (bar; the float64 paths agree to 1e-9), final pose <= 1e-3 deg / 1e-4 m.  Two workloads per
configuration: the bench's own scenes (synth defaults: un-normalised Sobel, 0.2-1 degree perturbation -- short steps,
almost every record reused) and scenes whose points do cross cells (grad_scale 1/64, 1-3 degrees).  The oracle needs
about 9 s per 4096-point query on one host thread; the scenes of a test run on separate threads."""
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest
import torch

import oracle
from test_gpu_edge_and_scale import assert_reuse_is_invisible, compare, fm, run_oracle, same_records, synth

pytestmark = pytest.mark.gpu

CONFIGS = {
    "cfg2": dict(N=4096, loss="squared_loss", lam=0.0, ratio=0.8, inlier=0.8, levels=None, outliers=0.0),
    "cfg3": dict(N=4096, loss="geman_mcclure_loss", lam=0.01, ratio=None, inlier=None, levels=None, outliers=0.1),
    "cfg4": dict(N=8192, loss="squared_loss", lam=0.01, ratio=None, inlier=None,
                 levels=[(512, 1024), (128, 512), (0, 128)], outliers=0.1),
}
WORKLOADS = {
    "bench": dict(),                                          # exactly what bench.py builds
    "moving": dict(grad_scale=1.0 / 64, rot_deg=(1.0, 3.0)),  # points change cells: records are re-gathered
}
N_ITERS = 20
SEEDS = (0, 1)  # bench.py's queries 0 and 1 of rank 0


def oracle_for(sc, cfg):
    sel = None
    if cfg["inlier"] is not None:
        inp = sc.reference_inputs()
        inl = oracle.find_inliers(inp["pts3D"].numpy(), sc.R0.numpy(), sc.t0.numpy(), inp["feature_map_query"].numpy(),
                                  inp["feature_ref"].numpy(), sc.K.numpy(), sc.im_width, sc.im_height, cfg["inlier"],
                                  loss=cfg["loss"])
        sel = np.nonzero(inl)[0]
    return run_oracle(sc, N_ITERS, cfg["loss"], cfg["lam"], cfg["ratio"], cfg["levels"], sel=sel), sel


@pytest.mark.parametrize("workload", sorted(WORKLOADS))
@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_full_size_trajectory_matches_oracle(name, workload):
    cfg = CONFIGS[name]
    scs = [synth().make_scene(q, N=cfg["N"], C=1024, Hf=128, Wf=128, device="cuda", outlier_frac=cfg["outliers"],
                              **WORKLOADS[workload]) for q in SEEDS]
    probs = [fm().Problem.from_scene(sc) for sc in scs]
    kw = dict(loss=cfg["loss"], lambda_=cfg["lam"], ratio_threshold=cfg["ratio"], levels=cfg["levels"],
              inlier_threshold=cfg["inlier"])
    out = fm().refine_batch(probs, N_ITERS, track=True, **kw)
    with ThreadPoolExecutor(len(scs)) as ex:
        oracles = list(ex.map(lambda sc: oracle_for(sc, cfg), scs))
    for b, (o, sel) in enumerate(oracles):
        assert sel is None or 0 < len(sel) < cfg["N"]
        compare("%s/%s[%d]" % (name, workload, b), out["results"][b], out["history"], o, b=b,
                use_ratio=cfg["ratio"] is not None, sel=sel)
        if cfg["levels"] is None:
            assert int(out["results"]["best_num_inliers"][b]) == int(o["best_num_inliers"])
            assert abs(float(out["results"]["best_cost"][b]) - o["best_cost"]) <= 1e-9 * abs(o["best_cost"])
    # what bench.py launches -- no history, records reused -- returns the same records, and so does the launch that
    # gathers every row at every evaluation
    plain = fm().refine_batch(probs, N_ITERS, **kw)
    assert same_records(plain["results"], out["results"])
    every = fm().refine_batch(probs, N_ITERS, track=True, gather_every_row=True, **kw)
    if cfg["inlier"] is None:
        assert_reuse_is_invisible(out, every, len(probs))
    else:  # (the fused selection pass streams four rows when its records are reused and two otherwise: counters differ)
        assert same_records(out["results"], every["results"], skip=("gathered_rows", "reused_rows"))
        for k in ("len", "Rs", "ts", "costs", "points2d", "mask", "kept"):
            assert np.array_equal(out["history"][k], every["history"][k], equal_nan=k in ("Rs", "ts", "costs")), k
    g, r = out["results"]["gathered_rows"].astype(np.float64), out["results"]["reused_rows"].astype(np.float64)
    if workload == "bench":
        assert (r / (g + r)).min() > 0.5, "the bench workload is expected to reuse most records"
    print("%s/%s: rows gathered %s of the algorithmic count" % (name, workload, np.round(g / (g + r), 3)))
    torch.cuda.synchronize()
