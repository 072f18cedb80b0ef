#!/usr/bin/env python
"""Three rounds of feature_pnp_multi's loop (optimize_feature_pnp.py:20-47: find_inliers + refinement at the pose the
previous round returned) on a batch of 256 RobotCar-shaped queries, one launch per round, with and without
fmpnp_options.keep_records: kernel ms (CUDA events) and bytes gathered per round.  Prints JSON lines."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "featuremetric-pnp_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import fmpnp_b200 as fm  # noqa: E402
from fmpnp_b200 import synth  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
gs = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
scs = [synth.make_scene(q, N=4096, C=1024, device="cuda", grad_scale=gs) for q in range(B)]
first = [fm.Problem.from_scene(sc) for sc in scs]
kw = dict(loss="squared_loss", lambda_=0.0, ratio_threshold=0.8, inlier_threshold=0.8)


def rounds(keep):
    ctx = fm.Context()
    probs, out = fm.ProblemBatch(first), []
    for k in range(3):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        res = fm.refine_batch(probs, 20, ctx=ctx, keep_records=keep and k > 0, **kw)["results"]
        b.record()
        torch.cuda.synchronize()
        out.append((a.elapsed_time(b), float(res["gathered_rows"].sum()) * 4 / 1e9, float(res["reused_rows"].sum()) * 4 / 1e9, res))
        probs = fm.ProblemBatch([fm.Problem(p.F, p.Gx, p.Gy, p.ref, p.pts3D, p.K, np.array(r["R"]), np.array(r["t"]),
                                            p.im_width, p.im_height) for p, r in zip(first, res)])  # descriptors marshalled outside the timed part
    ctx.close()
    return out


rounds(False)  # warm-up
plain, kept = rounds(False), rounds(True)
for k in range(3):
    same = all(plain[k][3][f].tobytes() == kept[k][3][f].tobytes() for f in plain[k][3].dtype.names
               if f not in ("gathered_rows", "reused_rows"))
    print(json.dumps({"round": k + 1, "queries": B, "grad_scale": gs,
                      "ms": round(plain[k][0], 3), "gathered_gb": round(plain[k][1], 2),
                      "keep_records_ms": round(kept[k][0], 3), "keep_records_gathered_gb": round(kept[k][1], 2),
                      "results_bit_identical": bool(same)}))
