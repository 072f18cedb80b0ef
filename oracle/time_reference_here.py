"""Times the UNMODIFIED reference (`/root/reference/featurePnP/model.py`, PyTorch CPU, float64, CHW) and the C
port (`oracle/fmpnp_oracle.c`) on the same full-shape synthetic query, in the build container.

Context for `cpu_baseline` / `--impl reference` in bench.py: the Python reference cannot travel to the GPU box,
so the timed CPU arm there is the C port; this script records how the two compare where both can run.
Writes tests/golden/reference_cpu_timing.json.  Test infrastructure only (needs /root/reference).

    python oracle/time_reference_here.py [N] [C] [map]
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "oracle"), os.path.join(ROOT, "featuremetric-pnp_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import oracle  # noqa: E402
import pyref  # noqa: E402
from fmpnp_b200 import synth  # noqa: E402


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    C = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    M = int(sys.argv[3]) if len(sys.argv) > 3 else 128
    n_iters = 20
    ref_model, ref_utils = pyref.load()
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    sc = synth.make_scene(0, N=N, C=C, Hf=M, Wf=M, device="cpu", outlier_frac=0.1)
    inp = sc.reference_inputs()
    out = {"shape": {"N": N, "C": C, "map": [M, M], "n_iters": n_iters, "loss": "geman_mcclure_loss", "lambda_": 0.01},
           "host": {"cores": threads, "torch": torch.__version__}}

    m = ref_model.sparseFeaturePnP(n_iters=n_iters, loss_fn=ref_utils.geman_mcclure_loss, lambda_=0.01)
    t0 = time.perf_counter()
    with torch.no_grad():
        R, t = m(inp["pts3D"], inp["feature_ref"], inp["feature_map_query"], inp["feature_grad_x"], inp["feature_grad_y"],
                 inp["K"], inp["im_width"], inp["im_height"], R_init=inp["R_init"], t_init=inp["t_init"])
    dt_ref = time.perf_counter() - t0
    out["reference_pytorch_cpu"] = {"seconds_per_query": dt_ref, "queries_per_s": 1.0 / dt_ref, "threads": threads}

    args = (inp["pts3D"].numpy(), inp["feature_ref"].numpy(), inp["feature_map_query"].numpy(),
            inp["feature_grad_x"].numpy(), inp["feature_grad_y"].numpy(), inp["K"].numpy(), sc.im_width, sc.im_height,
            inp["R_init"].numpy(), inp["t_init"].numpy(), n_iters)
    t0 = time.perf_counter()
    o = oracle.forward(*args, loss="geman_mcclure_loss", lambda_=0.01, ratio_threshold=None)
    dt_c = time.perf_counter() - t0
    out["c_port_one_thread"] = {"seconds_per_query": dt_c, "queries_per_s": 1.0 / dt_c}
    out["pose_difference"] = {"R_max_abs": float(np.abs(o["R"] - R.numpy()).max()),
                              "t_max_abs": float(np.abs(o["t"] - t.numpy()).max())}
    out["c_port_speedup_over_reference_per_thread"] = dt_ref / dt_c
    path = os.path.join(ROOT, "tests", "golden", "reference_cpu_timing.json")
    json.dump(out, open(path, "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
