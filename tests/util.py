"""Shared helpers for the tests: golden-fixture loading and tolerance checks."""
import glob
import math
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# parity bars from BASELINE.json north_star
COST_RTOL = 1e-4        # per-iteration cost, relative
ROT_TOL_DEG = 1e-3      # final pose, degrees
TRANS_TOL_M = 1e-4      # final pose, metres
# the oracle is a float64 restatement of the same arithmetic: it must sit far inside those bars
ORACLE_COST_RTOL = 1e-9
ORACLE_POSE_ATOL = 1e-9


def golden_names():
    """Fixtures of the optimizer itself (oracle/gen_golden.py).  The adapter / resampled-pyramid fixtures
    (oracle/gen_golden_adapter.py) have their own schema and tests."""
    names = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))
    return [n for n in names if n.startswith(("syn_", "toy"))]


def load_golden(name):
    d = dict(np.load(os.path.join(GOLDEN_DIR, name + ".npz"), allow_pickle=False))
    d["loss"] = str(d["loss"])
    d["n_iters"] = int(d["n_iters"])
    d["lambda_"] = float(d["lambda_"])
    rt = float(d["ratio_threshold"])
    d["ratio_threshold"] = None if math.isnan(rt) else rt
    ba = float(d["barron_alpha"])
    d["barron_alpha"] = 0.0 if math.isnan(ba) else ba
    d["levels"] = [tuple(int(v) for v in r) for r in d["levels"]] if d["levels"].size else None
    d["im_width"] = int(d["im_width"])
    d["im_height"] = int(d["im_height"])
    d["best_num_inliers"] = int(d["best_num_inliers"])
    return d


def chw64(m):
    """HWC fp32 (product layout) -> CHW float64 (reference / oracle layout); values unchanged."""
    return np.ascontiguousarray(np.transpose(m, (2, 0, 1)).astype(np.float64))


def rot_angle_deg(Ra, Rb):
    M = np.asarray(Ra) @ np.asarray(Rb).T
    c = max(-1.0, min(1.0, (np.trace(M) - 1.0) / 2.0))
    # acos loses precision near 0: use the skew part for small angles
    s = 0.5 * math.sqrt((M[2, 1] - M[1, 2]) ** 2 + (M[0, 2] - M[2, 0]) ** 2 + (M[1, 0] - M[0, 1]) ** 2)
    return math.degrees(math.atan2(s, c))


def assert_costs_close(a, b, rtol, what=""):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, "%s: cost history length %s vs %s" % (what, a.shape, b.shape)
    both_nan = np.isnan(a) & np.isnan(b)
    rel = np.abs(a - b) / np.maximum(np.abs(b), 1e-300)
    rel[both_nan] = 0.0
    assert np.all(rel <= rtol), "%s: max relative cost difference %.3e at entry %d" % (what, rel.max(), int(rel.argmax()))
