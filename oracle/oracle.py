"""ctypes front-end of the CPU oracle (`oracle/fmpnp_oracle.c`).  TEST INFRASTRUCTURE ONLY:
imported by `tests/`, by `__graft_entry__.smoke()` (as the checker) and by `bench.py`'s
cpu_baseline / `--impl reference` legs.  The product (`featuremetric-pnp_b200/`) never imports it.

Inputs follow the reference's layout: CHW float64 maps, N x C float64 descriptors (numpy).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libfmpnp_oracle.so")

LOSS_IDS = {"squared_loss": 0, "huber_loss": 1, "cauchy_loss": 2, "geman_mcclure_loss": 3, "barron_loss": 4}
OK, EMPTY_SUPPORT, NAN_STEP = 0, 1, 2


class Opts(C.Structure):
    _fields_ = [("n_iters", C.c_int32), ("loss_id", C.c_int32), ("barron_alpha", C.c_double),
                ("lambda0", C.c_double), ("use_ratio_test", C.c_int32), ("ratio_threshold", C.c_double)]


class Problem(C.Structure):
    _fields_ = [("pts3D", C.c_void_p), ("feature_ref", C.c_void_p), ("F", C.c_void_p), ("Gx", C.c_void_p),
                ("Gy", C.c_void_p), ("N", C.c_int32), ("C", C.c_int32), ("Hf", C.c_int32), ("Wf", C.c_int32),
                ("ref_ld", C.c_int32), ("K", C.c_double * 9), ("im_width", C.c_int32), ("im_height", C.c_int32)]


class Track(C.Structure):
    _fields_ = [("cap", C.c_int32), ("len", C.c_int32), ("Rs", C.c_void_p), ("ts", C.c_void_p),
                ("costs", C.c_void_p), ("points2d", C.c_void_p), ("mask", C.c_void_p), ("thr_mask", C.c_void_p)]


class Result(C.Structure):
    _fields_ = [("R", C.c_double * 9), ("t", C.c_double * 3), ("first_cost", C.c_double),
                ("best_cost", C.c_double), ("best_num_inliers", C.c_int32), ("status", C.c_int32),
                ("iters_run", C.c_int32), ("has_best", C.c_int32)]


def build(force=False):
    src = [os.path.join(_HERE, f) for f in ("fmpnp_oracle.c", "fmpnp_oracle.h", "Makefile")]
    stale = (not os.path.exists(_LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in src)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.oracle_forward.restype = C.c_int
        _lib.oracle_compute_cost.restype = C.c_int
        _lib.oracle_find_inliers.restype = C.c_int
        _lib.oracle_multilevel.restype = C.c_int
        _lib.oracle_forward_batch.restype = C.c_int
        _lib.oracle_optimizer_step.restype = C.c_int
    return _lib


def _f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class _Prob:
    """Keeps the numpy buffers alive next to the C struct."""

    def __init__(self, pts3D, feature_ref, F, Gx, Gy, K, im_width, im_height):
        self.pts = _f64(pts3D).reshape(-1, 3)
        self.ref = _f64(feature_ref)
        self.F = _f64(F)
        self.Gx = _f64(Gx) if Gx is not None else None
        self.Gy = _f64(Gy) if Gy is not None else None
        Cn, Hf, Wf = self.F.shape
        N = self.pts.shape[0]
        assert self.ref.shape == (N, Cn), (self.ref.shape, N, Cn)
        p = Problem()
        p.pts3D, p.feature_ref, p.F = _ptr(self.pts), _ptr(self.ref), _ptr(self.F)
        p.Gx, p.Gy = _ptr(self.Gx), _ptr(self.Gy)
        p.N, p.C, p.Hf, p.Wf, p.ref_ld = N, Cn, Hf, Wf, Cn
        p.K = (C.c_double * 9)(*_f64(K).reshape(9))
        p.im_width, p.im_height = int(im_width), int(im_height)
        self.c = p
        self.N = N


def make_opts(n_iters, loss="squared_loss", lambda_=0.01, ratio_threshold=None, barron_alpha=0.0):
    o = Opts()
    o.n_iters = int(n_iters)
    o.loss_id = LOSS_IDS[loss] if isinstance(loss, str) else int(loss)
    o.barron_alpha = float(barron_alpha)
    o.lambda0 = float(lambda_)
    o.use_ratio_test = 0 if ratio_threshold is None else 1
    o.ratio_threshold = 0.0 if ratio_threshold is None else float(ratio_threshold)
    return o


class _Trk:
    def __init__(self, cap, N):
        self.Rs = np.zeros((cap, 3, 3))
        self.ts = np.zeros((cap, 3))
        self.costs = np.zeros(cap)
        self.points2d = np.zeros((cap, N, 2), dtype=np.int32)
        self.mask = np.zeros((cap, N), dtype=np.uint8)
        self.thr = np.zeros((cap, N), dtype=np.uint8)
        t = Track()
        t.cap, t.len = cap, 0
        t.Rs, t.ts, t.costs = _ptr(self.Rs), _ptr(self.ts), _ptr(self.costs)
        t.points2d, t.mask, t.thr_mask = _ptr(self.points2d), _ptr(self.mask), _ptr(self.thr)
        self.c = t

    def as_dict(self, use_ratio):
        n = self.c.len
        return {"Rs": self.Rs[:n].copy(), "ts": self.ts[:n].copy(), "costs": self.costs[:n].copy(),
                "points2d": self.points2d[:n].copy(), "mask": self.mask[:n].astype(bool),
                "threshold_mask": self.thr[:n].astype(bool) if use_ratio else None}


def _res(r):
    return {"R": np.array(r.R[:]).reshape(3, 3), "t": np.array(r.t[:]), "first_cost": r.first_cost,
            "best_cost": r.best_cost, "best_num_inliers": r.best_num_inliers, "status": r.status,
            "iters_run": r.iters_run, "has_best": r.has_best}


def forward(pts3D, feature_ref, F, Gx, Gy, K, im_width, im_height, R0, t0, n_iters, loss="squared_loss",
            lambda_=0.01, ratio_threshold=None, barron_alpha=0.0, track=False):
    P = _Prob(pts3D, feature_ref, F, Gx, Gy, K, im_width, im_height)
    o = make_opts(n_iters, loss, lambda_, ratio_threshold, barron_alpha)
    R0, t0 = _f64(R0).reshape(9), _f64(t0).reshape(3)
    r = Result()
    trk = _Trk(n_iters + 1, P.N) if track else None
    lib().oracle_forward(C.byref(P.c), C.byref(o), _ptr(R0), _ptr(t0), C.byref(r),
                         C.byref(trk.c) if trk else None)
    out = _res(r)
    if trk:
        out["track"] = trk.as_dict(ratio_threshold is not None)
    return out


def compute_cost(pts3D, feature_ref, F, K, im_width, im_height, R, t, ratio_threshold=None):
    P = _Prob(pts3D, feature_ref, F, None, None, K, im_width, im_height)
    o = make_opts(1, "squared_loss", 0.0, ratio_threshold)
    R, t = _f64(R).reshape(9), _f64(t).reshape(3)
    cost = C.c_double()
    rc = lib().oracle_compute_cost(C.byref(P.c), C.byref(o), _ptr(R), _ptr(t), C.byref(cost))
    return None if rc == 1 else cost.value


def find_inliers(pts3D, R, t, F, feature_ref, K, im_width, im_height, threshold, loss="squared_loss",
                 barron_alpha=0.0):
    P = _Prob(pts3D, feature_ref, F, None, None, K, im_width, im_height)
    R, t = _f64(R).reshape(9), _f64(t).reshape(3)
    out = np.zeros(P.N, dtype=np.uint8)
    lib().oracle_find_inliers(C.byref(P.c), C.c_int32(LOSS_IDS[loss]), C.c_double(barron_alpha),
                              C.c_double(threshold), _ptr(R), _ptr(t), _ptr(out))
    return out.astype(bool)


def multilevel(levels, pts3D, feature_ref, F, Gx, Gy, K, im_width, im_height, R0, t0, n_iters,
               loss="squared_loss", lambda_=0.01, ratio_threshold=None, barron_alpha=0.0, track=False):
    P = _Prob(pts3D, feature_ref, F, Gx, Gy, K, im_width, im_height)
    o = make_opts(n_iters, loss, lambda_, ratio_threshold, barron_alpha)
    R0, t0 = _f64(R0).reshape(9), _f64(t0).reshape(3)
    lv = np.ascontiguousarray(np.asarray([(int(a), int(b)) for a, b in levels], dtype=np.int32))
    r = Result()
    lvl = (Result * len(lv))()
    ic = C.c_double()
    trk = _Trk((n_iters + 1) * len(lv), P.N) if track else None
    rc = lib().oracle_multilevel(C.byref(P.c), C.byref(o), _ptr(lv), C.c_int32(len(lv)), _ptr(R0), _ptr(t0),
                                 C.byref(r), C.byref(ic), lvl, C.byref(trk.c) if trk else None)
    out = _res(r)
    out["initial_cost"] = None if rc == 1 else ic.value
    out["levels"] = [_res(x) for x in lvl] if rc == 0 else []
    if trk:
        out["track"] = trk.as_dict(ratio_threshold is not None)
    return out


def forward_batch(problems, R0s, t0s, n_iters, loss="squared_loss", lambda_=0.01, ratio_threshold=None,
                  barron_alpha=0.0, levels=None, n_threads=1):
    """problems: list of dicts with keys pts3D, feature_ref, F, Gx, Gy, K, im_width, im_height."""
    Ps = [_Prob(p["pts3D"], p["feature_ref"], p["F"], p["Gx"], p["Gy"], p["K"], p["im_width"], p["im_height"])
          for p in problems]
    arr = (Problem * len(Ps))(*[p.c for p in Ps])
    o = make_opts(n_iters, loss, lambda_, ratio_threshold, barron_alpha)
    R0s = _f64(R0s).reshape(len(Ps), 9)
    t0s = _f64(t0s).reshape(len(Ps), 3)
    outs = (Result * len(Ps))()
    lv = None
    nl = 0
    if levels:
        lv = np.ascontiguousarray(np.asarray([(int(a), int(b)) for a, b in levels], dtype=np.int32))
        nl = len(lv)
    lib().oracle_forward_batch(arr, C.c_int32(len(Ps)), C.byref(o), _ptr(lv), C.c_int32(nl), _ptr(R0s), _ptr(t0s),
                               outs, C.c_int32(n_threads))
    return [_res(x) for x in outs]


def project(pts3D, R, t, K):
    pts = _f64(pts3D).reshape(-1, 3)
    N = pts.shape[0]
    P = np.zeros((N, 3))
    px = np.zeros((N, 2), dtype=np.int32)
    R, t, K = _f64(R).reshape(9), _f64(t).reshape(3), _f64(K).reshape(9)
    lib().oracle_project(_ptr(pts), C.c_int32(N), _ptr(R), _ptr(t), _ptr(K), _ptr(P), _ptr(px))
    return P, px


def loss(name, x, barron_alpha=0.0):
    x = _f64(x).reshape(-1)
    rho, w = np.zeros_like(x), np.zeros_like(x)
    lib().oracle_loss(C.c_int32(LOSS_IDS[name]), C.c_double(barron_alpha), _ptr(x), C.c_int32(x.size), _ptr(rho),
                      _ptr(w))
    return rho, w


def optimizer_step(g, H, lambda_, lr):
    g, H = _f64(g).reshape(6), _f64(H).reshape(36)
    d = np.zeros(6)
    lib().oracle_optimizer_step(_ptr(g), _ptr(H), C.c_double(lambda_), C.c_double(lr), _ptr(d))
    return d


def so3exp(w):
    w = _f64(w).reshape(3)
    dr = np.zeros(9)
    lib().oracle_so3exp(_ptr(w), _ptr(dr))
    return dr.reshape(3, 3)
