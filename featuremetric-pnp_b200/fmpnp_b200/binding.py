"""ctypes binding of include/fmpnp.h.  Marshals pointers only; all compute is in libfmpnp.so."""
import ctypes as C
import math
import os
import threading

import numpy as np
import torch

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MAX_LEVELS = 8

LOSS_IDS = {"squared_loss": 0, "huber_loss": 1, "cauchy_loss": 2, "geman_mcclure_loss": 3, "barron_loss": 4}
STATUS_OK, STATUS_EMPTY_SUPPORT, STATUS_NAN_STEP, STATUS_NO_INITIAL_COST = 0, 1, 2, 3
SOLVER_CHOLESKY, SOLVER_LU = 0, 1


class FmpnpError(RuntimeError):
    pass


# ---------------------------------------------------------------------------------------------
# struct mirrors (must match include/fmpnp.h)
# ---------------------------------------------------------------------------------------------
class CProblem(C.Structure):
    _fields_ = [("F", C.c_void_p), ("Gx", C.c_void_p), ("Gy", C.c_void_p), ("ref", C.c_void_p),
                ("pts3D", C.c_void_p), ("active", C.c_void_p), ("K", C.c_double * 9), ("R0", C.c_double * 9),
                ("t0", C.c_double * 3), ("N", C.c_int32), ("C", C.c_int32), ("Hf", C.c_int32), ("Wf", C.c_int32),
                ("im_width", C.c_int32), ("im_height", C.c_int32)]


class COptions(C.Structure):
    _fields_ = [("n_iters", C.c_int32), ("loss_id", C.c_int32), ("barron_alpha", C.c_double),
                ("lambda0", C.c_double), ("ratio_threshold", C.c_double), ("n_levels", C.c_int32),
                ("levels", (C.c_int32 * 2) * MAX_LEVELS), ("solver", C.c_int32), ("cluster_size", C.c_int32),
                ("gather_every_row", C.c_int32), ("pad_", C.c_int32), ("inlier_threshold", C.c_double),
                ("keep_records", C.c_int32), ("reserved", C.c_int32)]


class CResult(C.Structure):
    _fields_ = [("R", C.c_double * 9), ("t", C.c_double * 3), ("best_cost", C.c_double),
                ("initial_cost", C.c_double), ("best_num_inliers", C.c_int32), ("status", C.c_int32),
                ("n_evaluations", C.c_int32), ("level_status", C.c_int32 * MAX_LEVELS),
                ("gathered_rows", C.c_int64), ("reused_rows", C.c_int64)]


class CHistory(C.Structure):
    _fields_ = [("cap", C.c_int32), ("n_stride", C.c_int32), ("len", C.c_void_p), ("Rs", C.c_void_p),
                ("ts", C.c_void_p), ("costs", C.c_void_p), ("points2d", C.c_void_p), ("mask", C.c_void_p),
                ("kept", C.c_void_p)]


RESULT_DTYPE = np.dtype(CResult)

_lib = None
_lib_lock = threading.Lock()


def lib_path():
    return os.environ.get("FMPNP_LIB", os.path.join(_PKG_DIR, "lib", "libfmpnp.so"))


def lib():
    """Loads libfmpnp.so.  Fails loudly: there is no other implementation to fall back to."""
    global _lib
    with _lib_lock:
        if _lib is None:
            path = lib_path()
            if not os.path.exists(path):
                raise FmpnpError("CUDA library %s is missing: build it with `python -c 'import __graft_entry__ as g; "
                                 "g.build()'` or `make -C featuremetric-pnp_b200/csrc` (no CPU fallback exists)" % path)
            L = C.CDLL(path)
            L.fmpnp_version.restype = C.c_int
            L.fmpnp_create.argtypes = [C.POINTER(C.c_void_p), C.c_int]
            L.fmpnp_create.restype = C.c_int
            L.fmpnp_destroy.argtypes = [C.c_void_p]
            L.fmpnp_destroy.restype = C.c_int
            L.fmpnp_last_error.argtypes = [C.c_void_p]
            L.fmpnp_last_error.restype = C.c_char_p
            L.fmpnp_launch_count.argtypes = [C.c_void_p]
            L.fmpnp_launch_count.restype = C.c_int64
            L.fmpnp_last_geometry.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                              C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
            L.fmpnp_last_geometry.restype = C.c_int
            L.fmpnp_refine_batch.argtypes = [C.c_void_p, C.POINTER(CProblem), C.c_int32, C.POINTER(COptions),
                                             C.c_void_p, C.POINTER(CHistory), C.c_void_p]
            L.fmpnp_refine_batch.restype = C.c_int
            L.fmpnp_compute_cost_batch.argtypes = [C.c_void_p, C.POINTER(CProblem), C.c_int32, C.POINTER(COptions),
                                                   C.c_void_p, C.c_void_p, C.c_void_p]
            L.fmpnp_compute_cost_batch.restype = C.c_int
            L.fmpnp_compute_cost_batch_noalloc.argtypes = L.fmpnp_compute_cost_batch.argtypes
            L.fmpnp_compute_cost_batch_noalloc.restype = C.c_int
            L.fmpnp_find_inliers_batch.argtypes = [C.c_void_p, C.POINTER(CProblem), C.c_int32, C.c_int32, C.c_double,
                                                   C.c_double, C.c_void_p, C.c_int32, C.c_void_p]
            L.fmpnp_find_inliers_batch.restype = C.c_int
            L.fmpnp_pack_chw_to_hwc.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32,
                                                C.c_int32, C.c_void_p]
            L.fmpnp_pack_chw_to_hwc.restype = C.c_int
            L.fmpnp_sobel_pack.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
            L.fmpnp_sobel_pack.restype = C.c_int
            L.fmpnp_gather_descriptors.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                                   C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
            L.fmpnp_gather_descriptors.restype = C.c_int
            L.fmpnp_resize_bilinear_chw.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32,
                                                    C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
            L.fmpnp_resize_bilinear_chw.restype = C.c_int
            L.fmpnp_blur_chw.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                         C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
            L.fmpnp_blur_chw.restype = C.c_int
            L.fmpnp_shared_buffer_create.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]
            L.fmpnp_shared_buffer_open.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p)]
            L.fmpnp_shared_buffer_close.argtypes = [C.c_void_p, C.c_void_p]
            L.fmpnp_shared_buffer_destroy.argtypes = [C.c_void_p, C.c_void_p]
            L.fmpnp_copy_to_host.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
            for fn in (L.fmpnp_shared_buffer_create, L.fmpnp_shared_buffer_open, L.fmpnp_shared_buffer_close,
                       L.fmpnp_shared_buffer_destroy, L.fmpnp_copy_to_host):
                fn.restype = C.c_int
            _lib = L
    return _lib


class Context:
    """One fmpnp_ctx (workspace + launch geometry) bound to a CUDA device."""

    def __init__(self, device=None):
        if not torch.cuda.is_available():
            raise FmpnpError("no CUDA device: the feature-metric PnP refinement has no CPU fallback")
        self.device = torch.cuda.current_device() if device is None else torch.device(device).index or 0
        h = C.c_void_p()
        rc = lib().fmpnp_create(C.byref(h), int(self.device))
        if rc != 0:
            raise FmpnpError("fmpnp_create failed with %d" % rc)
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().fmpnp_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, rc, what):
        if rc != 0:
            msg = lib().fmpnp_last_error(self._h)
            raise FmpnpError("%s failed (%d): %s" % (what, rc, msg.decode() if msg else "?"))

    @property
    def launches(self):
        return int(lib().fmpnp_launch_count(self._h))

    @property
    def last_geometry(self):
        """(CTAs in the grid, CTAs per query, staging-ring bytes per CTA) of the last launch."""
        g, c, r, u = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
        lib().fmpnp_last_geometry(self._h, C.byref(g), C.byref(c), C.byref(r), C.byref(u))
        return dict(grid=g.value, cluster_size=c.value, ring_bytes=r.value, tasks_per_evaluation=u.value)


_default_ctx = {}


def default_context(device=None):
    if not torch.cuda.is_available():
        raise FmpnpError("no CUDA device: the feature-metric PnP refinement has no CPU fallback")
    dev = torch.cuda.current_device() if device is None else (torch.device(device).index or 0)
    if dev not in _default_ctx:
        _default_ctx[dev] = Context(dev)
    return _default_ctx[dev]


def loss_id_of(loss):
    """Maps a gin-registered loss (function object, functools.partial of barron_loss, or name) to
    (loss_id, barron_alpha)."""
    import functools
    if isinstance(loss, int):
        return int(loss), 0.0
    if isinstance(loss, str):
        name = loss.lstrip("@")
        if name in LOSS_IDS:
            return LOSS_IDS[name], 0.0
        raise FmpnpError("unknown loss %r" % loss)
    if isinstance(loss, functools.partial):
        name = getattr(loss.func, "__name__", "")
        if name == "barron_loss":
            alpha = loss.keywords.get("alpha", loss.args[0] if loss.args else None)
            if alpha is None:
                raise FmpnpError("barron_loss partial without alpha")
            return LOSS_IDS["barron_loss"], float(alpha)
        return loss_id_of(loss.func)
    alpha = getattr(loss, "fmpnp_barron_alpha", None)
    if alpha is not None:
        return LOSS_IDS["barron_loss"], float(alpha)
    name = getattr(loss, "__name__", None)
    if name in LOSS_IDS and name != "barron_loss":
        return LOSS_IDS[name], 0.0
    raise FmpnpError("loss_fn %r is not one of %s: arbitrary Python losses cannot run inside the CUDA kernel"
                     % (loss, sorted(LOSS_IDS)))


class SharedRecords:
    """Result records of ALL ranks in the memory of one GPU (fmpnp_shared_buffer_*): the owner creates the buffer and
    ships `handle` (64 bytes) to the other ranks, who open it; `block(first, count)` is what a rank passes to
    refine_batch(out=...) so that its kernels store their records straight into the owner's memory over NVLink.
    No collective, no extra kernel; the records of a launch are complete once the WRITING rank has synchronised
    its stream (the owner reads after a barrier)."""

    class _Range:
        def __init__(self, ptr):
            self._ptr = ptr

        def data_ptr(self):
            return self._ptr

    def __init__(self, n_records, handle=None, ctx=None):
        self.ctx = ctx or default_context()
        self.n_records, self.owner = int(n_records), handle is None
        p = C.c_void_p()
        if self.owner:
            buf = C.create_string_buffer(64)
            self.ctx.check(lib().fmpnp_shared_buffer_create(self.ctx._h, self.n_records * RESULT_DTYPE.itemsize,
                                                            C.byref(p), buf), "fmpnp_shared_buffer_create")
            self.handle = bytes(buf.raw)
        else:
            self.handle = bytes(handle)
            self.ctx.check(lib().fmpnp_shared_buffer_open(self.ctx._h, self.handle, C.byref(p)),
                           "fmpnp_shared_buffer_open")
        self.ptr = p.value

    def block(self, first, count):
        if first < 0 or first + count > self.n_records:
            raise FmpnpError("record block [%d, %d) outside the %d shared records" % (first, first + count, self.n_records))
        return SharedRecords._Range(self.ptr + first * RESULT_DTYPE.itemsize)

    def read(self):
        """All records as a host numpy structured array (any rank that has the buffer mapped may read)."""
        out = np.empty(self.n_records, dtype=RESULT_DTYPE)
        self.ctx.check(lib().fmpnp_copy_to_host(self.ctx._h, out.ctypes.data_as(C.c_void_p), C.c_void_p(self.ptr),
                                                out.nbytes, _stream_ptr()), "fmpnp_copy_to_host")
        return out

    def close(self):
        if getattr(self, "ptr", None):
            fn = lib().fmpnp_shared_buffer_destroy if self.owner else lib().fmpnp_shared_buffer_close
            fn(self.ctx._h, C.c_void_p(self.ptr))
            self.ptr = None


class Problem:
    """Inputs of one query (the product's layout: HWC fp32 maps).  Tensors live on the device -- or, for the maps
    and the descriptors, in PINNED host memory (`tensor.pin_memory()`): pinned memory is mapped into the device's
    address space, the kernels then read just the rows they gather over PCIe (zero-copy) instead of the caller
    uploading whole maps first."""

    def __init__(self, F, Gx, Gy, ref, pts3D, K, R0, t0, im_width, im_height, active=None):
        self.F, self.Gx, self.Gy, self.ref, self.pts3D, self.active = F, Gx, Gy, ref, pts3D, active
        self.K = np.asarray(K.cpu() if isinstance(K, torch.Tensor) else K, dtype=np.float64).reshape(9)
        self.R0 = np.asarray(R0.cpu() if isinstance(R0, torch.Tensor) else R0, dtype=np.float64).reshape(9)
        self.t0 = np.asarray(t0.cpu() if isinstance(t0, torch.Tensor) else t0, dtype=np.float64).reshape(3)
        self.im_width, self.im_height = int(im_width), int(im_height)
        for name, t_, dt in (("F", F, torch.float32), ("Gx", Gx, torch.float32), ("Gy", Gy, torch.float32),
                             ("ref", ref, torch.float32), ("pts3D", pts3D, torch.float64),
                             ("active", active, torch.uint8)):
            if t_ is None:
                continue
            ok = isinstance(t_, torch.Tensor) and t_.dtype == dt and t_.is_contiguous() and (
                t_.is_cuda or (name in ("F", "Gx", "Gy", "ref") and t_.is_pinned()))
            if not ok:
                raise FmpnpError("%s must be a contiguous CUDA %s tensor%s" % (
                    name, dt, " (or pinned host memory)" if name in ("F", "Gx", "Gy", "ref") else ""))
        if F.dim() != 3 or ref.dim() != 2 or pts3D.dim() != 2 or pts3D.shape[1] != 3:
            raise FmpnpError("expected F (Hf, Wf, C), ref (N, C), pts3D (N, 3)")
        self.Hf, self.Wf, self.C = (int(v) for v in F.shape)
        self.N = int(pts3D.shape[0])
        if tuple(ref.shape) != (self.N, self.C):
            raise FmpnpError("ref shape %s != (N=%d, C=%d)" % (tuple(ref.shape), self.N, self.C))
        for g in (Gx, Gy):
            if g is not None and tuple(g.shape) != tuple(F.shape):
                raise FmpnpError("gradient map shape %s != F shape %s" % (tuple(g.shape), tuple(F.shape)))
        if active is not None and tuple(active.shape) != (self.N,):
            raise FmpnpError("active mask must have shape (N,)")

    @classmethod
    def from_scene(cls, sc, active=None, R0=None, t0=None):
        return cls(sc.F, sc.Gx, sc.Gy, sc.ref, sc.pts3D, sc.K, sc.R0 if R0 is None else R0,
                   sc.t0 if t0 is None else t0, sc.im_width, sc.im_height, active=active)

    def fill(self, c):
        c.F, c.ref, c.pts3D = self.F.data_ptr(), self.ref.data_ptr(), self.pts3D.data_ptr()
        c.Gx = self.Gx.data_ptr() if self.Gx is not None else None
        c.Gy = self.Gy.data_ptr() if self.Gy is not None else None
        c.active = self.active.data_ptr() if self.active is not None else None
        c.K[:] = self.K.tolist()
        c.R0[:] = self.R0.tolist()
        c.t0[:] = self.t0.tolist()
        c.N, c.C, c.Hf, c.Wf = self.N, self.C, self.Hf, self.Wf
        c.im_width, c.im_height = self.im_width, self.im_height


class ProblemBatch:
    """A list of problems with its fmpnp_problem array filled in once.  The batched calls accept it in place of the
    list; a caller that refines the same queries repeatedly (or times the calls) then pays the marshalling of the
    descriptors (about 4 us per problem in Python) once instead of per call."""

    def __init__(self, problems):
        self.problems = list(problems)
        self.arr = (CProblem * len(self.problems))()
        for i, p in enumerate(self.problems):
            p.fill(self.arr[i])
        self.n_max = max([p.N for p in self.problems] + [1])

    def __len__(self):
        return len(self.problems)

    def __iter__(self):
        return iter(self.problems)

    def __getitem__(self, i):
        return self.problems[i]


def _problem_array(problems):
    return problems.arr if isinstance(problems, ProblemBatch) else ProblemBatch(problems).arr


def _options(n_iters, loss, lambda_, ratio_threshold, levels, solver, barron_alpha=None, cluster_size=0,
             gather_every_row=False, inlier_threshold=None, keep_records=False):
    o = COptions()
    o.keep_records = 1 if keep_records else 0
    o.cluster_size = int(cluster_size)
    o.gather_every_row = 1 if gather_every_row else 0
    o.inlier_threshold = 0.0 if inlier_threshold is None else float(inlier_threshold)
    lid, alpha = loss_id_of(loss)
    if barron_alpha is not None:
        alpha = float(barron_alpha)
    o.n_iters, o.loss_id, o.barron_alpha = int(n_iters), lid, alpha
    o.lambda0 = float(lambda_)
    o.ratio_threshold = float("nan") if ratio_threshold is None else float(ratio_threshold)
    levels = list(levels) if levels else []
    if len(levels) > MAX_LEVELS:
        raise FmpnpError("at most %d pyramid levels" % MAX_LEVELS)
    o.n_levels = len(levels)
    for i, lv in enumerate(levels):
        o.levels[i][0], o.levels[i][1] = int(lv[0]), int(lv[1])
    o.solver = {"cholesky": SOLVER_CHOLESKY, "lu": SOLVER_LU}[solver] if isinstance(solver, str) else int(solver)
    return o


def _stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def refine_batch(problems, n_iters, loss="squared_loss", lambda_=0.01, ratio_threshold=None, levels=None,
                 track=False, solver="cholesky", barron_alpha=None, ctx=None, sync=True, out=None, cluster_size=0,
                 gather_every_row=False, inlier_threshold=None, keep_records=False):
    """Batched sparseFeaturePnP.forward (levels=None) / multilevel_optimization (levels=[(s, e), ...]).

    keep_records=True: the same problems (arrays, sizes, order) as the previous refine_batch call of `ctx`, possibly at
    other poses: records that call left are reused where they still apply (fmpnp_options.keep_records).

    gather_every_row=True reads the rows of every supported point at every evaluation; the default keeps the
    record of a point that stays in its feature-map cell (bit-identical results, see include/fmpnp.h).
    inlier_threshold=t starts the launch with find_inliers(threshold=t, loss_fn=loss) at the initial pose and
    refines the points it keeps (== find_inliers_batch + refine_batch with that mask as `active`, one launch).

    Returns a dict: `results` (numpy structured array, one fmpnp_result per query, on the host if
    sync else None), `results_dev` (raw device tensor), and with track=True `history` (dict of
    host numpy arrays: len, Rs, ts, costs, points2d, mask, kept)."""
    ctx = ctx or default_context()
    B = len(problems)
    dev = torch.device("cuda", ctx.device)
    o = _options(n_iters, loss, lambda_, ratio_threshold, levels, solver, barron_alpha, cluster_size, gather_every_row,
                 inlier_threshold, keep_records)
    arr = _problem_array(problems)
    res_dev = out if out is not None else torch.empty((max(B, 1), RESULT_DTYPE.itemsize), dtype=torch.uint8, device=dev)
    hist, hbuf = None, None
    if track:
        cap = max(1, o.n_levels) * (int(n_iters) + 1)
        n_stride = max([p.N for p in problems] + [1])
        hbuf = dict(len=torch.zeros(B, dtype=torch.int32, device=dev),
                    Rs=torch.zeros((B, cap, 3, 3), dtype=torch.float64, device=dev),
                    ts=torch.zeros((B, cap, 3), dtype=torch.float64, device=dev),
                    costs=torch.zeros((B, cap), dtype=torch.float64, device=dev),
                    points2d=torch.zeros((B, cap, n_stride, 2), dtype=torch.int32, device=dev),
                    mask=torch.zeros((B, cap, n_stride), dtype=torch.uint8, device=dev),
                    kept=torch.zeros((B, cap, n_stride), dtype=torch.uint8, device=dev))
        hist = CHistory()
        hist.cap, hist.n_stride = cap, n_stride
        for k in ("len", "Rs", "ts", "costs", "points2d", "mask", "kept"):
            setattr(hist, k, hbuf[k].data_ptr())
    rc = lib().fmpnp_refine_batch(ctx._h, arr, B, C.byref(o), C.c_void_p(res_dev.data_ptr()),
                                  C.byref(hist) if hist is not None else None, _stream_ptr())
    ctx.check(rc, "fmpnp_refine_batch")
    ret = {"results_dev": res_dev, "results": None}
    if sync:
        if not isinstance(res_dev, torch.Tensor):
            raise FmpnpError("sync=True needs a torch tensor as `out` (a raw device range is read by its owner)")
        ret["results"] = res_dev[:B].cpu().numpy().view(RESULT_DTYPE).reshape(B)
        if B and int(ret["results"]["status"].min()) < 0:
            raise FmpnpError("fmpnp_refine_batch: the device scheduler gave up (FMPNP_STATUS_INCOMPLETE)")
        if track:
            ret["history"] = {k: v.cpu().numpy() for k, v in hbuf.items()}
    elif track:
        ret["history_dev"] = hbuf
    return ret


def compute_cost_batch(problems, ratio_threshold=None, ctx=None):
    """Batched sparseFeaturePnP.compute_cost at (R0, t0); returns (costs [B] with NaN == None, n_supported [B])."""
    ctx = ctx or default_context()
    B = len(problems)
    dev = torch.device("cuda", ctx.device)
    o = _options(1, "squared_loss", 0.0, ratio_threshold, None, "cholesky")
    arr = _problem_array(problems)
    costs = torch.empty(max(B, 1), dtype=torch.float64, device=dev)
    nsup = torch.empty(max(B, 1), dtype=torch.int32, device=dev)
    rc = lib().fmpnp_compute_cost_batch(ctx._h, arr, B, C.byref(o), C.c_void_p(costs.data_ptr()),
                                        C.c_void_p(nsup.data_ptr()), _stream_ptr())
    ctx.check(rc, "fmpnp_compute_cost_batch")
    return costs[:B], nsup[:B]


NOALLOC_MAX_QUERIES, NOALLOC_MAX_POINTS = 8, 4096


def compute_cost_batch_noalloc(problems, costs, n_supported=None, ratio_threshold=None, ctx=None):
    """fmpnp_compute_cost_batch_noalloc: compute_cost_batch into the caller's `costs` (float64 [>= B]) and
    `n_supported` (int32 [>= B] or None) CUDA tensors, without any device allocation -- the scratch is a static
    array inside the library (at most NOALLOC_MAX_QUERIES problems of NOALLOC_MAX_POINTS points per call; all calls
    on a device share it, so calls that overlap on the GPU race)."""
    ctx = ctx or default_context()
    B = len(problems)
    if not (costs.is_cuda and costs.dtype == torch.float64 and costs.is_contiguous() and costs.numel() >= B):
        raise FmpnpError("costs must be a contiguous CUDA float64 tensor with at least B elements")
    if n_supported is not None and not (n_supported.is_cuda and n_supported.dtype == torch.int32
                                        and n_supported.is_contiguous() and n_supported.numel() >= B):
        raise FmpnpError("n_supported must be a contiguous CUDA int32 tensor with at least B elements")
    o = _options(1, "squared_loss", 0.0, ratio_threshold, None, "cholesky")
    rc = lib().fmpnp_compute_cost_batch_noalloc(ctx._h, _problem_array(problems), B, C.byref(o),
                                                C.c_void_p(costs.data_ptr()),
                                                C.c_void_p(n_supported.data_ptr()) if n_supported is not None else None,
                                                _stream_ptr())
    ctx.check(rc, "fmpnp_compute_cost_batch_noalloc")
    return costs[:B], (n_supported[:B] if n_supported is not None else None)


def find_inliers_batch(problems, threshold, loss="squared_loss", barron_alpha=None, ctx=None, out=None):
    """Batched find_inliers at (R0, t0); returns a uint8 device tensor [B][max N] (`out` if given)."""
    ctx = ctx or default_context()
    B = len(problems)
    dev = torch.device("cuda", ctx.device)
    lid, alpha = loss_id_of(loss)
    if barron_alpha is not None:
        alpha = float(barron_alpha)
    if threshold is None or (isinstance(threshold, float) and math.isnan(threshold)):
        raise FmpnpError("find_inliers needs a threshold")
    n_stride = max([p.N for p in problems] + [1])
    arr = _problem_array(problems)
    if out is None:
        out = torch.zeros((max(B, 1), n_stride), dtype=torch.uint8, device=dev)
    else:
        if not (out.is_cuda and out.dtype == torch.uint8 and out.is_contiguous() and out.dim() == 2
                and out.shape[0] >= B and out.shape[1] >= n_stride):
            raise FmpnpError("find_inliers out= must be a contiguous CUDA uint8 [>=B][>=max N] tensor")
        n_stride = int(out.shape[1])
    rc = lib().fmpnp_find_inliers_batch(ctx._h, arr, B, lid, alpha, float(threshold), C.c_void_p(out.data_ptr()),
                                        n_stride, _stream_ptr())
    ctx.check(rc, "fmpnp_find_inliers_batch")
    return out[:B]


def pack_chw_to_hwc(src, ctx=None):
    """CHW float32/float64 CUDA tensor (the reference's layout) -> HWC float32 (the kernel's layout)."""
    ctx = ctx or default_context()
    if not (isinstance(src, torch.Tensor) and src.is_cuda and src.dim() == 3 and src.is_contiguous()
            and src.dtype in (torch.float32, torch.float64)):
        raise FmpnpError("pack_chw_to_hwc expects a contiguous CUDA float32/float64 (C, Hf, Wf) tensor")
    Cn, Hf, Wf = (int(v) for v in src.shape)
    dst = torch.empty((Hf, Wf, Cn), dtype=torch.float32, device=src.device)
    rc = lib().fmpnp_pack_chw_to_hwc(ctx._h, C.c_void_p(src.data_ptr()), 1 if src.dtype == torch.float64 else 0,
                                     C.c_void_p(dst.data_ptr()), Cn, Hf, Wf, _stream_ptr())
    ctx.check(rc, "fmpnp_pack_chw_to_hwc")
    return dst


def _check_chw(src, what):
    if not (isinstance(src, torch.Tensor) and src.is_cuda and src.dim() == 3 and src.is_contiguous()
            and src.dtype in (torch.float32, torch.float64)):
        raise FmpnpError("%s expects a contiguous CUDA float32/float64 (C, H, W) tensor" % what)
    return (int(v) for v in src.shape)


def sobel_pack(src, border="replicate", normalized=False, ctx=None):
    """CHW map -> (F, Gx, Gy) HWC float32: the reference's sobel_filter fused with the layout pack."""
    ctx = ctx or default_context()
    Cn, Hf, Wf = _check_chw(src, "sobel_pack")
    outs = [torch.empty((Hf, Wf, Cn), dtype=torch.float32, device=src.device) for _ in range(3)]
    rc = lib().fmpnp_sobel_pack(ctx._h, C.c_void_p(src.data_ptr()), 1 if src.dtype == torch.float64 else 0,
                                C.c_void_p(outs[0].data_ptr()), C.c_void_p(outs[1].data_ptr()),
                                C.c_void_p(outs[2].data_ptr()), Cn, Hf, Wf, {"replicate": 0, "zero": 1}[border],
                                1 if normalized else 0, _stream_ptr())
    ctx.check(rc, "fmpnp_sobel_pack")
    return tuple(outs)


def gather_descriptors(src, rows, cols, ctx=None):
    """feature_ref[n] = src[:, rows[n], cols[n]] as an (N, C) float32 CUDA tensor."""
    ctx = ctx or default_context()
    Cn, H, W = _check_chw(src, "gather_descriptors")
    rows = torch.as_tensor(rows).to(device=src.device, dtype=torch.int32).contiguous()
    cols = torch.as_tensor(cols).to(device=src.device, dtype=torch.int32).contiguous()
    N = int(rows.numel())
    if int(cols.numel()) != N:
        raise FmpnpError("rows / cols length mismatch")
    if N and (int(rows.min()) < -H or int(rows.max()) >= H or int(cols.min()) < -W or int(cols.max()) >= W):
        raise IndexError("descriptor cell outside the %d x %d reference map" % (H, W))
    dst = torch.empty((N, Cn), dtype=torch.float32, device=src.device)
    rc = lib().fmpnp_gather_descriptors(ctx._h, C.c_void_p(src.data_ptr()), 1 if src.dtype == torch.float64 else 0,
                                        C.c_void_p(rows.data_ptr()), C.c_void_p(cols.data_ptr()),
                                        C.c_void_p(dst.data_ptr()), Cn, H, W, N, _stream_ptr())
    ctx.check(rc, "fmpnp_gather_descriptors")
    return dst


def resize_bilinear_chw(src, Ho, Wo, ctx=None):
    """nn.functional.interpolate(src[None], size=(Ho, Wo), mode='bilinear')[0] in float64."""
    ctx = ctx or default_context()
    Cn, H, W = _check_chw(src, "resize_bilinear_chw")
    dst = torch.empty((Cn, int(Ho), int(Wo)), dtype=torch.float64, device=src.device)
    rc = lib().fmpnp_resize_bilinear_chw(ctx._h, C.c_void_p(src.data_ptr()), 1 if src.dtype == torch.float64 else 0,
                                         C.c_void_p(dst.data_ptr()), Cn, H, W, int(Ho), int(Wo), _stream_ptr())
    ctx.check(rc, "fmpnp_resize_bilinear_chw")
    return dst


def blur_chw(src, kernel, pad=1, ctx=None):
    """Depthwise conv2d of a CHW map with one k x k kernel and zero padding `pad`, in float64."""
    ctx = ctx or default_context()
    Cn, H, W = _check_chw(src, "blur_chw")
    kern = torch.as_tensor(kernel).to(device=src.device, dtype=torch.float64).contiguous()
    k = int(kern.shape[0])
    if kern.dim() != 2 or int(kern.shape[1]) != k:
        raise FmpnpError("blur kernel must be k x k")
    Ho, Wo = H + 2 * pad - k + 1, W + 2 * pad - k + 1
    if Ho < 1 or Wo < 1:
        raise FmpnpError("blur kernel larger than the padded map")
    dst = torch.empty((Cn, Ho, Wo), dtype=torch.float64, device=src.device)
    rc = lib().fmpnp_blur_chw(ctx._h, C.c_void_p(src.data_ptr()), 1 if src.dtype == torch.float64 else 0,
                              C.c_void_p(kern.data_ptr()), C.c_void_p(dst.data_ptr()), Cn, H, W, k, int(pad),
                              _stream_ptr())
    ctx.check(rc, "fmpnp_blur_chw")
    return dst
