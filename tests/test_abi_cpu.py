"""CPU: the C-ABI library loads without a GPU and exports every symbol include/fmpnp.h declares;
struct mirrors in the Python binding have the C sizes; the product fails loudly without CUDA."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ensure_built():
    import __graft_entry__ as g
    import fmpnp_b200 as fm
    if not os.path.exists(fm.lib_path()):
        g.build()
    return fm


def test_library_exports_every_declared_symbol():
    fm = _ensure_built()
    hdr = open(os.path.join(ROOT, "include", "fmpnp.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(fmpnp_[a-z_]+)\s*\(", hdr))
    assert {"fmpnp_create", "fmpnp_destroy", "fmpnp_refine_batch", "fmpnp_compute_cost_batch",
            "fmpnp_find_inliers_batch", "fmpnp_pack_chw_to_hwc", "fmpnp_last_error", "fmpnp_version",
            "fmpnp_launch_count", "fmpnp_struct_size", "fmpnp_last_geometry"} <= names
    L = ctypes.CDLL(fm.lib_path())
    for n in names:
        assert hasattr(L, n), "libfmpnp.so does not export " + n
    assert fm.lib().fmpnp_version() == 103


def test_struct_sizes_match_the_c_abi():
    fm = _ensure_built()
    from fmpnp_b200 import binding as b
    L = fm.lib()
    L.fmpnp_struct_size.restype = ctypes.c_int
    for which, cls in enumerate((b.CProblem, b.COptions, b.CResult, b.CHistory)):
        assert L.fmpnp_struct_size(which) == ctypes.sizeof(cls), cls.__name__


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    fm = _ensure_built()
    h = ctypes.c_void_p()
    assert fm.lib().fmpnp_create(ctypes.byref(h), 0) != 0  # FMPNP_ERR_NO_DEVICE
    with pytest.raises(fm.FmpnpError):
        fm.Context()
    import model
    m = model.sparseFeaturePnP(n_iters=1)
    with pytest.raises(fm.FmpnpError):
        m(torch.zeros(4, 3, dtype=torch.float64), torch.zeros(4, 2), torch.zeros(2, 8, 8), torch.zeros(2, 8, 8),
          torch.zeros(2, 8, 8), torch.eye(3, dtype=torch.float64), 64, 64)
    # the device-side adapter and its kernels have no CPU path either
    from collections import namedtuple
    from pose_prediction import optimize_feature_pnp as adapter
    P = namedtuple("P", "points_3d reference_inliers matrix")
    with pytest.raises(fm.FmpnpError):
        adapter.build_problem(torch.zeros(1, 2, 8, 8), torch.zeros(1, 2, 8, 8),
                              P(torch.zeros(4, 3).numpy(), torch.zeros(4, 2).numpy(), torch.eye(4).numpy()),
                              torch.eye(3), [64, 64])
    with pytest.raises(fm.FmpnpError):
        fm.sobel_pack(torch.zeros(2, 8, 8))


def test_loss_mapping():
    fm = _ensure_built()
    import functools
    import helpers.utils as hu
    assert fm.loss_id_of(hu.squared_loss) == (0, 0.0)
    assert fm.loss_id_of(hu.huber_loss) == (1, 0.0)
    assert fm.loss_id_of(hu.cauchy_loss) == (2, 0.0)
    assert fm.loss_id_of(hu.geman_mcclure_loss) == (3, 0.0)
    assert fm.loss_id_of(functools.partial(hu.barron_loss, alpha=-1.0)) == (4, -1.0)
    with pytest.raises(fm.FmpnpError):
        fm.loss_id_of(lambda x: x)


def test_header_is_plain_c_and_links_against_the_library(tmp_path):
    """include/fmpnp.h compiles as C99 (no C++ or CUDA types in the boundary) and a C program that takes the
    address of every declared entry point links against libfmpnp.so and runs (it only asks for the version and
    the struct sizes: no compute without a GPU)."""
    import shutil
    import subprocess
    fm = _ensure_built()
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    hdr = open(os.path.join(ROOT, "include", "fmpnp.h")).read()
    names = sorted(set(re.findall(r"\b(fmpnp_[a-z_]+)\s*\(", re.sub(r"/\*.*?\*/", "", hdr, flags=re.S))))
    src = tmp_path / "abi.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "fmpnp.h"\n'
                   "typedef void (*fn_t)(void);\nint main(void) {\n  fn_t fns[] = {%s};\n"
                   '  printf("%%d %%d %%d %%d %%d %%d\\n", fmpnp_version(), fmpnp_struct_size(0), fmpnp_struct_size(1),\n'
                   "         fmpnp_struct_size(2), (int)sizeof(fmpnp_problem), (int)(sizeof(fns) / sizeof(fns[0])));\n"
                   '  printf("%%d %%d %%d %%d %%d %%d\\n", (int)offsetof(fmpnp_options, gather_every_row),\n'
                   "         (int)offsetof(fmpnp_options, inlier_threshold), (int)sizeof(fmpnp_options),\n"
                   "         (int)offsetof(fmpnp_result, gathered_rows), (int)offsetof(fmpnp_result, reused_rows),\n"
                   "         (int)sizeof(fmpnp_result));\n"
                   "  return 0;\n}\n" % ", ".join("(fn_t)%s" % n for n in names))
    exe = tmp_path / "abi"
    libdir = os.path.dirname(fm.lib_path())
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                    str(src), "-o", str(exe), "-L", libdir, "-lfmpnp", "-Wl,-rpath," + libdir], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    assert int(out[0]) == 103 and int(out[1]) == int(out[4]) and int(out[5]) == len(names)
    # the ctypes mirrors place the fields added in version 101 where the C compiler does
    from fmpnp_b200 import binding as b
    import ctypes
    assert [int(v) for v in out[6:12]] == [b.COptions.gather_every_row.offset, b.COptions.inlier_threshold.offset,
                                           ctypes.sizeof(b.COptions), b.CResult.gathered_rows.offset,
                                           b.CResult.reused_rows.offset, ctypes.sizeof(b.CResult)]


def test_problem_batch_marshals_once():
    """ProblemBatch fills the fmpnp_problem array at construction; the batched calls take its array as is."""
    fm = _ensure_built()
    from fmpnp_b200 import binding as b

    class FakeProblem:  # what Problem.fill() needs, without CUDA tensors
        N, C, Hf, Wf, im_width, im_height = 5, 8, 4, 4, 64, 64

        def __init__(self, k):
            self.k = k

        def fill(self, c):
            c.N, c.C, c.Hf, c.Wf, c.im_width, c.im_height = self.N + self.k, self.C, self.Hf, self.Wf, 64, 64
            c.F = 0x1000 * (self.k + 1)

    batch = fm.ProblemBatch([FakeProblem(k) for k in range(3)])
    assert len(batch) == 3 and batch.n_max == 5 and batch[1].k == 1 and [p.k for p in batch] == [0, 1, 2]
    assert b._problem_array(batch) is batch.arr and [batch.arr[i].N for i in range(3)] == [5, 6, 7]
    assert batch.arr[2].F == 0x3000
