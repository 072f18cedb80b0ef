"""Import the UNMODIFIED reference optimizer (`/root/reference/featurePnP/model.py`) on CPU.

Test infrastructure only.  Works only where `/root/reference` exists (the build container);
the GPU box does not have it, so nothing under `tests/ -m gpu`, `smoke()` or `bench.py` may
call this at run time.  It is used (a) by `oracle/gen_golden.py` to produce the committed
fixtures under `tests/golden/`, and (b) by CPU tests that are skipped when the reference is
absent, to pin the C restatement (`oracle/fmpnp_oracle.c`) against the real thing.
"""
import importlib
import os
import sys

REFERENCE_ROOT = os.environ.get("FMPNP_REFERENCE_ROOT", "/root/reference")
_SHIM = os.path.join(os.path.dirname(os.path.abspath(__file__)), "refshim")


def available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "featurePnP", "model.py"))


def load():
    """Returns (model_module, utils_module) of the reference, imported under private names so
    they never collide with the product's drop-in `model` / `helpers.utils` modules."""
    if not available():
        raise RuntimeError("reference not present at %s" % REFERENCE_ROOT)
    saved_path = list(sys.path)
    saved_mods = {k: sys.modules.get(k) for k in ("model", "helpers", "helpers.utils", "gin", "kornia",
                                                   "kornia.filters")}
    try:
        for k in saved_mods:
            sys.modules.pop(k, None)
        sys.path.insert(0, os.path.join(REFERENCE_ROOT, "featurePnP"))
        sys.path.insert(0, _SHIM)
        # The reference's `helpers` directory has no __init__.py (a namespace package): any regular package called
        # `helpers` further down sys.path -- the product's drop-in one -- would win the import, and the reference's
        # model.py would silently run on the product's re-implementation of utils.py.  Load the reference's file by
        # path and register it under the names model.py imports.
        import importlib.util
        import types
        hdir = os.path.join(REFERENCE_ROOT, "featurePnP", "helpers")
        pkg = types.ModuleType("helpers")
        pkg.__path__ = [hdir]
        sys.modules["helpers"] = pkg
        spec = importlib.util.spec_from_file_location("helpers.utils", os.path.join(hdir, "utils.py"))
        ref_utils = importlib.util.module_from_spec(spec)
        sys.modules["helpers.utils"] = ref_utils
        spec.loader.exec_module(ref_utils)
        pkg.utils = ref_utils
        ref_model = importlib.import_module("model")
        assert os.path.realpath(ref_model.__file__).startswith(os.path.realpath(REFERENCE_ROOT)), ref_model.__file__
        assert os.path.realpath(ref_utils.__file__).startswith(os.path.realpath(REFERENCE_ROOT)), ref_utils.__file__
        assert sys.modules["helpers.utils"] is ref_utils and ref_model.squared_loss is ref_utils.squared_loss
    finally:
        sys.path[:] = saved_path
        for k, v in saved_mods.items():
            cur = sys.modules.pop(k, None)
            if k == "model":
                sys.modules["_fmpnp_ref_model"] = cur
            elif k == "helpers.utils":
                sys.modules["_fmpnp_ref_utils"] = cur
            if v is not None:
                sys.modules[k] = v
    return ref_model, ref_utils
