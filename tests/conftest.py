import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "featuremetric-pnp_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_cuda = torch.cuda.is_available()
    except Exception:
        has_cuda = False
    if has_cuda:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


# ----------------------------------------------------------------------------------------------
# Launch routes.  A single query is refined by a thread-block cluster by default; the task scheduler takes
# over from 5 queries on.  `route` re-runs a test with the process-wide default context replaced by one that
# forces the scheduler (FMPNP_UNITS is read when a context is created), so that small cases cover both.
# ----------------------------------------------------------------------------------------------
_route_ctx = {}


def _scheduler_context(units):
    import fmpnp_b200 as fm
    if units not in _route_ctx:
        os.environ["FMPNP_UNITS"] = str(units)
        try:
            _route_ctx[units] = fm.Context()
        finally:
            del os.environ["FMPNP_UNITS"]
    return _route_ctx[units]


@pytest.fixture(params=["auto", "tasks2", "tasks128"])
def route(request):
    if request.param == "auto":
        yield request.param
        return
    import torch
    from fmpnp_b200 import binding
    dev = torch.cuda.current_device()
    binding.default_context()
    saved = binding._default_ctx[dev]
    binding._default_ctx[dev] = _scheduler_context(int(request.param[5:]))
    try:
        yield request.param
    finally:
        binding._default_ctx[dev] = saved
