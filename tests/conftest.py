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
# Launch routes.  The task scheduler is the default for every batch size; an explicit cluster_size (or
# FMPNP_SCHEDULER=0) selects one CTA / thread-block cluster per query.  `route` re-runs a test with the process-wide
# default context replaced by one that fixes the tasks per evaluation (FMPNP_UNITS is read when a context is created),
# so that small cases cover several task counts as well.
# ----------------------------------------------------------------------------------------------
_route_ctx = {}


def _scheduler_context(units):
    import fmpnp_b200 as fm
    if units not in _route_ctx:
        # "light": the light-evaluation route (csrc/fmpnp_light.cuh) forced on for every batch size -- by default it is
        # taken only by batches of half an SM count of queries or more with a ratio test
        # (FMPNP_LIGHT_MISS_DEN=1: and kept on however many points an evaluation has to gather)
        env = {"FMPNP_LIGHT": "1", "FMPNP_LIGHT_MISS_DEN": "1"} if units == "light" else {"FMPNP_UNITS": str(units)}
        os.environ.update(env)
        try:
            _route_ctx[units] = fm.Context()
        finally:
            for k in env:
                del os.environ[k]
    return _route_ctx[units]


@pytest.fixture(params=["auto", "tasks2", "tasks128", "light"])
def route(request):
    if request.param == "auto":
        yield request.param
        return
    import torch
    from fmpnp_b200 import binding
    dev = torch.cuda.current_device()
    binding.default_context()
    saved = binding._default_ctx[dev]
    binding._default_ctx[dev] = _scheduler_context("light" if request.param == "light" else int(request.param[5:]))
    try:
        yield request.param
    finally:
        binding._default_ctx[dev] = saved
