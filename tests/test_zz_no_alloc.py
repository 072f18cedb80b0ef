"""GPU: fmpnp_compute_cost_batch_noalloc -- compute_cost with its scratch in a file-scope device array of the
library (no cudaMalloc / cudaFree in the call; include/fmpnp.h) -- against fmpnp_compute_cost_batch, through the
C ABI, on two calls in a row that share the array.  Collected last on purpose."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
TIGHT = 1e-9  # the bar of the other compute_cost tests; the two entry points are expected to agree bit for bit
KS = ((100.0, 0.0, 127.055733), (0.0, 100.0, 124.546844), (0.0, 0.0, 1.0))
SMALL = dict(Hf=32, Wf=32, im_width=256, im_height=256, K=KS)


def _problems(seeds, N, Cn, far=()):
    import fmpnp_b200 as fm
    from fmpnp_b200 import synth
    probs = []
    for s in seeds:
        sc = synth.make_scene(s, N=N + 7 * (s % 3), C=Cn, device="cpu", outlier_frac=0.1, grad_scale=1.0 / 64, **SMALL)
        t0 = sc.t0 + torch.tensor([1000.0, 0.0, 0.0], dtype=torch.float64) if s in far else sc.t0  # nothing supported
        probs.append(fm.Problem(sc.F.cuda(), None, None, sc.ref.cuda(), sc.pts3D.cuda(), sc.K, sc.R0, t0, sc.im_width,
                                sc.im_height))
    return probs


def _check(probs, ratio, costs, nsup):
    import fmpnp_b200 as fm
    B = len(probs)
    want_c, want_n = fm.compute_cost_batch(probs, ratio_threshold=ratio)
    costs.fill_(-7.0)
    nsup.fill_(-7)
    mem = torch.cuda.memory_allocated()
    got_c, got_n = fm.compute_cost_batch_noalloc(probs, costs, nsup, ratio_threshold=ratio)
    assert torch.cuda.memory_allocated() == mem
    torch.cuda.synchronize()
    want_c, got_c = want_c.cpu().numpy(), got_c.cpu().numpy()
    assert np.array_equal(got_n.cpu().numpy(), want_n.cpu().numpy())
    assert np.array_equal(np.isnan(got_c), np.isnan(want_c))
    ok = ~np.isnan(want_c)
    assert np.all(np.abs(got_c[ok] - want_c[ok]) <= TIGHT * np.abs(want_c[ok])), (got_c, want_c)
    assert np.array_equal(got_c[ok], want_c[ok]), "the two entry points reduce in the same fixed order"
    assert (costs[B:] == -7.0).all() and (nsup[B:] == -7).all()  # nothing written past B


def test_noalloc_compute_cost_matches_the_allocating_entry_point_twice_in_a_row():
    import fmpnp_b200 as fm
    costs = torch.empty(fm.binding.NOALLOC_MAX_QUERIES + 2, dtype=torch.float64, device="cuda")
    nsup = torch.empty(fm.binding.NOALLOC_MAX_QUERIES + 2, dtype=torch.int32, device="cuda")
    # first call: a full batch of the largest kind the array is sized for in these tests, raw-SSD ratio test on
    _check(_problems(range(8), 900, 128), 0.7, costs, nsup)
    # second call straight after it: fewer, smaller problems with other channel counts (the array still holds the
    # first call's records, flags and a spent counter), one of them with no supported point (cost NaN == None)
    _check(_problems([11, 12, 13], 300, 64, far=(12,)), None, costs, nsup)
    # and the first shape again on a side stream, ordered by the stream itself
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        _check(_problems([3, 4], 500, 256), 0.9, costs, nsup)
    torch.cuda.current_stream().wait_stream(side)


def test_noalloc_rejects_what_does_not_fit():
    import fmpnp_b200 as fm
    from fmpnp_b200 import binding
    costs = torch.empty(16, dtype=torch.float64, device="cuda")
    with pytest.raises(fm.FmpnpError, match="at most %d" % binding.NOALLOC_MAX_QUERIES):
        fm.compute_cost_batch_noalloc(_problems(range(9), 40, 16), costs)
    with pytest.raises(fm.FmpnpError, match="at most %d" % binding.NOALLOC_MAX_POINTS):
        fm.compute_cost_batch_noalloc(_problems([0], binding.NOALLOC_MAX_POINTS + 1, 4), costs)
    # the limits themselves are accepted (N = max points, C = 4: a small map, a full-size point list)
    big = _problems([0], binding.NOALLOC_MAX_POINTS, 4)
    assert big[0].N == binding.NOALLOC_MAX_POINTS
    got, _ = fm.compute_cost_batch_noalloc(big, costs)
    want, _ = fm.compute_cost_batch(big)
    assert np.array_equal(got.cpu().numpy(), want.cpu().numpy())
    # the header's limits are the binding's
    import os
    hdr = open(os.path.join(os.path.dirname(binding._PKG_DIR), "include", "fmpnp.h")).read()
    assert "#define FMPNP_NOALLOC_MAX_QUERIES %d" % binding.NOALLOC_MAX_QUERIES in hdr
    assert "#define FMPNP_NOALLOC_MAX_POINTS %d" % binding.NOALLOC_MAX_POINTS in hdr
