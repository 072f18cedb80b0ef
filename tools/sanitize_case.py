"""Small end-to-end cases for compute-sanitizer (memcheck / racecheck / synccheck): every streaming mode of the
refinement kernel (short rows, whole rows, chunked rows, unaligned generic path), clusters of 1 / 4 / 16 CTAs,
find_inliers, compute_cost and the adapter kernels.  Usage: compute-sanitizer --tool racecheck python tools/sanitize_case.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "featuremetric-pnp_b200"))
import torch  # noqa: E402

import fmpnp_b200 as fm  # noqa: E402
from fmpnp_b200 import synth  # noqa: E402

KS = ((50.0, 0.0, 63.5), (0.0, 50.0, 62.3), (0.0, 0.0, 1.0))
dev = torch.device("cuda:0")
for C, levels in ((1024, [(512, 1024), (128, 512), (0, 128)]), (20, [(8, 20), (0, 8)]), (5, None), (640, None)):
    sc = synth.make_scene(3, N=300, C=C, Hf=16, Wf=16, im_width=128, im_height=128, K=KS, device="cpu", outlier_frac=0.1,
                          grad_scale=1.0 / 64)
    p = fm.Problem(sc.F.to(dev), sc.Gx.to(dev), sc.Gy.to(dev), sc.ref.to(dev), sc.pts3D.to(dev), sc.K, sc.R0, sc.t0, 128, 128)
    for S in (1, 4, 16):
        r = fm.refine_batch([p] * (3 if S == 1 else 1), 3, loss="geman_mcclure_loss", lambda_=0.01, ratio_threshold=0.9,
                            levels=levels, track=True, cluster_size=S)["results"]
        assert int(r["status"][0]) == 0
    # task scheduler (default route), and the light-evaluation route forced on (FMPNP_LIGHT is read when a context is made)
    r = fm.refine_batch([p] * 5, 3, loss="squared_loss", lambda_=0.0, ratio_threshold=0.8, inlier_threshold=0.8 if levels is None else None,
                        levels=levels)["results"]
    assert int(r["status"][0]) >= 0
    os.environ["FMPNP_LIGHT"] = "1"
    os.environ["FMPNP_LIGHT_MISS_DEN"] = "1"
    lctx = fm.Context()
    del os.environ["FMPNP_LIGHT"], os.environ["FMPNP_LIGHT_MISS_DEN"]
    r2 = fm.refine_batch([p] * 5, 3, loss="squared_loss", lambda_=0.0, ratio_threshold=0.8, inlier_threshold=0.8 if levels is None else None,
                         levels=levels, ctx=lctx)["results"]
    assert r2["R"].tobytes() == r["R"].tobytes()
    lctx.close()
    fm.find_inliers_batch([p, p], 0.8, loss="cauchy_loss")
    fm.compute_cost_batch([p], ratio_threshold=0.7)
x = torch.rand((37, 19, 23), device=dev)
fm.sobel_pack(x)
fm.sobel_pack(x.double(), border="zero", normalized=True)
fm.pack_chw_to_hwc(x)
fm.gather_descriptors(x, torch.tensor([0, 5, -1]), torch.tensor([3, 22, -2]))
fm.blur_chw(fm.resize_bilinear_chw(x, 11, 30), torch.full((3, 3), 1.0 / 9))
torch.cuda.synchronize()
print("sanitize_case ok")
