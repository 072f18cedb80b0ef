"""CPU: the N > 1 path (query sharding + final result gather) on world_size-2 gloo.  The compute
behind each rank's block is the oracle here (no GPU in this suite); what is under test is the
partition and the gather."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from util import chw64, load_golden


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_host_near_gpu_is_a_no_op_without_nvml_or_gpu():
    from fmpnp_b200.sharding import host_near_gpu
    before = os.sched_getaffinity(0)
    with host_near_gpu(0) as h:
        assert h.cores is None or h.cores <= len(before)
    assert os.sched_getaffinity(0) == before


def test_shard_range_covers_everything():
    from fmpnp_b200.sharding import shard_range
    for B in (0, 1, 5, 16, 17, 256, 16384):
        for G in (1, 2, 3, 4, 8):
            seen = []
            for r in range(G):
                lo, hi = shard_range(B, r, G)
                assert 0 <= lo <= hi <= B
                seen += list(range(lo, hi))
            assert seen == list(range(B)), (B, G)
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _worker(rank, world, port, names, q):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    root = os.path.dirname(here)
    for p in (here, os.path.join(root, "oracle"), os.path.join(root, "featuremetric-pnp_b200")):
        sys.path.insert(0, p)
    import oracle
    from fmpnp_b200.binding import RESULT_DTYPE
    from fmpnp_b200.sharding import gather_result_records, records_to_numpy, shard_range
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    B = len(names)
    lo, hi = shard_range(B, rank, world)
    rec = np.zeros(hi - lo, dtype=RESULT_DTYPE)
    for i, name in enumerate(names[lo:hi]):
        g = load_golden(name)
        o = oracle.forward(g["pts3D"], g["ref"].astype(np.float64), chw64(g["F"]), chw64(g["Gx"]), chw64(g["Gy"]), g["K"],
                           g["im_width"], g["im_height"], g["R0"], g["t0"], g["n_iters"], loss=g["loss"],
                           lambda_=g["lambda_"], ratio_threshold=g["ratio_threshold"], barron_alpha=g["barron_alpha"])
        rec[i]["R"] = o["R"].reshape(9)
        rec[i]["t"] = o["t"]
        rec[i]["best_cost"] = o["best_cost"]
        rec[i]["best_num_inliers"] = o["best_num_inliers"]
        rec[i]["status"] = o["status"]
    local = torch.from_numpy(rec.view(np.uint8).reshape(hi - lo, RESULT_DTYPE.itemsize).copy())
    allrec = records_to_numpy(gather_result_records(local, B), RESULT_DTYPE)
    # the asynchronous form (what bench.py uses so that ranks do not run in lock-step) returns the same records
    again, work = gather_result_records(local, B, async_op=True)
    if work is not None:
        work.wait()
    assert np.array_equal(records_to_numpy(again, RESULT_DTYPE).view(np.uint8), allrec.view(np.uint8))
    if rank == 0:
        q.put((allrec["best_cost"].tolist(), allrec["best_num_inliers"].tolist(), allrec["R"].tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_gather_matches_single_process():
    names = ["syn_squared_lm", "syn_geman_lm", "syn_cauchy_lm", "syn_huber_lm", "syn_gn_ratio08"]  # 5 over 2 ranks: ragged
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, names, q)) for r in range(2)]
    for p in procs:
        p.start()
    costs, inl, Rs = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for i, name in enumerate(names):
        g = load_golden(name)
        assert abs(costs[i] - g["best_cost"]) <= 1e-9 * abs(g["best_cost"]), name
        assert inl[i] == g["best_num_inliers"], name
        np.testing.assert_allclose(np.array(Rs[i]).reshape(3, 3), g["out_R"], atol=1e-9)
