#!/usr/bin/env python
"""bench.py -- refined queries / second of the feature-metric PnP hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[1]): per GPU, 256 synthetic
RobotCar-shaped queries x 4096 points, C = 1024 hypercolumn channels, 128 x 128 maps
(F, Gx, Gy HWC fp32 = 201 MB per query, 51.5 GB per GPU: far larger than L2), 1024 x 1024
image, Gauss-Newton (lambda_ = 0) with squared loss, ratio_threshold = 0.8 and an initial inlier
mask from find_inliers(threshold = 0.8), 20 LM iterations.

One step = one pass of the hot path over the batch = ONE launch of fmpnp::refine_tasks_kernel: the
find_inliers selection runs as the first pass of the refinement launch (fmpnp_options.inlier_threshold;
--two-calls times the two separate launches).  With N > 1 every rank's kernel stores its result
records straight into rank 0's memory over NVLink peer access (the path's only exchange: the final
result gather; --gather nccl times an NCCL all-gather per step instead).  `value` has the inputs
resident in HBM; `e2e` goes through the same C ABI with HOST (pinned) buffers, host<->device traffic
inside the timed region.  Weak scaling: every rank owns its own 256 queries and their maps.

`roofline.achieved` / `frac` count the bytes the timed launch really gathered (the kernels' own row
counters) -- a physical bandwidth; `frac_contract` is SURVEY 8(d)'s algorithmic-byte figure, which
exceeds 1 because records of points that stay in their cell are reused; `every_row_*` is the same
launch with fmpnp_options.gather_every_row = 1 (the HBM-bound regime), timed in the same run.

Other workloads (not the driver's line; kept under profiles/):
    --config geman_lm | pyramid          cfg3 / cfg4 shapes
    --grad-scale S --rot-deg LO HI       scenes whose points cross cells (the default scenes take short steps)
    --total-queries T                    a JOB of T queries split over the ranks (shard_range, strong scaling) and
                                         streamed through HBM in chunks of --queries; --config cmu_slice = cfg5
                                         (cfg3 shape, T = 16384)

`--impl reference` times the reference algorithm's CPU port (oracle/, float64, CHW maps, the
reference's own operation order) on the host cores -- the Python reference itself cannot travel
to the GPU box (it lives in /root/reference of the build container).
"""
import argparse
import json
import math
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "featuremetric-pnp_b200"))

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "refined queries/sec (20 LM iters, 4k pts, C=1024)"
UNIT = "queries/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    # workload overrides (exploration only; the defaults are the contract)
    ap.add_argument("--queries", type=int, default=256, help="queries per GPU (resident batch / chunk size)")
    ap.add_argument("--points", type=int, default=None, help="points per query (default 4096; 8192 for the pyramid)")
    ap.add_argument("--channels", type=int, default=1024)
    ap.add_argument("--map", type=int, default=128)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--config", default="gn_inlier", choices=["gn_inlier", "geman_lm", "pyramid", "cmu_slice"])
    ap.add_argument("--grad-scale", type=float, default=1.0,
                    help="scale of the Sobel maps (1 = un-normalised 3x3 Sobel, the reference's; 0.125 = kornia's "
                         "normalised kernel; 1/64 = gradients per pixel for 8-px cells: long steps)")
    ap.add_argument("--rot-deg", type=float, nargs=2, default=[0.2, 1.0], metavar=("LO", "HI"),
                    help="initial rotation error range in degrees (SURVEY 8(d): 0.2 .. 1)")
    ap.add_argument("--total-queries", type=int, default=0,
                    help="job mode: this many queries in total, split over the ranks and streamed through HBM in "
                         "chunks of --queries (strong scaling); 0 = the resident batch of --queries per GPU")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="N > 1: result records stored by the kernel into rank 0's memory (peer), or an NCCL "
                         "all-gather per step (nccl)")
    ap.add_argument("--sustain-s", type=float, default=2.0, help="length of the sustained diagnostic loop (0 = off)")
    ap.add_argument("--e2e-queries", type=int, default=0, help="queries per e2e step (0 = the whole batch)")
    ap.add_argument("--host-numa", default="local", choices=["local", "interleave"],
                    help="pinned host buffers of the e2e legs: on the GPU's NUMA node, or interleaved over all nodes")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-adapter-e2e", action="store_true")
    ap.add_argument("--no-zero-copy", action="store_true", help="skip the zero-copy e2e variants")
    ap.add_argument("--no-every-row", action="store_true", help="skip the gather_every_row arm")
    ap.add_argument("--cpu-queries", type=int, default=0, help="cpu_baseline sample size (0 = auto)")
    ap.add_argument("--two-calls", action="store_true",
                    help="cfg2: fmpnp_find_inliers_batch + fmpnp_refine_batch(active = mask) instead of one launch with "
                         "fmpnp_options.inlier_threshold")
    ap.add_argument("--gather-every-row", action="store_true",
                    help="time the refinement with fmpnp_options.gather_every_row = 1 (no record reuse) as the main arm")
    ap.add_argument("--outliers", type=float, default=None, help="outlier fraction (default: 0 for cfg2, 0.1 otherwise)")
    a = ap.parse_args()
    if a.points is None:
        a.points = 8192 if a.config == "pyramid" else 4096
    if a.config == "cmu_slice" and a.total_queries == 0:
        a.total_queries = 16384
    return a


def workload(args):
    shape = "%d points, C=%d, %dx%d maps" % (args.points, args.channels, args.map, args.map)
    if args.total_queries:
        what = "%d queries in total, sharded over the GPUs and streamed through HBM in chunks of %d" % (
            args.total_queries, args.queries)
    else:
        what = "%d queries per GPU" % args.queries
    c = dict(workload="cfg2: synthetic RobotCar-shaped batch, %s x %s, GN + inlier mask" % (what, shape),
             queries_per_gpu=args.queries, points=args.points, channels=args.channels, map=[args.map, args.map],
             image=[1024, 1024], n_iters=args.iters, storage="HWC fp32", arithmetic="f64",
             l2="inputs (%.1f GB per GPU) far exceed the 126 MB L2; no flush needed" % (
                 args.queries * (3 * args.map * args.map + args.points) * args.channels * 4 / 1e9))
    if args.config == "gn_inlier":
        c.update(loss="squared_loss", lambda_=0.0, ratio_threshold=0.8, inlier_mask="find_inliers(threshold=0.8)",
                 levels=None)
    elif args.config in ("geman_lm", "cmu_slice"):
        tag = "cfg3" if args.config == "geman_lm" else "cfg5"
        c.update(workload=c["workload"].replace("cfg2", tag).replace("GN + inlier mask", "Geman-McClure LM"),
                 loss="geman_mcclure_loss", lambda_=0.01, ratio_threshold=None, inlier_mask=None, levels=None)
    else:
        C = args.channels
        c.update(workload=c["workload"].replace("cfg2", "cfg4").replace("GN + inlier mask", "3-level pyramid"),
                 loss="squared_loss", lambda_=0.01, ratio_threshold=None, inlier_mask=None,
                 levels=[[C // 2, C], [C // 8, C // 2], [0, C // 8]])
    if args.grad_scale != 1.0 or list(args.rot_deg) != [0.2, 1.0]:
        c.update(grad_scale=args.grad_scale, rot_deg=list(args.rot_deg))
        c["workload"] += " (grad_scale %g, initial rotation %g-%g deg)" % (args.grad_scale, args.rot_deg[0], args.rot_deg[1])
    return c


def scene_kwargs(args):
    outl = args.outliers if args.outliers is not None else (0.0 if args.config == "gn_inlier" else 0.1)
    return dict(N=args.points, C=args.channels, Hf=args.map, Wf=args.map, outlier_frac=outl, grad_scale=args.grad_scale,
                rot_deg=tuple(args.rot_deg))


# ---------------------------------------------------------------------------------------------
# clocks sampler (NVML), runs during the timed region
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, device_index):
        self.samples, self.power, self.reasons, self.max_mhz = [], [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            uuid = torch.cuda.get_device_properties(device_index).uuid
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.nv = pynvml
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                self.power.append(self.nv.nvmlDeviceGetPowerUsage(self.h) / 1e3)
                fn = getattr(self.nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                    self.nv.nvmlDeviceGetCurrentClocksThrottleReasons
                bits = int(fn(self.h))
                for b, n in self.REASONS.items():
                    if bits & b and n != "gpu_idle":
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop.wait(0.004)  # a step takes a few milliseconds

    def __enter__(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples), "power_w": float(np.median(self.power)) if self.power else None}


# ---------------------------------------------------------------------------------------------
# CPU arm: the oracle port (reference algorithm, float64, CHW) on host threads
# ---------------------------------------------------------------------------------------------
def scene_to_reference_problem(sc):
    inp = sc.reference_inputs()
    return dict(pts3D=inp["pts3D"].numpy(), feature_ref=inp["feature_ref"].numpy(), F=inp["feature_map_query"].numpy(),
                Gx=inp["feature_grad_x"].numpy(), Gy=inp["feature_grad_y"].numpy(), K=inp["K"].numpy(),
                im_width=sc.im_width, im_height=sc.im_height), inp["R_init"].numpy(), inp["t_init"].numpy()


def host_threads():
    """Host threads the CPU arm uses: every core this process may run on, capped at 64 (each thread holds one
    full-shape query in the reference's layout: ~0.45 GB of float64 CHW maps)."""
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    return max(1, min(cores, 64))


def cpu_problems(args, cfg, n_queries, scenes=None):
    """`n_queries` full-shape queries in the reference's layout (float64, CHW), inlier mask applied like cfg2 does.
    Inputs are synthesised on the GPU when there is one (input generation only), else on the host."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    from fmpnp_b200 import synth
    gen_dev = "cuda" if torch.cuda.is_available() else "cpu"
    probs, R0s, t0s = [], [], []
    for q in range(n_queries):
        if scenes is not None:
            sc = scenes[q]
        else:
            sc = synth.make_scene(q, device=gen_dev, **scene_kwargs(args))
        p, R0, t0 = scene_to_reference_problem(sc)
        if cfg["inlier_mask"]:
            inl = oracle.find_inliers(p["pts3D"], R0, t0, p["F"], p["feature_ref"], p["K"], p["im_width"], p["im_height"],
                                      0.8, loss=cfg["loss"])
            p["pts3D"] = np.ascontiguousarray(p["pts3D"][inl])
            p["feature_ref"] = np.ascontiguousarray(p["feature_ref"][inl])
        probs.append(p)
        R0s.append(R0)
        t0s.append(t0)
        del sc
    return probs, np.stack(R0s), np.stack(t0s)


def cpu_pass(cfg, problems, threads):
    """One timed pass of the CPU implementation over `problems`; returns (seconds, its results)."""
    import oracle
    probs, R0s, t0s = problems
    t = time.perf_counter()
    out = oracle.forward_batch(probs, R0s, t0s, cfg["n_iters"], loss=cfg["loss"], lambda_=cfg["lambda_"],
                               ratio_threshold=cfg["ratio_threshold"], levels=cfg["levels"], n_threads=threads)
    return time.perf_counter() - t, out


def parity_vs_cpu(cpu_results, gpu_records):
    """The CPU arm's final poses / costs against the GPU launch's records of the SAME queries (north-star bars:
    pose within 1e-3 deg and 1e-4 m, cost within 1e-4 relative, inlier counts equal)."""
    rot, tr, rc, same_n = 0.0, 0.0, 0.0, True
    for o, r in zip(cpu_results, gpu_records):
        Rg, tg = np.array(r["R"]).reshape(3, 3), np.array(r["t"])
        M = Rg @ np.asarray(o["R"]).T
        s = 0.5 * math.sqrt((M[2, 1] - M[1, 2]) ** 2 + (M[0, 2] - M[2, 0]) ** 2 + (M[1, 0] - M[0, 1]) ** 2)
        rot = max(rot, math.degrees(math.atan2(s, max(-1.0, min(1.0, (np.trace(M) - 1.0) / 2.0)))))
        tr = max(tr, float(np.abs(tg - np.asarray(o["t"])).max()))
        if o["has_best"]:
            rc = max(rc, abs(float(r["best_cost"]) - o["best_cost"]) / max(abs(o["best_cost"]), 1e-300))
            same_n = same_n and int(r["best_num_inliers"]) == int(o["best_num_inliers"])
    n = min(len(cpu_results), len(gpu_records))
    return {"queries": n, "max_rot_deg": rot, "max_trans_m": tr, "max_rel_best_cost": rc,
            "best_num_inliers_equal": bool(same_n),
            "within_north_star": bool(rot <= 1e-3 and tr <= 1e-4 and rc <= 1e-4 and same_n)}


def run_reference(args, cfg, rank):
    if rank != 0:
        return
    threads = host_threads()
    per_step = threads  # one full-shape query per thread per step
    problems = cpu_problems(args, cfg, per_step)
    if args.warmup > 0:  # one warm-up pass is enough on the CPU
        cpu_pass(cfg, (problems[0][:2], problems[1][:2], problems[2][:2]), min(threads, 2))
    tot_q, tot_t = 0, 0.0
    for _ in range(max(1, args.steps)):
        tot_t += cpu_pass(cfg, problems, threads)[0]
        tot_q += per_step
    v = tot_q / tot_t
    sample = "%d full-shape queries per step (seeds 0..%d of the same workload), one per thread, %d step(s)" % (
        per_step, per_step - 1, max(1, args.steps))
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(1, args.steps), "higher_is_better": True,
            "scaling": "strong" if args.total_queries else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": cfg,
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "oracle/ C port of featurePnP/model.py (float64, CHW, reference operation order) on host threads; "
                    "the Python reference cannot travel to the GPU box"}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# helpers of our arm
# ---------------------------------------------------------------------------------------------
def max_over_ranks(x, world, dev):
    if world == 1:
        return float(x)
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def interleave_host_memory(on):
    """set_mempolicy(MPOL_INTERLEAVE over every node) for the allocations that follow (pinned buffers), or back to
    the default policy.  Returns False when the syscall is unavailable."""
    import ctypes
    try:
        libc = ctypes.CDLL(None, use_errno=True)
        nodes = [int(d[4:]) for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()]
        if not on or len(nodes) < 2:
            return libc.syscall(238, 0, None, 0) == 0 and not on  # MPOL_DEFAULT
        mask = ctypes.c_ulong(sum(1 << n for n in nodes))
        return libc.syscall(238, 3, ctypes.byref(mask), ctypes.c_ulong(max(nodes) + 2)) == 0  # MPOL_INTERLEAVE
    except Exception:
        return False


def timed_launches(fn, steps, events=None):
    """Runs fn(i) `steps` times between two events on the current stream; returns elapsed ms."""
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        fn(i)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b)


# ---------------------------------------------------------------------------------------------
# job mode: T queries in total, sharded over the ranks, streamed through HBM in chunks
# ---------------------------------------------------------------------------------------------
def run_job(args, cfg, rank, local_rank, world, dev, fm, sharding, synth):
    """BASELINE configs 3-5 at their stated sizes: more queries than fit in HBM at once.  Each rank owns the
    contiguous block shard_range gives it (the reference shards a slice the same way: run.py:38-39,
    submit_cmu_auto.sh:17-43) and refines it chunk by chunk: a chunk's inputs are produced in HBM (here: synthesised
    -- in the pipeline the CNN leaves the hypercolumns there), one launch refines it, its records land in the job's
    result array (rank 0's memory when N > 1).  The clock covers the refinement launches (CUDA events, summed, max
    over ranks): inputs are resident when a chunk's launch starts, as in the resident mode; wall time with input
    generation is reported next to it."""
    import torch.distributed as dist
    from fmpnp_b200.binding import RESULT_DTYPE
    ctx = fm.default_context()
    T = args.total_queries
    lo, hi = sharding.shard_range(T, rank, world)
    chunk = args.queries
    kw = scene_kwargs(args)
    levels = cfg["levels"]
    fused = bool(cfg["inlier_mask"])
    shared = sharding.shared_result_records(T, ctx=ctx) if world > 1 else None
    res_local = torch.empty((max(hi - lo, 1), RESULT_DTYPE.itemsize), dtype=torch.uint8, device=dev)
    # the chunk's map sets stay resident (one per slot, distinct per rank); every query of the job has its own points,
    # descriptors and initial pose (seeded by its global index) and views the map set of its slot
    t0 = time.perf_counter()
    maps = []
    for s in range(min(chunk, hi - lo)):
        g = torch.Generator(device=dev)
        g.manual_seed(99991 + rank * chunk + s)
        maps.append(synth.make_map(args.channels, args.map, args.map, g, dev, args.grad_scale))
    torch.cuda.synchronize()
    gen_maps_s = time.perf_counter() - t0

    def refine(probs, out):
        fm.refine_batch(probs, cfg["n_iters"], loss=cfg["loss"], lambda_=cfg["lambda_"],
                        ratio_threshold=cfg["ratio_threshold"], levels=levels, ctx=ctx, sync=False, out=out,
                        gather_every_row=args.gather_every_row, inlier_threshold=0.8 if fused else None)

    kern_ms, n_chunks, gen_s = [], 0, 0.0
    wall0 = time.perf_counter()
    events = []
    with ClockSampler(local_rank) as clk:
        for c0 in range(lo, hi, chunk):
            c1 = min(hi, c0 + chunk)
            tg = time.perf_counter()
            scenes = [synth.make_scene(q, device=dev, shared_map=maps[q - c0], **kw) for q in range(c0, c1)]
            probs = fm.ProblemBatch([fm.Problem.from_scene(sc) for sc in scenes])
            torch.cuda.synchronize()
            gen_s += time.perf_counter() - tg
            if n_chunks == 0:  # warm-up launch on the first chunk (untimed)
                refine(probs, res_local[: c1 - c0])
                torch.cuda.synchronize()
            out = shared.block(c0, c1 - c0) if shared is not None else res_local[c0 - lo: c1 - lo]
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            refine(probs, out)
            b.record()
            events.append((a, b))
            torch.cuda.synchronize()  # the chunk's inputs are released below
            n_chunks += 1
            del scenes, probs
        torch.cuda.synchronize()
    wall_s = time.perf_counter() - wall0
    kern_ms = [a.elapsed_time(b) for a, b in events]
    if world > 1:
        dist.barrier()
    busy_ms = float(sum(kern_ms))
    job_ms = max_over_ranks(busy_ms, world, dev)
    wall_max = max_over_ranks(wall_s, world, dev)
    value = T / (job_ms / 1e3)
    # resident rate on this GPU for comparison: the last chunk's shape, repeated
    line = None
    if rank == 0:
        recs = shared.read() if shared is not None else res_local[: hi - lo].cpu().numpy().view(RESULT_DTYPE).reshape(-1)
        ok = int((recs["status"] >= 0).sum())
        gath = float(recs["gathered_rows"].astype(np.float64).sum()) * 4.0
        reus = float(recs["reused_rows"].astype(np.float64).sum()) * 4.0
        peak = measured_peak()[0]
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": n_chunks, "warmup": 1,
                "ms_per_step": job_ms / max(n_chunks, 1), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": dict(cfg, total_queries=T, chunk_queries=chunk,
                               maps="each rank keeps %d map sets resident; every query has its own points, descriptors "
                                    "and initial pose and views the map set of its chunk slot" % len(maps)),
                "job": {"queries": T, "queries_finished": ok, "chunks_per_rank": n_chunks, "refine_ms_max_over_ranks": job_ms,
                        "wall_s_with_input_generation": wall_max, "input_generation_s_rank0": gen_s + gen_maps_s,
                        "rank0_kernel_ms_per_chunk": [round(x, 3) for x in kern_ms[:64]]},
                "roofline": {"bound": "hbm", "kernel": "fmpnp::refine_tasks_kernel", "unit": "GB/s", "peak": peak,
                             "achieved": gath / world / (job_ms / 1e3) / 1e9, "frac": gath / world / (job_ms / 1e3) / 1e9 / peak,
                             "frac_contract": (gath + reus) / world / (job_ms / 1e3) / 1e9 / peak,
                             "reused_fraction": reus / max(gath + reus, 1.0), "traffic": None,
                             "note": "per-GPU average over the job (bytes of all ranks / N / time)"},
                "clocks": clk.summary(), "gpu_launches": n_chunks,
                "status_histogram": {str(k): int(v) for k, v in zip(*np.unique(recs["status"], return_counts=True))}}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        if shared is not None:
            shared.close()
        dist.destroy_process_group()


def measured_peak():
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        return float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def main():
    args = parse()
    cfg = workload(args)
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, cfg, rank)
        return

    import torch.distributed as dist
    import fmpnp_b200 as fm
    from fmpnp_b200 import sharding, synth
    from fmpnp_b200.binding import RESULT_DTYPE

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    if args.total_queries:
        run_job(args, cfg, rank, local_rank, world, dev, fm, sharding, synth)
        return
    ctx = fm.default_context()

    B = args.queries
    levels = cfg["levels"]
    kw = scene_kwargs(args)
    outl = kw["outlier_frac"]
    # ---- synthetic inputs, generated in HBM (seeded per global query index)
    t_gen = time.perf_counter()
    scenes = [synth.make_scene(rank * B + q, device=dev, **kw) for q in range(B)]
    active = torch.ones((B, args.points), dtype=torch.uint8, device=dev) if cfg["inlier_mask"] else None
    # descriptors marshalled once (fmpnp_problem array in host memory); every call still stages them to the device
    probs = fm.ProblemBatch([fm.Problem.from_scene(sc, active=active[i] if active is not None else None)
                             for i, sc in enumerate(scenes)])
    probs_noact = fm.ProblemBatch([fm.Problem.from_scene(sc) for sc in scenes])
    torch.cuda.synchronize()
    gen_s = time.perf_counter() - t_gen
    res_dev = torch.empty((B, RESULT_DTYPE.itemsize), dtype=torch.uint8, device=dev)

    fused = bool(cfg["inlier_mask"]) and not args.two_calls
    # ---- N > 1: where the records of the timed steps go
    shared, gather_mode = None, "none"
    if world > 1:
        gather_mode = args.gather
        if gather_mode == "peer":
            try:
                shared = sharding.shared_result_records(world * B, ctx=ctx)
            except Exception as e:  # CUDA IPC unavailable on this box: the collective route
                shared, gather_mode = None, "nccl (peer buffer unavailable: %r)" % (e,)
            ok = torch.tensor([1 if shared is not None else 0], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok.item()) == 0 and shared is not None:
                shared.close()
                shared, gather_mode = None, "nccl (peer buffer unavailable on another rank)"
    out_timed = shared.block(rank * B, B) if shared is not None else res_dev

    def launch(out, every_row=args.gather_every_row):
        """find_inliers -> refinement of the points it keeps: one launch (inlier_threshold), or with --two-calls the
        find_inliers kernel (active mask written in place) followed by the refinement kernel."""
        if cfg["inlier_mask"] and not fused:
            fm.find_inliers_batch(probs_noact, 0.8, loss=cfg["loss"], ctx=ctx, out=active)
        fm.refine_batch(probs_noact if fused else probs, cfg["n_iters"], loss=cfg["loss"], lambda_=cfg["lambda_"],
                        ratio_threshold=cfg["ratio_threshold"], levels=levels, ctx=ctx, sync=False, out=out,
                        gather_every_row=every_row, inlier_threshold=0.8 if fused else None)

    pending = []

    def step(i=None, every_row=args.gather_every_row, kev=None, out=None):
        out = out_timed if out is None else out
        if i is not None and kev is not None:
            kev[i][0].record()
        launch(out, every_row)
        if i is not None and kev is not None:
            kev[i][1].record()
        if world > 1 and shared is None:  # NCCL route: one all-gather of the records per step, on NCCL's own stream
            pending.append(sharding.gather_result_records(res_dev, world * B, async_op=True)[1])

    def drain():
        t = time.perf_counter()
        for w in pending:
            if w is not None:
                w.wait()
        del pending[:]
        return time.perf_counter() - t

    def timed_region(steps, every_row=args.gather_every_row):
        """EXACTLY `steps` steps between a barrier + synchronize on both sides; returns (max-over-ranks ms, this rank's
        ms, per-launch ms, seconds spent waiting for the gathers)."""
        kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(steps):
            step(i, every_row, kev)
        wait_s = drain()  # NCCL route: every step's gather has completed before the clock stops
        b.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        mine = a.elapsed_time(b)
        return max_over_ranks(mine, world, dev), mine, [x.elapsed_time(y) for x, y in kev], wait_s

    # one untimed launch into local memory: the row counters of a launch (identical for every step: nothing carries
    # over between launches), then the warm-up steps
    launch(res_dev)
    for _ in range(max(args.warmup, 0)):
        step()
    drain()
    torch.cuda.synchronize()
    res = res_dev.cpu().numpy().view(RESULT_DTYPE).reshape(B).copy()
    launches0 = ctx.launches
    with ClockSampler(local_rank) as clk:
        total_ms, my_ms, kern_ms, wait_s = timed_region(args.steps)
    launches = ctx.launches - launches0
    geo = ctx.last_geometry  # of the timed refinement launches
    value = world * B * args.steps / (total_ms / 1e3)

    # ---- the gathered records (N > 1): complete, and this rank's block equals its local records
    gather_check = None
    if world > 1:
        if shared is not None:
            allrec = shared.read() if rank == 0 else None
        else:
            allrec = sharding.records_to_numpy(sharding.gather_result_records(res_dev, world * B), RESULT_DTYPE)
        if rank == 0:
            mine = allrec[:B]
            gather_check = {"records": int(len(allrec)), "all_finished": bool((allrec["status"] >= 0).all()),
                            "rank0_block_equals_local": bool(all(
                                np.ascontiguousarray(mine[f]).tobytes() == np.ascontiguousarray(res[f]).tobytes()
                                for f in res.dtype.names))}

    # ---- roofline of the dominant kernel (refine_tasks_kernel), from the kernel's own row counters
    gath_bytes = float(res["gathered_rows"].astype(np.float64).sum()) * 4.0    # rows the launch actually read
    reused_bytes = float(res["reused_rows"].astype(np.float64).sum()) * 4.0    # rows it did not have to read again
    alg_bytes = gath_bytes + reused_bytes  # every supported point's 4 rows at every evaluation
    pass_bytes, first_eval_bytes = 0.0, None
    _, nsup_all = fm.compute_cost_batch(probs_noact, ctx=ctx)  # supported points at the initial pose
    nsup_sum = float(nsup_all.double().sum().item())
    if fused:  # the inlier pass is not one of the contract's n_iters + 1 evaluations: leave its rows out of the contract figure
        rows = 4 if (levels is None and not args.gather_every_row) else 2
        pass_bytes = nsup_sum * args.channels * rows * 4.0
        alg_bytes -= pass_bytes
    if levels is None:  # what one full evaluation of the batch must read: every supported point's four rows, once
        first_eval_bytes = nsup_sum * args.channels * 4 * 4.0
    peak, peak_src = measured_peak()
    k_ms = float(np.mean(kern_ms))
    achieved = gath_bytes / (k_ms / 1e3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    # the ncu capture was taken on the default workload: it says nothing about another one
    captured = (args.config == "gn_inlier" and args.queries == 256 and args.points == 4096 and args.channels == 1024 and
                args.map == 128 and cfg["n_iters"] == 20 and args.grad_scale == 1.0 and list(args.rot_deg) == [0.2, 1.0] and
                not args.two_calls)
    if os.path.exists(tpath) and captured:
        try:
            traffic = json.load(open(tpath)).get("refine_kernel_dram_bytes_per_launch_every_row" if args.gather_every_row
                                                 else "refine_kernel_dram_bytes_per_launch")
        except Exception:
            traffic = None
    kname = "fmpnp::refine_tasks_kernel" if geo["tasks_per_evaluation"] else "fmpnp::refine_kernel"
    roofline = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src, "kernel_ms": k_ms,
                "gathered_bytes_per_launch": gath_bytes,
                "reused_fraction": reused_bytes / max(gath_bytes + reused_bytes, 1.0),
                "frac_contract": alg_bytes / (k_ms / 1e3) / 1e9 / peak,
                "algorithmic_bytes_per_launch": alg_bytes, "inlier_pass_bytes_per_launch": pass_bytes,
                "gather_floor_ms": first_eval_bytes / peak / 1e6 if first_eval_bytes else None,
                "frac_of_gather_floor": first_eval_bytes / peak / 1e6 / k_ms if first_eval_bytes else None,
                "share_of_step": k_ms * args.steps / my_ms, "geometry": geo,
                "kernel_ms_all": [round(x, 3) for x in kern_ms],
                "note": "achieved / frac = bytes the timed launch read (fmpnp_result.gathered_rows x 4) / its CUDA-event "
                        "time / the measured COPY peak (a read-only stream can exceed it slightly).  The reference's "
                        "lookup is nearest-cell (model.py:74-97): a point that stays in its cell keeps its record "
                        "(bit-identical), so most of SURVEY 8(d)'s algorithmic bytes are never read -- frac_contract "
                        "counts them anyway and exceeds 1.  gather_floor_ms = one evaluation's rows at the peak: the "
                        "least this launch can take; every_row_* = the same launch with gather_every_row = 1, the "
                        "HBM-bound regime.  Nothing carries over between launches: every timed step gathers its first "
                        "evaluation in full"}

    # ---- the same launch with every row gathered at every evaluation (no record reuse): the HBM-bound regime
    if not args.gather_every_row and not args.no_every_row:
        for _ in range(2):
            step(every_row=True, out=res_dev)
        drain()
        t2, _, k2s, _ = timed_region(args.steps, every_row=True)
        launch(res_dev, every_row=True)
        torch.cuda.synchronize()
        res2 = res_dev.cpu().numpy().view(RESULT_DTYPE).reshape(B)
        k2 = float(np.mean(k2s))
        b2 = float(res2["gathered_rows"].astype(np.float64).sum()) * 4.0
        same = all(np.ascontiguousarray(res[f]).tobytes() == np.ascontiguousarray(res2[f]).tobytes()
                   for f in res.dtype.names if f not in ("gathered_rows", "reused_rows"))
        roofline.update({"every_row_value": world * B * args.steps / (t2 / 1e3), "every_row_kernel_ms": k2,
                         "every_row_bytes_per_launch": b2, "every_row_achieved": b2 / (k2 / 1e3) / 1e9,
                         "every_row_frac": b2 / (k2 / 1e3) / 1e9 / peak, "every_row_bit_identical": bool(same)})
        launch(res_dev)  # (res_dev back to the default arm's records for the comparisons below)
        torch.cuda.synchronize()

    # ---- sustained diagnostic: the same step back to back for a couple of seconds (clocks / power settle)
    sustained = None
    if args.sustain_s > 0:
        n_sus = int(min(5000, max(args.steps, math.ceil(args.sustain_s * 1e3 / max(total_ms / args.steps, 1e-3)))))
        with ClockSampler(local_rank) as clk_s:
            ts, _, ks, _ = timed_region(n_sus)
        sustained = {"steps": n_sus, "value": world * B * n_sus / (ts / 1e3), "ms_per_step": ts / n_sus,
                     "kernel_ms": float(np.mean(ks)), "clocks": clk_s.summary()}

    # ---- per-rank view (N > 1): which GPU the maximum comes from, and what the gather cost
    per_rank = None
    if world > 1:
        mine = {"rank": rank, "kernel_ms": k_ms, "kernel_ms_max": float(np.max(kern_ms)), "region_ms": my_ms,
                "gather_wait_s": wait_s, "clocks": clk.summary(),
                "sustained_kernel_ms": sustained["kernel_ms"] if sustained else None}
        allr = [None] * world
        dist.all_gather_object(allr, mine)
        ks = [r["kernel_ms"] for r in allr]
        per_rank = {"kernel_ms": [round(x, 4) for x in ks], "region_ms": [round(r["region_ms"], 3) for r in allr],
                    "sm_mhz": [r["clocks"].get("sm_mhz") for r in allr], "power_w": [r["clocks"].get("power_w") for r in allr],
                    "reasons": [r["clocks"].get("reasons") for r in allr],
                    "gather_wait_s": [round(r["gather_wait_s"], 6) for r in allr],
                    "sustained_kernel_ms": [r["sustained_kernel_ms"] for r in allr],
                    "kernel_ms_min_median_max": [float(np.min(ks)), float(np.median(ks)), float(np.max(ks))],
                    "slowest_rank_over_fastest": float(np.max(ks) / np.min(ks)),
                    "step_over_slowest_kernel": total_ms / args.steps / float(np.max(ks))}

    # context for the fraction: what plain library kernels stream on THIS GPU right now (not the contract's peak)
    try:
        free_b, _ = torch.cuda.mem_get_info(dev)
        n_el = int(min(4 << 30, free_b // 3) // 4)
        xs = torch.empty(n_el, dtype=torch.float32, device=dev).fill_(1.0)
        ys = torch.empty(n_el // 2, dtype=torch.float32, device=dev)

        def best_ms(fn, reps=4):
            fn()
            torch.cuda.synchronize()
            best = 1e30
            for _ in range(reps):
                a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a_.record()
                fn()
                b_.record()
                torch.cuda.synchronize()
                best = min(best, a_.elapsed_time(b_))
            return best
        roofline["context_torch_sum_read_gbs"] = n_el * 4 / best_ms(lambda: xs.sum()) / 1e6
        roofline["context_torch_copy_gbs"] = 2 * (n_el // 2) * 4 / best_ms(lambda: ys.copy_(xs[: n_el // 2])) / 1e6
        del xs, ys
    except Exception as e:
        roofline["context_error"] = repr(e)

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": dict(cfg, outlier_frac=outl, launches_per_step=1 + int(bool(cfg["inlier_mask"]) and not fused),
                           noise_sigma=0.002, result_gather=gather_mode),
            "roofline": roofline, "clocks": clk.summary(),
            "gpu_launches": int(launches), "input_generation_s": gen_s,
            "status_histogram": {str(k): int(v) for k, v in zip(*np.unique(res["status"], return_counts=True))},
            "mean_evaluations_per_query": float(res["n_evaluations"].mean())}
    if sustained:
        line["sustained"] = sustained
    if per_rank:
        line["per_rank"] = per_rank
    if gather_check:
        line["gather_check"] = gather_check

    # ---- e2e: same C ABI, inputs start in pinned HOST memory, results end in host memory
    near = sharding.host_near_gpu(local_rank)
    near.__enter__()  # pinned buffers on the GPU's NUMA node (matters when several ranks copy at once)
    interleaved = interleave_host_memory(True) if args.host_numa == "interleave" else False
    if not args.no_e2e:
        names = ("F", "Gx", "Gy", "ref", "pts3D")
        E = min(args.e2e_queries or B, B)
        t_pin = time.perf_counter()
        host = None
        while host is None:
            try:  # one pinned allocation per array kind; a view per query
                big = {k: torch.empty((E,) + tuple(getattr(scenes[0], k).shape), dtype=getattr(scenes[0], k).dtype,
                                      pin_memory=True) for k in names}
                for i in range(E):
                    for k in names:
                        big[k][i].copy_(getattr(scenes[i], k), non_blocking=True)
                torch.cuda.synchronize()
                host = [{k: big[k][i] for k in names} for i in range(E)]
            except RuntimeError:
                if E <= 32:
                    raise
                big = None
                E = max(32, E // 2)
        pin_s = time.perf_counter() - t_pin
        EF = min(E, 32)  # the full-upload variant needs a second copy of the inputs in HBM: a smaller step
        devbuf = [{k: torch.empty_like(getattr(scenes[i], k)) for k in names} for i in range(EF)]
        dev_small = [{k: torch.empty_like(getattr(scenes[i], k)) for k in ("ref", "pts3D")} for i in range(E)]
        act_e = torch.ones((E, args.points), dtype=torch.uint8, device=dev) if cfg["inlier_mask"] else None

        def batch_of(F, Gx, Gy, ref, pts, n, act=None):
            return fm.ProblemBatch([fm.Problem(F(i), Gx(i) if Gx else None, Gy(i) if Gy else None, ref(i), pts(i),
                                               scenes[i].K, scenes[i].R0, scenes[i].t0, 1024, 1024,
                                               active=act[i] if act is not None else None) for i in range(n)])

        eres = torch.empty((E, RESULT_DTYPE.itemsize), dtype=torch.uint8, device=dev)
        eres_host = torch.empty((E, RESULT_DTYPE.itemsize), dtype=torch.uint8).pin_memory()
        d2h = eres_host.numel()

        def refine_e2e(pb, pb_act, pb_na, n):
            if cfg["inlier_mask"] and not fused:
                fm.find_inliers_batch(pb_na, 0.8, loss=cfg["loss"], ctx=ctx, out=act_e)
            fm.refine_batch(pb if fused else pb_act, cfg["n_iters"], loss=cfg["loss"], lambda_=cfg["lambda_"],
                            ratio_threshold=cfg["ratio_threshold"], levels=levels, ctx=ctx, sync=False, out=eres,
                            inlier_threshold=0.8 if fused else None)
            eres_host[:n].copy_(eres[:n], non_blocking=True)

        def time_e2e(fn, n, ksteps):
            for _ in range(2):
                fn()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            ms = timed_launches(lambda i: fn(), ksteps)
            return max_over_ranks(ms, world, dev)

        # (1) every input uploaded to HBM first (EF queries per step)
        d_of = lambda k: (lambda i: devbuf[i][k])
        fprobs = batch_of(d_of("F"), d_of("Gx"), d_of("Gy"), d_of("ref"), d_of("pts3D"), EF)
        fprobs_act = batch_of(d_of("F"), d_of("Gx"), d_of("Gy"), d_of("ref"), d_of("pts3D"), EF, act_e) if act_e is not None else fprobs
        fprobs_na = batch_of(d_of("F"), None, None, d_of("ref"), d_of("pts3D"), EF)
        h2d_full = sum(host[i][k].numel() * host[i][k].element_size() for i in range(EF) for k in names)

        def full_step():
            for i in range(EF):
                for k in names:
                    devbuf[i][k].copy_(host[i][k], non_blocking=True)
            refine_e2e(fprobs, fprobs_act, fprobs_na, EF)

        ksteps = max(2, min(args.steps, 5))
        f_ms = time_e2e(full_step, EF, ksteps)
        line["e2e_full_upload"] = {"value": world * EF * ksteps / (f_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": int(h2d_full),
                                   "d2h_bytes_per_step": int(EF * RESULT_DTYPE.itemsize), "queries_per_step": EF, "steps": ksteps,
                                   "h2d_gbs": h2d_full * ksteps / (f_ms / 1e3) / 1e9,
                                   "note": "HWC fp32 maps + descriptors + points copied from pinned host memory to HBM, "
                                           "then the refinement launch, then the records D2H; all inside the timed region"}
        line["e2e"] = line["e2e_full_upload"]
        del devbuf, fprobs, fprobs_act, fprobs_na
        # (2) the maps LEFT in pinned host memory (zero-copy): the kernels fetch the rows they gather over PCIe
        # themselves, and with record reuse that is about one evaluation's worth of rows instead of three whole maps.
        # Variant "maps": F, Gx, Gy stay on the host, descriptors + points are uploaded (sequential DMA);
        # variant "all": the descriptors stay on the host as well.
        if not args.no_zero_copy:
            h_of = lambda k: (lambda i: host[i][k])
            s_of = lambda k: (lambda i: dev_small[i][k])
            zsteps = max(2, min(args.steps, 3 if E > 64 else 5))
            zc = {}
            for variant in ("maps", "all"):
                ref_of = s_of("ref") if variant == "maps" else h_of("ref")
                zp = batch_of(h_of("F"), h_of("Gx"), h_of("Gy"), ref_of, s_of("pts3D"), E)
                zp_act = batch_of(h_of("F"), h_of("Gx"), h_of("Gy"), ref_of, s_of("pts3D"), E, act_e) if act_e is not None else zp
                zp_na = batch_of(h_of("F"), None, None, ref_of, s_of("pts3D"), E)
                up = ("pts3D", "ref") if variant == "maps" else ("pts3D",)

                def zc_step():
                    for k in up:  # (one strided-free copy per query: the destination tensors are separate)
                        for i in range(E):
                            dev_small[i][k].copy_(host[i][k], non_blocking=True)
                    refine_e2e(zp, zp_act, zp_na, E)

                z_ms = time_e2e(zc_step, E, zsteps)
                zres = eres_host.numpy().view(RESULT_DTYPE).reshape(E)
                same = all(np.ascontiguousarray(zres[f]).tobytes() == np.ascontiguousarray(res[:E][f]).tobytes()
                           for f in ("R", "t", "best_cost", "best_num_inliers", "status"))
                host_maps = 3 if variant == "maps" else 4          # of the 4 rows of a gather, those that cross PCIe
                pcie = float(zres["gathered_rows"].astype(np.float64).sum()) * 4.0 * host_maps / 4.0
                if cfg["inlier_mask"] and not fused:  # (fused: the pass's rows are in gathered_rows already)
                    pcie += float(nsup_all[:E].double().sum().item()) * args.channels * 4.0 * (1 if variant == "maps" else 2)
                pcie += sum(host[i][k].numel() * host[i][k].element_size() for i in range(E) for k in up)
                zc[variant] = {"value": world * E * zsteps / (z_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": int(pcie),
                               "d2h_bytes_per_step": int(d2h), "h2d_gbs": pcie * zsteps / (z_ms / 1e3) / 1e9,
                               "results_equal_to_resident_run": bool(same)}
            best = max(zc, key=lambda k: zc[k]["value"])
            line["e2e"] = dict(zc[best], queries_per_step=E, steps=zsteps, variant=best, pinned_setup_s=pin_s,
                               host_numa="interleave" if interleaved else "local",
                               note="same C ABI, inputs in pinned HOST memory, results back in host memory, all inside "
                                    "the timed region.  The maps are not uploaded: pinned memory is mapped into the "
                                    "device's address space and the kernels read the rows they gather over PCIe "
                                    "(h2d_bytes_per_step = rows read + the uploaded arrays, counted from the kernels' "
                                    "row counters); `e2e_full_upload` is the same step with every input copied to HBM "
                                    "first, `e2e_zero_copy` has both variants")
            line["e2e_zero_copy"] = zc
        del host, dev_small, big

        # ---- e2e one level up (the reference's feature_pnp): hypercolumns in pinned host memory -> H2D ->
        # Sobel + pack kernel, descriptor gather kernel -> the same refinement -> results D2H
        if not args.no_adapter_e2e:
            A = min(E, 32)
            qh_host = [scenes[i].F.permute(2, 0, 1).contiguous().cpu().pin_memory() for i in range(A)]
            pts_host = [scenes[i].pts3D.cpu().pin_memory() for i in range(A)]
            cells = []
            for i in range(A):  # reference keypoints: the ground-truth cell of every point in the query map
                P = scenes[i].pts3D.cpu()
                u = (scenes[i].K[0, 0] * P[:, 0] / P[:, 2] + scenes[i].K[0, 2]).clamp(0, 1023)
                v = (scenes[i].K[1, 1] * P[:, 1] / P[:, 2] + scenes[i].K[1, 2]).clamp(0, 1023)
                cells.append(((v.long() * args.map // 1024).int().to(dev), (u.long() * args.map // 1024).int().to(dev)))
            qh_dev = [torch.empty_like(q, device=dev) for q in qh_host]
            rh_dev = [torch.empty_like(q, device=dev) for q in qh_host]
            pts_dev = [torch.empty_like(q, device=dev) for q in pts_host]
            ares = torch.empty((A, RESULT_DTYPE.itemsize), dtype=torch.uint8, device=dev)
            ares_host = torch.empty((A, RESULT_DTYPE.itemsize), dtype=torch.uint8).pin_memory()
            a_h2d = sum(2 * q.numel() * 4 for q in qh_host) + sum(q.numel() * 8 for q in pts_host)

            def adapter_step():
                aprobs = []
                for i in range(A):
                    qh_dev[i].copy_(qh_host[i], non_blocking=True)
                    rh_dev[i].copy_(qh_host[i], non_blocking=True)  # the reference image's hypercolumn
                    pts_dev[i].copy_(pts_host[i], non_blocking=True)
                    F_, Gx_, Gy_ = fm.sobel_pack(qh_dev[i], ctx=ctx)
                    ref_ = fm.gather_descriptors(rh_dev[i], cells[i][0], cells[i][1], ctx=ctx)
                    aprobs.append(fm.Problem(F_, Gx_, Gy_, ref_, pts_dev[i], scenes[i].K, scenes[i].R0, scenes[i].t0, 1024, 1024))
                fm.refine_batch(aprobs, cfg["n_iters"], loss=cfg["loss"], lambda_=cfg["lambda_"],
                                ratio_threshold=cfg["ratio_threshold"], levels=levels, ctx=ctx, sync=False, out=ares)
                ares_host.copy_(ares, non_blocking=True)

            for _ in range(2):
                adapter_step()
            torch.cuda.synchronize()
            a_ms = timed_launches(lambda i: adapter_step(), ksteps)
            line["e2e_adapter"] = {"value": A * ksteps / (a_ms / 1e3), "unit": UNIT + " per GPU", "h2d_bytes_per_step": int(a_h2d),
                                   "d2h_bytes_per_step": int(ares_host.numel()), "queries_per_step": A,
                                   "note": "feature_pnp level: query + reference hypercolumns (CHW fp32) and points in pinned "
                                           "host memory -> H2D -> fmpnp_sobel_pack + fmpnp_gather_descriptors -> refinement "
                                           "-> D2H; rank-local, not aggregated"}
            # ---- the pipeline call itself: hypercolumns already on the GPU (where the CNN leaves them), one
            # query per call as s2dhm does: build_problem (Sobel + pack + descriptor gather) -> refinement -> pose
            if levels is None:
                import collections
                import model as dropin
                from helpers import utils as hu
                from pose_prediction import optimize_feature_pnp as adapter
                Pred = collections.namedtuple("Pred", "points_3d reference_inliers matrix")
                preds = []
                for i in range(min(8, A)):
                    P = scenes[i].pts3D.cpu()
                    u = (scenes[i].K[0, 0] * P[:, 0] / P[:, 2] + scenes[i].K[0, 2]).clamp(0, 1023)
                    v = (scenes[i].K[1, 1] * P[:, 1] / P[:, 2] + scenes[i].K[1, 2]).clamp(0, 1023)
                    T = np.eye(4)
                    T[:3, :3], T[:3, 3] = scenes[i].R0.numpy(), scenes[i].t0.numpy()
                    preds.append(Pred(P.numpy(), torch.stack([u, v], 1).numpy(), T))
                mdl = dropin.sparseFeaturePnP(n_iters=cfg["n_iters"], loss_fn=getattr(hu, cfg["loss"]),
                                              lambda_=cfg["lambda_"], ratio_threshold=cfg["ratio_threshold"])

                def pipeline_query(i):
                    prob = adapter.build_problem(qh_dev[i], rh_dev[i], preds[i], scenes[i].K, [1024, 1024])
                    return mdl.forward_packed(prob)

                pipeline_query(0)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                for i in range(len(preds)):
                    pipeline_query(i)
                torch.cuda.synchronize()
                dt = time.perf_counter() - t0
                line["pipeline_call"] = {"ms_per_query": 1e3 * dt / len(preds), "queries": len(preds),
                                         "note": "one query per call, hypercolumns resident on the GPU: build_problem "
                                                 "(fmpnp_sobel_pack + fmpnp_gather_descriptors) -> forward_packed -> "
                                                 "CPU float64 pose; wall clock"}
            del qh_host, qh_dev, rh_dev

    # ---- e2e at the reference's own Python call: sparseFeaturePnP.forward, one query per call, CPU float64 CHW
    # tensors in (what s2dhm hands over today), CPU float64 pose out.  Rank-local, cfg3/cfg2 losses only.
    if not args.no_e2e and levels is None:
        import model as dropin
        from helpers import utils as hu
        D = min(4, B)
        ins = [scenes[i].reference_inputs() for i in range(D)]
        m = dropin.sparseFeaturePnP(n_iters=cfg["n_iters"], loss_fn=getattr(hu, cfg["loss"]), lambda_=cfg["lambda_"],
                                    ratio_threshold=cfg["ratio_threshold"])

        def dropin_query(x):
            dropin.clear_pack_cache()
            return m(x["pts3D"], x["feature_ref"], x["feature_map_query"], x["feature_grad_x"], x["feature_grad_y"],
                     x["K"], x["im_width"], x["im_height"], R_init=x["R_init"], t_init=x["t_init"])

        dropin_query(ins[0])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for x in ins:
            dropin_query(x)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        per_q = sum(x[k].numel() * x[k].element_size() for x in ins[:1]
                    for k in ("pts3D", "feature_ref", "feature_map_query", "feature_grad_x", "feature_grad_y"))
        line["e2e_dropin"] = {"value": D / dt, "unit": UNIT + " per GPU", "h2d_bytes_per_query": int(per_q), "queries": D,
                              "note": "model.sparseFeaturePnP.forward per query: pageable CPU float64 CHW tensors -> float32 "
                                      "in pinned staging (host threads) -> H2D -> pack kernel -> refinement -> CPU float64 "
                                      "pose; wall clock, no inlier mask.  Bound by the host reading 436 MB of float64 per "
                                      "query, not by the GPU"}
        del ins
    if interleaved:
        interleave_host_memory(False)
    near.__exit__(None, None, None)
    if "e2e" in line:
        line["e2e"]["host_cores_near_gpu"] = near.cores

    # ---- CPU baseline beside it (rank 0, N = 1 only): the oracle port on a bounded sample, and its final poses
    # against the records of the launch that was timed (same scenes)
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = host_threads()
        nq = min(args.cpu_queries or threads, B)
        try:
            problems = cpu_problems(args, cfg, nq, scenes=scenes)
            dt, cpu_out = cpu_pass(cfg, problems, threads)
            line["cpu_baseline"] = {"value": nq / dt, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": "%d full-shape queries (seeds 0..%d of the same workload), one per thread, "
                                              "%.1f s wall" % (nq, nq - 1, dt),
                                    "note": "C port of the reference's algorithm; the unmodified PyTorch reference needs "
                                            "11.8 s per query of this shape with 8 threads (build container, "
                                            "tests/golden/reference_cpu_timing.json) and cannot travel to this box"}
            line["parity_vs_cpu_arm"] = parity_vs_cpu(cpu_out, res[:nq])
        except Exception as e:  # the baseline must never take the GPU number down with it
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": "failed: %r" % (e,)}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        if shared is not None:
            shared.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
