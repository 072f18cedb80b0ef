#!/usr/bin/env python
"""Turns ncu outputs brought back in gpurun_out/ into the small text summaries committed under profiles/.

    python tools/summarize_profiles.py launches gpurun_out/launches_r01.csv profiles/r01_launch_list.md
    python tools/summarize_profiles.py kernel   gpurun_out/prof.ncu-rep     profiles/r01_refine_kernel.md [alg_bytes]
"""
import collections
import csv
import json
import os
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_bytes.sum.per_second", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active", "sm__cycles_elapsed.avg.per_second",
]


def ms(value, unit):
    v = float(value.replace(",", ""))
    return {"ns": v / 1e6, "us": v / 1e3, "ms": v, "s": v * 1e3}.get(unit, v)


def launches(src, dst):
    rows = list(csv.reader(open(src)))
    h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    H = rows[h]
    ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
    agg, seq = collections.OrderedDict(), []
    for r in rows[h + 1:]:
        if len(r) <= vi:
            continue
        name = r[ki].split("(")[0]
        v = ms(r[vi], r[ui])
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        seq.append((name, v))
    tot = sum(a[1] for a in agg.values())
    ours = [(n, v) for n, v in seq if n.startswith("fmpnp::")]
    tot_ours = sum(v for _, v in ours)
    with open(dst, "w") as f:
        f.write("# ncu launch list (`--metrics gpu__time_duration.sum --clock-control none`)\n\n")
        f.write("source: `%s` — %d launches, %.2f ms of device time in total.\n" % (os.path.basename(src), len(seq), tot))
        f.write("Per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes.\n\n")
        f.write("## library kernels of this repo (the timed region launches only these)\n\n| # | kernel | ms |\n|---|---|---|\n")
        for i, (n, v) in enumerate(ours):
            f.write("| %d | `%s` | %.3f |\n" % (i, n, v))
        f.write("\nshare inside the hot-path step (fmpnp kernels only):\n\n| kernel | launches | ms | share |\n|---|---|---|---|\n")
        oa = collections.OrderedDict()
        for n, v in ours:
            a = oa.setdefault(n, [0, 0.0])
            a[0] += 1
            a[1] += v
        for n, (c, v) in oa.items():
            f.write("| `%s` | %d | %.3f | %.4f |\n" % (n, c, v, v / tot_ours if tot_ours else 0))
        f.write("\n## everything else in the process (synthetic-input generation by torch, outside the timed region)\n\n")
        f.write("| kernel | launches | ms | share of process |\n|---|---|---|---|\n")
        for n, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:15]:
            f.write("| `%s` | %d | %.3f | %.4f |\n" % (n[:90], c, v, v / tot))


def kernel(src, dst, alg_bytes=None):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    H, U, V = rows[0], rows[1], rows[2]
    d = {h: (U[i], V[i]) for i, h in enumerate(H)}
    stalls = sorted(((h, float(v[1].replace(",", ""))) for h, v in d.items()
                     if "issue_stalled" in h and h.endswith("per_issue_active.ratio")), key=lambda x: -x[1])
    with open(dst, "w") as f:
        f.write("# ncu --set full: `%s`\n\nsource report: `%s` (kernel `%s`)\n\n| metric | unit | value |\n|---|---|---|\n" % (
            os.path.basename(dst), os.path.basename(src), d.get("Kernel Name", ("", "?"))[1]))
        for k in KEYS:
            if k in d:
                f.write("| `%s` | %s | %s |\n" % (k, d[k][0], d[k][1]))
        rd = d.get("dram__bytes_read.sum")
        wr = d.get("dram__bytes_write.sum")
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}
        if rd and wr:
            traffic = float(rd[1].replace(",", "")) * scale[rd[0]] + float(wr[1].replace(",", "")) * scale[wr[0]]
            f.write("\nDRAM traffic per launch = read + write = **%.4e B**" % traffic)
            if alg_bytes:
                f.write("; algorithmic bytes of the same launch = %.4e B; ratio = **%.4f**" % (alg_bytes, traffic / alg_bytes))
            f.write("\n")
        f.write("\n## warp stall reasons (average stalled warps per issue-active cycle)\n\n| reason | ratio |\n|---|---|\n")
        for h, v in stalls[:10]:
            f.write("| %s | %.3f |\n" % (h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v))
    return d


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3])
    else:
        kernel(sys.argv[2], sys.argv[3], float(sys.argv[4]) if len(sys.argv) > 4 else None)
