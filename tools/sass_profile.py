#!/usr/bin/env python
"""Per-opcode and hottest-instruction summary of an ncu report's SASS page.
    python tools/sass_profile.py gpurun_out/x.ncu-rep [min_count_millions]"""
import collections, csv, subprocess, sys

def load(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    H = rows[1]
    isrc, iex, isamp = H.index("Source"), H.index("Instructions Executed"), H.index("# Samples")
    data = []
    for r in rows[2:]:
        try:
            data.append((int(r[iex]), r[isrc], int(r[isamp] or 0)))
        except (ValueError, IndexError):
            pass
    return data

def opcode(src):
    t = src.split()
    op = t[1] if t[0].startswith("@") else t[0]
    return op.rstrip(";").split(".")[0]

if __name__ == "__main__":
    data = load(sys.argv[1])
    tot = sum(d[0] for d in data)
    tots = sum(d[2] for d in data)
    by, bys = collections.Counter(), collections.Counter()
    for n, s, sm in data:
        by[opcode(s)] += n
        bys[opcode(s)] += sm
    print("total warp instructions %.3e, samples %d" % (tot, tots))
    for k, v in by.most_common(18):
        print("%-10s %6.2f%% inst  %6.2f%% samples" % (k, 100 * v / tot, 100 * bys[k] / tots))
    if len(sys.argv) > 2:
        thr = float(sys.argv[2]) * 1e6
        for i, (n, s, sm) in enumerate(data):
            if n >= thr:
                print(i, "%.0fM" % (n / 1e6), sm, s[:100])
