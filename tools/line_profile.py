#!/usr/bin/env python
"""Source-line view of an ncu capture made with --import-source on: joins the SASS page of the report (stall samples,
instructions executed per SASS instruction) with nvdisasm's line table of the same kernel in the library that was
profiled.
    python tools/line_profile.py gpurun_out/x.ncu-rep [kernel substring] [top N]
Prints the share of samples / executed warp instructions per source file and per source line, with the leading stall
reasons of each line."""
import collections, csv, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "featuremetric-pnp_b200", "lib", "libfmpnp.so")


def line_table(kernel):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", LIB], cwd=tmp, capture_output=True)
    cubin = [f for f in os.listdir(tmp) if f.startswith("fmpnp_kernels.") and f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
    fn, cur, out = None, None, []
    for l in txt.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
        if m:
            fn, cur = m.group(1), None
            continue
        if fn is None or kernel not in fn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        if re.search(r"^\s+/\*[0-9a-f]+\*/", l):
            out.append(cur)
    return out


def sass_rows(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    H = rows[1]
    return H, rows[2:]


if __name__ == "__main__":
    rep = sys.argv[1]
    kernel = sys.argv[2] if len(sys.argv) > 2 else "refine_tasks_kernel"
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    lt = line_table(kernel)
    H, rows = sass_rows(rep)
    assert len(lt) == len(rows), (len(lt), len(rows), "the library is not the one that was profiled")
    isamp, iex = H.index("# Samples"), H.index("Instructions Executed")
    stall = [(i, h) for i, h in enumerate(H) if h.startswith("stall_") and "Not Issued" not in h]
    by_line = collections.defaultdict(lambda: [0, 0, collections.Counter()])
    by_file = collections.defaultdict(lambda: [0, 0])
    ts = te = 0
    for loc, r in zip(lt, rows):
        s, e = int(r[isamp] or 0), int(r[iex] or 0)
        ts += s
        te += e
        b = by_line[loc]
        b[0] += s
        b[1] += e
        for i, h in stall:
            v = int(r[i] or 0)
            if v:
                b[2][h[6:]] += v
        f = by_file[loc[0] if loc else None]
        f[0] += s
        f[1] += e
    print("samples %d, warp instructions executed %.4g" % (ts, te))
    for f, (s, e) in sorted(by_file.items(), key=lambda x: -x[1][0]):
        print("  %-28s %5.1f%% samples %5.1f%% instructions" % (f, 100.0 * s / ts, 100.0 * e / te))
    print("top lines by samples:")
    for loc, (s, e, st) in sorted(by_line.items(), key=lambda x: -x[1][0])[:top]:
        lead = ", ".join("%s %.0f%%" % (k, 100.0 * v / max(1, sum(st.values()))) for k, v in st.most_common(3))
        print("  %-24s %5.1f%% samples %5.1f%% instr   %s" % ("%s:%d" % loc if loc else "?", 100.0 * s / ts, 100.0 * e / te, lead))
