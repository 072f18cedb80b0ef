#!/usr/bin/env python
"""Opcode histogram and hot-loop excerpt of a kernel in the built library (no GPU needed).
    python tools/sass_histogram.py [kernel substring] > profiles/rNN/sass_<kernel>.md
Uses cuobjdump / nvdisasm on featuremetric-pnp_b200/lib/libfmpnp.so."""
import collections, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "featuremetric-pnp_b200", "lib", "libfmpnp.so")


def disasm(kernel):
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
        m = re.search(r"^\s+/\*([0-9a-f]+)\*/\s+(.*?);", l)
        if m:
            out.append((cur, m.group(2).strip()))
    return out


def opcode(ins, full=False):
    t = ins.split()
    op = t[1] if t[0].startswith("@") else t[0]
    return op if full else op.split(".")[0]


if __name__ == "__main__":
    kernel = sys.argv[1] if len(sys.argv) > 1 else "refine_tasks_kernel"
    seq = disasm(kernel)
    print("# SASS of `%s` (sm_100a, %d instructions; tools/sass_histogram.py on the built libfmpnp.so)\n" % (kernel, len(seq)))
    by = collections.Counter(opcode(i) for _, i in seq)
    print("## Opcode histogram (static)\n")
    print("| opcode | count |\n|---|---|")
    for k, v in by.most_common(40):
        print("| %s | %d |" % (k, v))
    full = collections.Counter(opcode(i, True) for _, i in seq)
    print("\nData movement / tensor opcodes of interest: " + ", ".join(
        "%s %d" % (k, sum(v for o, v in full.items() if o.startswith(k))) for k in
        ("LDGSTS", "UBLKCP", "UTMALDG", "UTCHMMA", "LDTM", "HMMA", "DFMA", "DADD", "DMUL", "F2F", "LDG", "STG", "LDS", "STS", "LDL", "STL")))
    print("\nLDGSTS variants: " + ", ".join("%s %d" % kv for kv in full.most_common() if kv[0].startswith("LDGSTS")))
    # hot loop: the longest run of instructions attributed to copy_rows_full / acc1 (fmpnp_stream.cuh)
    print("\n## Row copy of a long-row item (`copy_rows_full`, fmpnp_stream.cuh): one address per row, immediate offsets\n\n```")
    shown = 0
    for i, (loc, ins) in enumerate(seq):
        if "LDGSTS" in ins and shown < 24 and "0x" in ins:
            print(ins)
            shown += 1
    print("```\n\n## Channel accumulation (`acc4` / `acc1`): LDS.128, 4 F2F.F64.F32 + 1 DADD + 6 DFMA per channel\n\n```")
    start = next((i for i, (loc, ins) in enumerate(seq) if loc and loc[0] == "fmpnp_stream.cuh" and ins.startswith("F2F.F64.F32")), None)
    if start is not None:
        for loc, ins in seq[max(0, start - 6):start + 40]:
            print(ins)
    print("```")
