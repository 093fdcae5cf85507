#!/usr/bin/env python
"""Per-source-line instruction counts and stall samples for one kernel of an .ncu-rep.

    python profiles/tools/line_profile.py <report.ncu-rep> <kernel substring> [lib.so]

ncu's CLI `--page source --csv` only prints the SASS view; this joins it (by instruction offset)
with `nvdisasm -g` line info from the cubin inside the shared library, then aggregates executed
warp instructions and stall samples per CUDA source line and per SASS opcode.
"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def sass_lines(lib, kernel):
    tmp = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL)
    cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], cwd=tmp, capture_output=True, text=True).stdout
    out, on, cur = {}, False, ("?", 0)
    for ln in txt.splitlines():
        if ln.startswith(".text."):
            on = kernel in ln
            continue
        if not on:
            continue
        m = re.match(r'\s*//## File "(.*)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            out[int(m.group(1), 16)] = (cur, m.group(2).strip())
    return out


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    lib = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "safelife_k2_b200", "libsafelife_b200.so")
    csvtxt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(csvtxt.splitlines()))
    lines = sass_lines(lib, kernel)
    per_line = collections.Counter()
    per_line_samples = collections.Counter()
    per_op = collections.Counter()
    launches = 0
    hdr, base, active = None, None, False
    for r in rows:
        if not r:
            continue
        if r[0] == "Kernel Name":
            active = kernel in r[1]
            launches += active
            hdr = base = None
            continue
        if r[0] == "Address":
            hdr = r
            continue
        if not active or hdr is None or launches != 1:
            continue
        addr = int(r[0], 16)
        if base is None:
            base = addr
        n = int(r[hdr.index("Instructions Executed")] or 0)
        s = int(r[hdr.index("# Samples")] or 0)
        loc, _ = lines.get(addr - base, (("?", 0), ""))
        per_line[loc] += n
        per_line_samples[loc] += s
        op = r[1].split()[0] if not r[1].strip().startswith("@") else r[1].split()[1]
        per_op[op.split(".")[0]] += n
    tot = sum(per_line.values())
    tots = sum(per_line_samples.values())
    print("kernel %s: %d warp instructions, %d stall samples (first profiled launch)" % (kernel, tot, tots))
    print("\n-- by source line (>=0.5%% of instructions or samples) --")
    for loc, n in sorted(per_line.items(), key=lambda kv: (kv[0][0], kv[0][1])):
        s = per_line_samples[loc]
        if n >= 0.005 * tot or s >= 0.005 * tots:
            print("%-22s:%4d  inst %9d (%5.1f%%)  samples %6d (%5.1f%%)" % (loc[0], loc[1], n, 100 * n / tot, s, 100 * s / max(tots, 1)))
    print("\n-- by file --")
    byfile = collections.Counter()
    for loc, n in per_line.items():
        byfile[loc[0]] += n
    for f, n in byfile.most_common():
        print("%-28s %9d (%5.1f%%)" % (f, n, 100 * n / tot))
    print("\n-- by opcode --")
    for op, n in per_op.most_common(25):
        print("%-10s %9d (%5.1f%%)" % (op, n, 100 * n / tot))


if __name__ == "__main__":
    main()
