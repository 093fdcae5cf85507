#!/usr/bin/env python
"""Dynamic SASS listing of one kernel of an .ncu-rep with source lines attached.

    python profiles/tools/sass_dump.py <report.ncu-rep> <kernel substring> <lib.so> > listing.txt

One line per SASS instruction of the first profiled launch:
    offset  file:line  executed-warp-instructions  stall-samples  instruction
"""
import csv
import subprocess
import sys

from line_profile import sass_lines


def main():
    rep, kernel, lib = sys.argv[1:4]
    lines = sass_lines(lib, kernel)
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    hdr = base = None
    active, launches = False, 0
    for r in csv.reader(txt.splitlines()):
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
        print("%05x %s:%d %d %d  %s" % (addr - base, loc[0], loc[1], n, s, r[1]))


if __name__ == "__main__":
    main()
