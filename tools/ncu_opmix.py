#!/usr/bin/env python
"""Per-trellis-step SASS opcode mix of a kernel from an .ncu-rep captured with --import-source on:
    python tools/ncu_opmix.py REPORT.ncu-rep WARPS_PER_LAUNCH STEPS
Prints executed warp-instructions per warp and trellis step, by opcode (the numbers DESIGN.md's pipe-balance
notes quote)."""
import collections
import csv
import re
import subprocess
import sys


def main():
    rep, warps, steps = sys.argv[1], float(sys.argv[2]), float(sys.argv[3])
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    hdr = rows[1]
    i_s, i_e = hdr.index("Source"), hdr.index("Instructions Executed")
    hist, tot = collections.Counter(), 0
    for r in rows[2:]:
        try:
            c = int(r[i_e])
        except (ValueError, IndexError):
            continue
        m = re.match(r"(@!?U?P\d+\s+)?([A-Za-z0-9_.]+)", r[i_s].strip())
        hist[m.group(2) if m else "?"] += c
        tot += c
    print(f"{rep}: {tot / warps / steps:.1f} warp-instructions per warp and step")
    for op, c in hist.most_common(16):
        print(f"  {op:24s} {c / warps / steps:7.1f}")


if __name__ == "__main__":
    main()
