#!/usr/bin/env python
"""Opcode mix of one kernel from an ncu report's source page (SASS view), weighted by executed warp instructions.

    python tools/sass_mix.py gpurun_out/x.ncu-rep k_mac_ra [launch_index]
IMAD.WIDE occupies the fmaheavy pipe for 4 cycles per warp, IMAD for 2 (tools/int_pipe_probe.cu)."""
import collections
import csv
import io
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern, "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h = next(i for i, r in enumerate(rows) if "Source" in r)
ix = {n: i for i, n in enumerate(rows[h])}
mix = collections.Counter()
tot = 0
for r in rows[h + 1:]:
    if len(r) <= ix["Instructions Executed"]:
        continue
    if r[ix["Source"]] == "Source" or r[0] == "Kernel Name":
        break   # next launch of the same kernel
    s = r[ix["Source"]].strip()
    if s.startswith("@"):
        s = s.split(None, 1)[1]
    op = s.split()[0].rstrip(";")
    n = int(r[ix["Instructions Executed"]] or 0)
    parts = op.split(".")
    key = parts[0]
    if key == "IMAD":
        for sub in ("WIDE", "MOV", "IADD", "SHL", "HI"):
            if sub in parts:
                key = "IMAD." + sub
                break
    mix[key] += n
    tot += n
print("kernel", rows[0][1][:100] if rows and len(rows[0]) > 1 else kern)
print("total warp instructions %d" % tot)
heavy = 4 * mix["IMAD.WIDE"] + 4 * mix["IMAD.HI"] + 2 * (mix["IMAD"] + mix["IMAD.MOV"] + mix["IMAD.IADD"] + mix["IMAD.SHL"])
for k, v in mix.most_common(28):
    print("  %-12s %12d  %5.1f%%" % (k, v, 100.0 * v / tot))
print("fmaheavy pipe cycles (4*WIDE + 2*IMAD*) = %d  (wide part %.1f%%)" % (heavy, 100.0 * 4 * mix["IMAD.WIDE"] / max(1, heavy)))
