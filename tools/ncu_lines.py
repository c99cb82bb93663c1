#!/usr/bin/env python
"""Stall samples and executed instructions per CUDA source line of one kernel of an ncu report.
    python tools/ncu_lines.py report.ncu-rep kernel_regex [top]"""
import csv
import io
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kern,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
fil, hdr, out, ix = None, None, [], {}
for r in csv.reader(io.StringIO(raw)):
    if not r:
        continue
    if r[0] == "File Path":
        fil = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        ix = {n: i for i, n in enumerate(r)}
        continue
    if hdr and r[0].isdigit():
        num = lambda v: int(v) if v.isdigit() else 0   # noqa: E731
        s = num(r[ix["# Samples"]])
        ie = num(r[ix["Instructions Executed"]])
        st = {n[6:]: int(r[i]) for n, i in ix.items() if n.startswith("stall_") and "Not Issued" not in n and i < len(r) and r[i].isdigit()}
        out.append((fil, int(r[0]), s, ie, r[1].strip()[:100], st))
tot = sum(o[2] for o in out)
toti = sum(o[3] for o in out)
print("samples", tot, "warp instructions", toti)
for o in sorted(out, key=lambda x: -x[2])[:top]:
    st = sorted(o[5].items(), key=lambda kv: -kv[1])[:3]
    print("%-13s %4d %5.1f%% samp %5.1f%% inst  %-60s %s" % (o[0], o[1], 100 * o[2] / max(1, tot), 100 * o[3] / max(1, toti), o[4][:60], st))
