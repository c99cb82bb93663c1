#!/usr/bin/env python
"""Static opcode summary of the shipped library: `cuobjdump -sass lumenos_b200/liblumenos_b200.so`, per kernel the counts of the
mnemonics that carry the design claims -- UBLKCP (cp.async.bulk: 1-D TMA bulk copies), SYNCS (mbarrier), IMAD.WIDE (the 32x32->64
multiplies of the 64-bit modular arithmetic), IADD3 / .X carry chains, LDG/STG/LDS/STS, and tensor-core ops (must be 0).

    python tools/sass_summary.py > profiles/sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "lumenos_b200", "liblumenos_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
arch = sorted(set(re.findall(r"arch = (sm_\w+)", out)))
kern, stats = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = re.sub(r"\(.*", "", kern).replace("lmn::", "").replace("void ", "")
        stats[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(1)
        c = stats[kern]
        c["total"] += 1
        base = op.split(".")[0]
        if base == "UBLKCP": c["UBLKCP"] += 1
        elif base == "SYNCS": c["SYNCS"] += 1
        elif op.startswith("IMAD.WIDE"): c["IMAD.WIDE"] += 1
        elif base == "IMAD": c["IMAD"] += 1
        elif base in ("IADD3", "IADD", "UIADD3"):
            c["IADD3"] += 1
            if ".X" in op: c["IADD3.X"] += 1
        elif base in ("LDG", "STG", "LDS", "STS", "LDL", "STL"): c[base] += 1
        elif base in ("HMMA", "IMMA", "DMMA", "UTCHMMA", "UTCIMMA", "UTCQMMA", "UTMALDG", "UTMASTG"): c["tensor/TMA-map"] += 1
cols = ["total", "IMAD.WIDE", "IMAD", "IADD3", "IADD3.X", "UBLKCP", "SYNCS", "LDG", "STG", "LDS", "STS", "LDL", "STL", "tensor/TMA-map"]
print("# static SASS opcode counts of %s (%s)" % (os.path.relpath(so, ROOT), ", ".join(arch)))
print("# %-46s " % "kernel" + " ".join("%9s" % c for c in cols))
tot = collections.Counter()
for kname, c in stats.items():
    print("%-48s " % kname[:48] + " ".join("%9d" % c[x] for x in cols))
    tot.update(c)
print("%-48s " % "ALL KERNELS" + " ".join("%9d" % tot[x] for x in cols))
