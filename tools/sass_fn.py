#!/usr/bin/env python
"""Static opcode count of one function of liblumenos_b200.so (cuobjdump -sass), optionally between two addresses.
    python tools/sass_fn.py <substring of mangled name> [lo_hex hi_hex]"""
import collections, re, subprocess, sys, os
so = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "lumenos_b200", "liblumenos_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout.splitlines()
key = sys.argv[1]
lo = int(sys.argv[2], 16) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3], 16) if len(sys.argv) > 3 else 1 << 60
on, mix, lines = False, collections.Counter(), []
for l in txt:
    if "Function :" in l:
        on = key in l
        if on: print(l.strip())
        continue
    if not on: continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if not m: continue
    addr, ins = int(m.group(1), 16), m.group(2).strip()
    if addr < lo or addr >= hi: continue
    lines.append((addr, ins))
    if ins.startswith("@"): ins = ins.split(None, 1)[1]
    op = ins.split()[0]
    p = op.split(".")
    k = p[0]
    if k == "IMAD":
        for sub in ("WIDE", "MOV", "IADD", "SHL", "HI"):
            if sub in p: k = "IMAD." + sub; break
    mix[k] += 1
tot = sum(mix.values())
print("instructions", tot)
for k, v in mix.most_common(20): print("  %-12s %6d" % (k, v))
heavy = 4 * mix["IMAD.WIDE"] + 4 * mix["IMAD.HI"] + 2 * (mix["IMAD"] + mix["IMAD.MOV"] + mix["IMAD.IADD"] + mix["IMAD.SHL"])
print("fmaheavy cycles/warp (static) =", heavy)
if os.environ.get("DUMP"):
    for a, i in lines: print("%05x  %s" % (a, i))
