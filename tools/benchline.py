#!/usr/bin/env python
"""prints the key numbers of bench.py JSON lines (the last line of each file is the JSON line)"""
import json
import sys
for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    e = d.get("e2e") or {}
    print(f, "| s/step %.4f e2e %s | parity %s | roofline %.3f | merkle %s | enc %s dec %s" % (
        d["value"], e.get("value"), str(d.get("parity_check"))[:3], d["roofline"]["frac"], (d.get("commit_merkle") or {}).get("ms"),
        (d.get("encrypt_matrix") or {}).get("ms"), (d.get("client_decrypt") or {}).get("ms")))
    print("   ", {k: round(v["frac"], 3) for k, v in d.get("kernel_rooflines", {}).items()})
