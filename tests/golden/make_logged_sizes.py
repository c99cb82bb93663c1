"""Generates tests/golden/logged_sizes.json from the reference's own server logs (run in the build container, where
/root/reference exists): the "Marshaled MatR / MatZ / QueriedCols" lines (fhe/ligero.go:672,682,692) and "Marshaled encrypted
proof length" (cmd/server/main.go:250) of results/{baseline,experimental}/server/bench_*.txt, i.e. humanize.Bytes() of the
serialized sizes -- the only reference-held data that constrains Lattigo's ciphertext framing."""
import glob
import json
import os
import re

REF = "/root/reference/results"
out = []
for kind in ("baseline", "experimental"):
    for f in sorted(glob.glob(os.path.join(REF, kind, "server", "bench_*.txt"))):
        rows, cols, logn = map(int, re.match(r"bench_(\d+)x(\d+)_(\d+)\.txt", os.path.basename(f)).groups())
        rec = {"kind": kind, "rows": rows, "cols": cols, "logN": logn, "file": os.path.relpath(f, "/root/reference")}
        for line in open(f):
            m = re.match(r"Marshaled (MatR|MatZ|QueriedCols|encrypted proof length): (.*)$", line.strip())
            if m:
                rec[{"encrypted proof length": "proof"}.get(m.group(1), m.group(1))] = m.group(2)
            m = re.match(r"Number of queried columns: (\d+)", line.strip())
            if m:
                rec["queries"] = int(m.group(1))
        out.append(rec)
# "Marshaled keys length" of the client logs (cmd/client/main.go:136): humanize.Bytes of the JSON body that carries pk, rlk, the Galois keys of
# GaloisElementsForInnerSum(1, rows) and (experimental runs) the ring-switch evk, each as base64 of its MarshalBinary
for rec in out:
    f = os.path.join(REF, rec["kind"], "client", os.path.basename(rec["file"]))
    for line in open(f):
        m = re.match(r"Marshaled keys length: (.*)$", line.strip())
        if m:
            rec["keys"] = m.group(1)
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "logged_sizes.json"), "w"), indent=1)
print(len(out), "logs")
