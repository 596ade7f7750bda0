#!/usr/bin/env python
"""Instruction mix + stall hot spots of the first kernel in an .ncu-rep (source page, SASS)."""
import collections, csv, io, subprocess, sys
path = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 20
out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
print(rows[0][1])
hdr = rows[1]
iS, iE, iW = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
ops, stalls, body, n = collections.Counter(), collections.Counter(), [], 0
for r in rows[2:]:
    if r[0] == "Kernel Name": break
    if len(r) <= iE: continue
    src = r[iS].strip(); ex = int(r[iE] or 0); st = int(r[iW] or 0)
    t = src.split(); op = t[0] if not t[0].startswith("@") else t[1]
    op = ".".join(op.split(".")[:2]) if op.startswith(("LD", "ST")) else op.split(".")[0]
    ops[op] += ex; stalls[op] += st; n += ex; body.append((ex, st, src))
mx = max(e for e, _, _ in body)
print("total warp-inst", n, " hot-loop trips (max line count)", mx, " inst/trip %.1f" % (n / mx))
for op, c in ops.most_common(topn): print(f"{op:12s} {c:10d} {100*c/n:5.1f}%  per-trip {c/mx:6.1f}  stall-samples {stalls[op]}")
print("--- top stall lines")
for ex, st, src in sorted(body, key=lambda t: -t[1])[:topn]: print(st, ex, src)
