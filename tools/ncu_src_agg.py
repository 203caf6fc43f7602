#!/usr/bin/env python
"""Aggregates an `ncu --page source --csv` export: executed warp instructions and stall samples per opcode.
usage: ncu -i X.ncu-rep --page source --csv > src.csv ; python tools/ncu_src_agg.py src.csv [kernel_index]"""
import csv, collections, re, sys
rows = list(csv.reader(open(sys.argv[1])))
want = int(sys.argv[2]) if len(sys.argv) > 2 else 0
kernels, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name": cur = {"name": r[1], "hdr": None, "data": []}; kernels.append(cur); continue
    if cur is None: continue
    if r and r[0] == "Address": cur["hdr"] = r; continue
    if cur["hdr"] and len(r) == len(cur["hdr"]) and r[0].startswith("0x"): cur["data"].append(r)
k = kernels[want]
hdr = k["hdr"]; ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = collections.Counter(); byop = collections.defaultdict(collections.Counter); execs = collections.Counter(); n = 0
for r in k["data"]:
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+)", r[ix["Source"]].strip())
    op = m.group(2) if m else "?"
    ie = int(r[ix["Instructions Executed"]] or 0); execs[op] += ie; n += ie
    for s in stalls:
        v = int(r[ix[s]] or 0); tot[s] += v; byop[op][s] += v
allsum = sum(tot.values())
print(k["name"]); print("static instrs", len(k["data"]), "executed warp instrs", n)
print("stall samples:", ", ".join(f"{a[6:]}={b} ({100*b/allsum:.1f}%)" for a, b in tot.most_common(12)))
for op, v in execs.most_common(34):
    s = sum(byop[op].values())
    top = ", ".join(f"{a[6:]}:{c}" for a, c in byop[op].most_common(3))
    print(f"  {op:10s} {v:9d} {100*v/n:5.1f}%  samples {s:6d} ({100*s/allsum:4.1f}%)  {top}")
