#!/usr/bin/env python3
"""Aggregate `ncu --page source --csv` output over all captured launches of a kernel: stall samples and
executed warp instructions per SASS instruction, printed in program order for the hottest window(s).

  ncu -i REPORT.ncu-rep --page source --csv > src.csv ; python tools/ncu_sass_hot.py src.csv [min_samples]
"""
import collections, csv, sys
rows = csv.reader(open(sys.argv[1]))
thr = int(sys.argv[2]) if len(sys.argv) > 2 else 1
hdr = None
samp = collections.Counter(); execd = collections.Counter(); text = {}; order = []
nk = 0
for r in rows:
    if not r: continue
    if r[0] == "Kernel Name": nk += 1; continue
    if r[0] == "Address": hdr = {n: i for i, n in enumerate(r)}; continue
    if hdr is None or len(r) < len(hdr): continue
    a = r[hdr["Address"]]
    if a not in text: text[a] = r[hdr["Source"]]; order.append(a)
    samp[a] += int(r[hdr["# Samples"]] or 0)
    execd[a] += int(r[hdr["Instructions Executed"]] or 0)
tot = sum(samp.values())
print(f"{nk} launches, {tot} samples, {sum(execd.values())} warp instructions")
cum = 0
for a in order:
    if execd[a] == 0 and samp[a] == 0: continue
    cum += samp[a]
    flag = "#" * min(60, int(600 * samp[a] / max(tot, 1)))
    if samp[a] >= thr or len(sys.argv) > 3:
        print(f"{a[-5:]} {samp[a]:6d} {100*cum/tot:5.1f}% x{execd[a]//max(nk,1):5d} {text[a][:70]:70s} {flag}")
