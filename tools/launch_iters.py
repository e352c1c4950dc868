#!/usr/bin/env python3
"""Per-iteration view of an ncu launch list (ncu --metrics gpu__time_duration.sum --csv of `bench.py --profile`, single
stream: ASLR_SOLVE_LANES=1): kernel times of every solver iteration of the second (warm) solve and the totals per kernel.
usage: tools/launch_iters.py launches.csv [every]"""
import csv
import re
import sys


def main():
    path = sys.argv[1]
    every = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rows = []
    for r in csv.DictReader(lines):
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void aslr::", "").replace("<unnamed>::", "")
        rows.append((name[:14], float(r["Metric Value"]) / 1e3, r["Grid Size"]))
    starts = [i for i, r in enumerate(rows) if r[0].startswith("init_kernel")]
    sol = rows[starts[1] + 1:] if len(starts) > 1 else rows
    nxt = [i for i, r in enumerate(sol) if r[0].startswith("init_kernel")]
    if nxt:
        sol = sol[:nxt[0]]
    its, cur = [], []
    for r in sol:
        if r[0].startswith("solver_calcdif") and cur:
            its.append(cur)
            cur = []
        if r[0].startswith(("solver_calcdif", "backward", "rollout", "linesearch", "trial_cost", "accept")):
            cur.append(r)
    if cur:
        its.append(cur)
    tot = {}
    for k, it in enumerate(its):
        if k < 12 or k % every == 0 or k > len(its) - 3:
            print(k, " ".join("%s:%.0f" % (n[:9], v) for n, v, g in it), "| sum %.0f us" % sum(v for n, v, g in it))
        for n, v, g in it:
            tot[n] = tot.get(n, 0) + v
    print({k: round(v / 1e3, 2) for k, v in tot.items()}, "total ms %.2f" % (sum(tot.values()) / 1e3))


if __name__ == "__main__":
    main()
