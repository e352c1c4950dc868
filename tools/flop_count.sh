#!/bin/bash
# Exact FP64 operation counts of the per-knot arithmetic -> profiles/flop_counts.json (see tools/flop_count.cpp)
set -e
cd "$(dirname "$0")/.."
tmp=$(mktemp -d)
python - "$tmp" <<'PY'
import sys
sys.path.insert(0, ".")
from aslr_to_b200 import workloads as W
d = sys.argv[1]
for tag, w in (("sea2", W.two_dof_sea()), ("vsa2", W.two_dof_vsa_modified()), ("sea7", W.chain7_batch(B=1, T=2)),
               ("vsa7", W.chain7_vsa_batch(B=1, T=2))):
    open("%s/desc_%s.bin" % (d, tag), "wb").write(bytes(W.to_desc(w)))
PY
g++ -O1 -std=c++17 -I include -I aslr_to_b200/csrc -o "$tmp/flop_count" tools/flop_count.cpp
"$tmp/flop_count" "$tmp" > profiles/flop_counts.json
rm -rf "$tmp"
cat profiles/flop_counts.json
