#!/usr/bin/env python3
"""Condense ncu output into the small text files committed under profiles/.

  python tools/ncu_summary.py launches gpurun_out/launches.csv            > profiles/rNN_launches.md
  python tools/ncu_summary.py full gpurun_out/prof.ncu-rep [more.ncu-rep] > profiles/rNN_ncu_full.md
  python tools/ncu_summary.py figures a.ncu-rep [b.ncu-rep] "what was captured"   -> profiles/ncu_figures.json

`figures` writes the ncu-derived numbers bench.py quotes (DRAM bytes per launch, FP64 pipe utilisation, registers,
spill sectors per kernel) together with the content hash of the kernel sources they were captured from; bench.py
refuses to print them when the sources have changed since (run it right after the capture, before editing).

`launches` reads the CSV of `ncu --metrics gpu__time_duration.sum --csv`; `full` reads `.ncu-rep`
files of `ncu --set full` through `ncu -i ... --page raw --csv` (works without a GPU).
"""
import collections
import csv
import io
import subprocess
import sys

FULL_METRICS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__occupancy_limit_registers", "occupancy limit (regs), blocks/SM"),
    ("launch__occupancy_limit_shared_mem", "occupancy limit (smem), blocks/SM"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active % of peak"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("smsp__warps_eligible.avg.per_cycle_active", "eligible warps / cycle / SMSP"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "active lanes / instruction"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "FP64 pipe % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", "local-memory load sectors (spills)"),
    ("l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum", "local-memory store sectors (spills)"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
]
STALL_PREFIX = "smsp__average_warps_issue_stalled_"
STALL_SUFFIX = "_per_issue_active.ratio"


def short_name(full):
    return full.replace("void ", "").split("(")[0]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui, gi, bi = (hdr.index(k) for k in ("Kernel Name", "Metric Value", "Metric Unit", "Grid Size", "Block Size"))
    agg = collections.OrderedDict()
    seq = []
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[ui], 1.0)
        name = short_name(r[ki])
        a = agg.setdefault(name, [0, 0.0, r[gi], r[bi]])
        a[0] += 1
        a[1] += v
        seq.append((name, v))
    tot = sum(a[1] for a in agg.values())
    print("| kernel | launches | total us | share | grid | block |")
    print("|---|---|---|---|---|---|")
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("| `%s` | %d | %.1f | %.1f %% | %s | %s |" % (name, a[0], a[1], 100 * a[1] / tot, a[2], a[3]))
    print("\ntotal %.1f us over %d launches (cold-cache, serialised by ncu: compare shares, not absolutes)\n" % (tot, len(seq)))
    print("Per-launch durations in launch order (us), every 5th launch of each kernel:\n")
    for name in agg:
        vals = [v for n, v in seq if n == name]
        if len(vals) >= 10:
            print("* `%s`: %s" % (name, " ".join("%.0f" % v for v in vals[::5])))


def full(paths):
    for path in paths:
        txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        rows = [r for r in rows if len(r) > 20]
        hdr, units, data = rows[0], rows[1], rows[2:]
        ki = hdr.index("Kernel Name")
        print("## %s\n" % path.split("/")[-1])
        names = ["`%s`" % short_name(r[ki]) for r in data]
        print("| metric | " + " | ".join(names) + " |")
        print("|---|" + "---|" * len(names))
        for key, label in FULL_METRICS:
            if key not in hdr:
                continue
            i = hdr.index(key)
            u = units[i]
            print("| %s (%s) | " % (label, u or "-") + " | ".join(_fmt(r[i]) for r in data) + " |")
        print("\nWarp stall reasons (average warps stalled per issued instruction):\n")
        print("| reason | " + " | ".join(names) + " |")
        print("|---|" + "---|" * len(names))
        for i, h in enumerate(hdr):
            if h.startswith(STALL_PREFIX) and h.endswith(STALL_SUFFIX):
                vals = [float(r[i].replace(",", "") or 0) for r in data]
                if max(vals) >= 0.05:
                    print("| %s | " % h[len(STALL_PREFIX):-len(STALL_SUFFIX)] + " | ".join("%.2f" % v for v in vals) + " |")
        print()


def _unit_scale(u):
    return {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)


def figures(paths, what):
    import json
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import __graft_entry__ as ge
    kernels = {}
    for path in paths:
        _figures_of(path, kernels)
    out = {"source_hash": ge._source_hash(), "capture": what, "reports": [os.path.basename(q) for q in paths],
           "kernels": kernels}
    dst = os.path.join(root, "profiles", "ncu_figures.json")
    with open(dst, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", dst, "for source hash", out["source_hash"][:12])


def _figures_of(path, kernels):
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(io.StringIO(txt)) if len(r) > 20]
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")

    def val(r, key):
        if key not in hdr:
            return None
        i = hdr.index(key)
        try:
            return float(r[i].replace(",", "")) * _unit_scale(units[i])
        except ValueError:
            return None
    for r in data:
        name = short_name(r[ki]).replace("aslr::", "")
        base = name.split("<")[0]
        entry = {"instance": name, "grid": r[hdr.index("Grid Size")] if "Grid Size" in hdr else None,
                 "duration_us": val(r, "gpu__time_duration.sum"),
                 "dram_bytes_per_launch": (val(r, "dram__bytes_read.sum") or 0.0) + (val(r, "dram__bytes_write.sum") or 0.0),
                 "fp64_pipe_pct": val(r, "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
                 "issue_slots_busy_pct": val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                 "warps_active_pct": val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
                 "registers": val(r, "launch__registers_per_thread"),
                 "warp_instructions": val(r, "smsp__inst_executed.sum"),
                 "spill_sectors": (val(r, "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum") or 0.0) +
                                  (val(r, "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum") or 0.0)}
        kernels.setdefault(base, entry)  # the first launch of each kernel in the capture


def _fmt(s):
    try:
        v = float(s.replace(",", ""))
    except ValueError:
        return s
    if v == int(v) and abs(v) < 1e15:
        return "%d" % v
    return "%.3f" % v


if __name__ == "__main__":
    if len(sys.argv) < 3 or sys.argv[1] not in ("launches", "full", "figures"):
        raise SystemExit(__doc__)
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    elif sys.argv[1] == "figures":
        figures([q for q in sys.argv[2:] if q.endswith(".ncu-rep")], " ".join(q for q in sys.argv[2:] if not q.endswith(".ncu-rep")))
    else:
        full(sys.argv[2:])
