#!/usr/bin/env python3
"""Where does a kernel spend its issue slots, by SOURCE LINE (inlined device functions included).

  python tools/ncu_lines.py REPORT.ncu-rep KERNEL_REGEX OBJECT.o [top] [MANGLED_REGEX]

KERNEL_REGEX selects the kernel in the report (demangled name); MANGLED_REGEX (default: the same) selects
the function in the object, e.g. backward_cta_kernelILi28ELi7E.

`ncu --page source --csv` gives per-SASS-instruction samples; `nvdisasm -g -c` of the object the library
was linked from gives the file:line of every SASS instruction (needs -lineinfo).  Both come from the same
build, so the instruction offsets line up.  Output: top source lines by warp-stall samples, with the
dominant stall reason, and totals per file.
"""
import collections
import csv
import io
import re
import os
import subprocess
import sys


def sass_lines(obj, kernel_regex):
    """offset(hex str, lower, no leading zeros) -> 'file:line' for the first matching function."""
    if obj.endswith(".o"):  # host object with an embedded fatbin: pull the cubin out first
        import glob
        import tempfile
        tmp = tempfile.mkdtemp()
        subprocess.run(["cuobjdump", "-xelf", "all", obj], cwd=tmp, capture_output=True)
        obj = glob.glob(tmp + "/*.cubin")[0]
    txt = subprocess.run(["nvdisasm", "-g", "-c", obj], capture_output=True, text=True).stdout
    out, cur, active = {}, None, False
    fn_re = re.compile(r"^\s*//-+ \.text\.(\S+)")
    loc_re = re.compile(r'//## File "([^"]+)", line (\d+)')
    ins_re = re.compile(r"^\s*/\*([0-9a-f]{4,})\*/")
    done = False
    for ln in txt.splitlines():
        m = fn_re.match(ln)
        if m:
            if active:
                done = True
            active = (not done) and re.search(kernel_regex, m.group(1)) is not None
            cur = None
            continue
        if not active:
            continue
        m = loc_re.search(ln)
        if m:
            cur = "%s:%s" % (m.group(1).split("/")[-1], m.group(2))
            continue
        m = ins_re.match(ln)
        if m and cur:
            out[int(m.group(1), 16)] = cur
    return out


def main():
    rep, kre, obj = sys.argv[1], sys.argv[2], os.path.abspath(sys.argv[3])
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    ai, si, ei = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    lines = sass_lines(obj, sys.argv[5] if len(sys.argv) > 5 else kre)
    data = []
    for r in rows[hdr_i + 1:]:
        if len(r) <= ei or r[0] == "Address":
            break  # first kernel instance only
        try:
            addr, samples, execd = int(r[ai], 16), int(r[si]), int(r[ei])
        except ValueError:
            continue
        data.append((addr, samples, execd, r))
    if not data:
        raise SystemExit("no instructions found for %s" % kre)
    base = min(d[0] for d in data)
    by_line = collections.defaultdict(lambda: [0, 0, collections.Counter()])
    by_file = collections.Counter()
    tot_s = tot_e = 0
    for addr, samples, execd, r in data:
        loc = lines.get(addr - base, "?")
        a = by_line[loc]
        a[0] += samples
        a[1] += execd
        for i, h in stall_cols:
            try:
                a[2][h] += int(r[i])
            except ValueError:
                pass
        by_file[loc.split(":")[0]] += samples
        tot_s += samples
        tot_e += execd
    print("kernel %s: %d samples, %d warp instructions, %d SASS instructions, %d with line info" %
          (kre, tot_s, tot_e, len(data), sum(1 for d in data if (d[0] - base) in lines)))
    print("\nsamples by file:")
    for f, n in by_file.most_common():
        print("  %-24s %6.1f %%" % (f, 100.0 * n / max(tot_s, 1)))
    print("\n%-28s %8s %7s %10s  %s" % ("source line", "samples", "%", "warp-inst", "top stall reasons"))
    for loc, (s, e, st) in sorted(by_line.items(), key=lambda kv: -kv[1][0])[:top]:
        why = ", ".join("%s %d" % (k.replace("stall_", ""), v) for k, v in st.most_common(3) if v)
        print("%-28s %8d %6.1f%% %10d  %s" % (loc, s, 100.0 * s / max(tot_s, 1), e, why))


if __name__ == "__main__":
    main()
