#!/usr/bin/env python
"""Joins an `ncu --page source --csv` dump (SASS rows of ONE kernel launch) with the line table of
`nvdisasm -g -c <cubin>` and prints per-source-line totals: warp instructions, thread instructions,
stall samples, L1 tag requests. Usage:

    cuobjdump -xelf all ray_tracer_b200/libb200pt.so            # -> wavefront.sm_100a.cubin
    nvdisasm -g -c wavefront.sm_100a.cubin > wavefront.sass
    ncu -i prof.ncu-rep --page source --csv --kernel-name regex:k_extend --launch-count 1 > src.csv
    python profiles/tools/sass_by_line.py wavefront.sass <mangled kernel substring> src.csv [top] [launch index in the dump]
"""
import csv
import re
import sys
from collections import defaultdict


def line_table(sass_path, kernel):
    """-> list of (file, line) per instruction index of the kernel's .text section."""
    out, on, cur = [], False, ("?", 0)
    for ln in open(sass_path):
        if ln.startswith(".text."):
            on = kernel in ln
            continue
        if not on:
            continue
        if ln.startswith("//-----") or ln.lstrip().startswith(".section"):
            if out:
                break
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            out.append((cur, m.group(2).strip()))
    return out


def main():
    sass, kernel, src = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    table = line_table(sass, kernel)
    rows = list(csv.reader(open(src)))
    # the dump holds one block per profiled launch of the kernel: "Kernel Name" row, header row, SASS rows
    which = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
    hi = starts[which]
    hdr = rows[hi]
    col = {h: i for i, h in enumerate(hdr)}
    data = []
    for r in rows[hi + 1:]:
        if not r or r[0] in ("Kernel Name", "Address"):
            break
        data.append(r)
    if len(data) != len(table):
        print("warning: %d profile rows vs %d sass instructions" % (len(data), len(table)), file=sys.stderr)
    agg = defaultdict(lambda: [0, 0, 0, 0, 0])
    tot = [0, 0, 0, 0, 0]
    for i, r in enumerate(data):
        key = table[i][0] if i < len(table) else ("?", 0)
        vals = [int(float(r[col[c]] or 0)) if c in col else 0      # kernels without local memory have no such column
                for c in ("Instructions Executed", "Thread Instructions Executed", "# Samples", "L1 Tag Requests Global",
                          "L2 Theoretical Sectors Local")]
        for k, v in enumerate(vals):
            agg[key][k] += v
            tot[k] += v
    print("total warp-inst %d thread-inst %d (%.2f threads/inst) samples %d l1tag %d local-sectors %d" %
          (tot[0], tot[1], tot[1] / max(tot[0], 1), tot[2], tot[3], tot[4]))
    print("%-26s %7s %7s %6s %7s %7s" % ("file:line", "winst%", "smpl%", "thr", "l1tag%", "lsect%"))
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1][2])[:top]:
        print("%-26s %7.2f %7.2f %6.1f %7.2f %7.2f" % ("%s:%d" % key, 100 * v[0] / max(tot[0], 1), 100 * v[2] / max(tot[2], 1),
                                                       v[1] / max(v[0], 1), 100 * v[3] / max(tot[3], 1), 100 * v[4] / max(tot[4], 1)))


if __name__ == "__main__":
    main()
