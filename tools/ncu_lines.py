"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump per CUDA source line.
usage: ncu -i rep --page source --csv --print-source cuda,sass > src.csv ; python tools/ncu_lines.py src.csv [top]"""
import csv
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    cur = None
    agg = defaultdict(lambda: defaultdict(float))
    text = {}
    header = None
    for row in csv.reader(open(path)):
        if not row:
            continue
        if row[0] == "File Path":
            cur = row[1].split("/")[-1]
            continue
        if row[0] == "Line No":
            header = {k: i for i, k in enumerate(row)}
            header["Source"] = 1
            continue
        if row[0] == "Function Name" or header is None:
            continue
        try:
            line = int(row[0])
        except ValueError:
            continue
        if row[2] not in ("", "-"):   # a SASS row below its CUDA line: counted through the CUDA row already
            continue
        key = (cur, line)
        text[key] = row[1].strip()[:100]
        for k in ("# Samples", "Instructions Executed", "Thread Instructions Executed", "stall_no_inst", "stall_long_sb",
                  "stall_short_sb", "stall_barrier", "stall_wait", "stall_mio", "stall_lg", "stall_math", "stall_branch_resolving",
                  "L1 Wavefronts Shared Excessive"):
            i = header.get(k)
            if i is not None and i < len(row) and row[i]:
                try:
                    agg[key][k] += float(row[i])
                except ValueError:
                    pass
    tot = sum(v["# Samples"] for v in agg.values()) or 1.0
    toti = sum(v["Instructions Executed"] for v in agg.values()) or 1.0
    print("total samples %d, warp instructions %d" % (tot, toti))
    names = ("stall_no_inst", "stall_long_sb", "stall_short_sb", "stall_barrier", "stall_wait", "stall_mio", "stall_lg", "stall_math", "stall_branch_resolving")
    print("stall totals: " + ", ".join("%s %.1f%%" % (n[6:], 100 * sum(v[n] for v in agg.values()) / tot) for n in names))
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1]["# Samples"])[:top]:
        thr = v["Thread Instructions Executed"] / v["Instructions Executed"] if v["Instructions Executed"] else 0
        dom = max(names, key=lambda n: v[n])
        print("%5.1f%% smp %5.1f%% inst thr %4.1f %-14s %s:%d  %s" % (100 * v["# Samples"] / tot, 100 * v["Instructions Executed"] / toti,
                                                                  thr, dom[6:], key[0], key[1], text[key]))


if __name__ == "__main__":
    main()


def regions(path, spec):
    """spec: list of (name, file, lo, hi); prints instruction and sample share per region."""
    cur, header = None, None
    acc = defaultdict(lambda: [0.0, 0.0, 0.0])
    for row in csv.reader(open(path)):
        if not row:
            continue
        if row[0] == "File Path":
            cur = row[1].split("/")[-1]
            continue
        if row[0] == "Line No":
            header = {k: i for i, k in enumerate(row)}
            continue
        if header is None or row[0] == "Function Name" or row[2] not in ("", "-"):
            continue
        try:
            line = int(row[0])
        except ValueError:
            continue
        name = "other:" + cur
        for n, f, lo, hi in spec:
            if f == cur and lo <= line <= hi:
                name = n
                break
        try:
            acc[name][0] += float(row[header["# Samples"]] or 0)
            acc[name][1] += float(row[header["Instructions Executed"]] or 0)
            acc[name][2] += float(row[header["Thread Instructions Executed"]] or 0)
        except ValueError:
            pass
    ts = sum(v[0] for v in acc.values()) or 1
    ti = sum(v[1] for v in acc.values()) or 1
    for n, v in sorted(acc.items(), key=lambda kv: -kv[1][0]):
        print("%-28s %5.1f%% samples %5.1f%% warp-inst  %4.1f thr/inst" % (n, 100 * v[0] / ts, 100 * v[1] / ti, v[2] / v[1] if v[1] else 0))
