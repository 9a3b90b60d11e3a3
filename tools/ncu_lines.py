#!/usr/bin/env python
"""Per-source-line view of an ncu report captured with --import-source on:
share of warp instructions, average active lanes and stall samples per line.

  ncu -i k.ncu-rep --page source --csv --print-source cuda,sass > /tmp/src.csv
  python tools/ncu_lines.py /tmp/src.csv [top_n]
"""
import collections
import csv
import sys


def num(x):
    try:
        return float(x)
    except ValueError:
        return 0.0


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    agg = collections.defaultdict(lambda: [0.0, 0.0, 0.0, 0.0])
    cur, hdr = None, None
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if r and r[0] == "Line No":
            hdr = {n: i for i, n in enumerate(r)}
            continue
        if hdr is None or len(r) < 10 or not r[0].isdigit():
            continue
        a = agg[(cur, int(r[0]), r[1].strip()[:100])]
        a[0] += num(r[hdr["Instructions Executed"]])
        a[1] += num(r[hdr["Thread Instructions Executed"]])
        a[2] += num(r[hdr["# Samples"]])
        a[3] += num(r[hdr["stall_long_sb"]])
    tot = sum(a[0] for a in agg.values()) or 1.0
    tots = sum(a[2] for a in agg.values()) or 1.0
    print("warp instructions %.0f, stall samples %.0f" % (tot, tots))
    byfile = collections.defaultdict(lambda: [0.0, 0.0, 0.0])
    for (f, _, _), a in agg.items():
        for k in range(3):
            byfile[f][k] += a[k]
    for f, a in sorted(byfile.items(), key=lambda x: -x[1][0]):
        print("%-22s inst %5.1f%%  lanes %5.1f  samples %5.1f%%" % (f, 100 * a[0] / tot, a[1] / max(a[0], 1), 100 * a[2] / tots))
    print()
    for (f, ln, src), a in sorted(agg.items(), key=lambda x: -x[1][2])[:top]:
        print("%-18s %4d inst %5.1f%% lanes %5.1f smp %5.1f%% long_sb %4.1f%% | %s"
              % (f, ln, 100 * a[0] / tot, a[1] / max(a[0], 1), 100 * a[2] / tots, 100 * a[3] / tots, src))


if __name__ == "__main__":
    main()
