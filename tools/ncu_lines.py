#!/usr/bin/env python
"""Per-source-line instruction / stall-sample shares from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:K > f.csv`.
usage: ncu_lines.py f.csv [min_percent]"""
import csv
import sys


def main(path, minpct=0.8):
    rows = list(csv.reader(open(path)))
    k = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
    hdr = rows[k]
    iI, iS = hdr.index("Instructions Executed"), hdr.index("# Samples")
    fname = ""
    lines = []
    for r in rows[:k]:
        if r and r[0] == "File Path":
            fname = r[1]
    for r in rows[k + 1:]:
        if r and r[0] == "File Path":
            fname = r[1]
        if len(r) < len(hdr) or not r[0].isdigit():
            continue
        try:
            lines.append((fname.split("/")[-1], int(r[0]), int(r[iI] or 0), int(r[iS] or 0), r[1].strip()))
        except ValueError:
            pass                                     # a source line whose embedded quotes broke the CSV row
    ti = sum(l[2] for l in lines) or 1
    ts = sum(l[3] for l in lines) or 1
    print(f"warp instructions {ti:.4g}, samples {ts}")
    for f, ln, i, s, src in lines:
        if 100.0 * i / ti >= minpct or 100.0 * s / ts >= minpct:
            print(f"{100.0 * i / ti:5.1f}% inst {100.0 * s / ts:5.1f}% smp  {f}:{ln}: {src[:100]}")


if __name__ == "__main__":
    main(sys.argv[1], float(sys.argv[2]) if len(sys.argv) > 2 else 0.8)
