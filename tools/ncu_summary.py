#!/usr/bin/env python
"""Summarise an .ncu-rep (ncu -i ... --page raw --csv) as a markdown table: one row per captured launch."""
import csv
import subprocess
import sys

COLS = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu %"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma %"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ %"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "warp inst")]


def main(path):
    out = subprocess.check_output(["ncu", "-i", path, "--page", "raw", "--csv"], text=True, stderr=subprocess.DEVNULL)
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print("| kernel | grid | " + " | ".join(n for _, n in COLS) + " |")
    print("|---|---|" + "---|" * len(COLS))
    for r in rows[2:]:
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")
        cells = []
        for key, _ in COLS:
            if key in idx:
                v = r[idx[key]]
                u = units[idx[key]]
                try:
                    f = float(v.replace(",", ""))
                    v = f"{f:.1f}" if f < 1e4 else f"{f:.3g}"
                except ValueError:
                    pass
                cells.append(f"{v} {u}".strip() if u in ("us", "ms", "Mbyte", "Kbyte", "Gbyte", "byte") else v)
            else:
                cells.append("-")
        print(f"| `{name}` | {r[idx['Grid Size']]} | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
