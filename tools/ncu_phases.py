#!/usr/bin/env python
"""Per-phase breakdown of one kernel from `ncu -i X.ncu-rep --page source --csv --print-source sass`.

A phase = the SASS between two BAR.SYNC instructions in program order (the kernels here are straight-line
phase sequences separated by __syncthreads()).  For every phase: thread instructions (optionally per pixel),
share of stall samples, the top stall reasons and the top opcodes.

usage: ncu_phases.py sass.csv [units]      units = divisor for the per-unit column (e.g. output pixels of the launch)
"""
import csv
import sys
from collections import Counter


def main(path, units=None):
    rows = list(csv.reader(open(path)))
    k = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[k]
    ix = {h: i for i, h in enumerate(hdr)}
    stall_cols = [h for h in hdr if h.startswith("stall_")]
    phases = [dict(inst=0, samples=0, ops=Counter(), stalls=Counter())]
    for r in rows[k + 1:]:
        if r and r[0] == "Address":
            break                                    # the page repeats the kernel: one copy is enough
        if len(r) < len(hdr):
            continue
        src = r[ix["Source"]].strip()
        toks = src.split()
        op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
        op = op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("VIMNMX", "LDS", "STS", "LDG", "STG", "IMAD")) and "." in op else "")
        ph = phases[-1]
        ti = int(r[ix["Thread Instructions Executed"]] or 0)
        ph["inst"] += ti
        ph["ops"][op] += ti
        ph["samples"] += int(r[ix["# Samples"]] or 0)
        for c in stall_cols:
            v = r[ix[c]]
            if v and v != "0":
                ph["stalls"][c[6:]] += int(v)
        if src.startswith("BAR.SYNC") or " BAR.SYNC" in src:
            phases.append(dict(inst=0, samples=0, ops=Counter(), stalls=Counter()))
    tot_i = sum(p["inst"] for p in phases) or 1
    tot_s = sum(p["samples"] for p in phases) or 1
    u = float(units) if units else None
    print(f"total thread instructions {tot_i:.4g}" + (f" = {tot_i / u:.1f} per unit" if u else "") + f", samples {tot_s}")
    for n, p in enumerate(phases):
        line = f"phase {n}: inst {100.0 * p['inst'] / tot_i:5.1f}%"
        if u:
            line += f" ({p['inst'] / u:6.1f}/unit)"
        line += f"  samples {100.0 * p['samples'] / tot_s:5.1f}%"
        st = ", ".join(f"{k} {100.0 * v / tot_s:.1f}" for k, v in p["stalls"].most_common(4))
        ops = ", ".join(f"{k} {v / u:.1f}" if u else f"{k} {100.0 * v / tot_i:.1f}%" for k, v in p["ops"].most_common(8))
        print(line + f"\n    stalls(% of all samples): {st}\n    ops: {ops}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None)
