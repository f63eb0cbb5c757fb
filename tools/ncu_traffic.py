#!/usr/bin/env python
"""From an `ncu --set full` report of one bench step: per-launch duration / DRAM bytes CSV and profiles/traffic.json
(DRAM bytes per frame of the dominant kernel group, used by bench.py's roofline.traffic)."""
import csv
import json
import subprocess
import sys


def main(rep, frames, out_csv, out_json):
    out = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True, stderr=subprocess.DEVNULL)
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}

    def val(r, key):
        v = float(r[ix[key]].replace(",", ""))
        u = units[ix[key]]
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1, "ms": 1e3, "ns": 1e-3, "s": 1e6}.get(u, 1)

    tot = {}
    with open(out_csv, "w") as f:
        f.write("kernel,grid,gpu__time_duration_us,dram_read_bytes,dram_write_bytes,alu_pipe_pct,issue_active_pct\n")
        for r in rows[2:]:
            name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "").strip()
            t, rd, wr = val(r, "gpu__time_duration.sum"), val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
            f.write(f"{name},\"{r[ix['Grid Size']]}\",{t:.1f},{rd:.0f},{wr:.0f},"
                    f"{r[ix['sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active']]},"
                    f"{r[ix['smsp__issue_active.avg.pct_of_peak_sustained_active']]}\n")
            k = name.split("<")[0]
            if k == "pyr_level_tma_kernel":               # K1 = the staged-load kernel (level 0) + the persistent TMA kernel (levels > 0)
                k = "pyr_level_kernel"
            a = tot.setdefault(k, [0.0, 0.0, 0.0, 0, 0.0, 0.0])
            a[0] += t; a[1] += rd; a[2] += wr; a[3] += 1
            a[4] += t * float(r[ix['sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active']])      # time-weighted
            a[5] += t * float(r[ix['smsp__issue_active.avg.pct_of_peak_sustained_active']])
    total_t = sum(a[0] for a in tot.values())
    summary = {k: {"launches": a[3], "time_us": a[0], "share": a[0] / total_t, "dram_read_bytes": a[1], "dram_write_bytes": a[2],
                   "alu_pipe_pct": a[4] / a[0], "issue_active_pct": a[5] / a[0]}
               for k, a in tot.items()}
    p = tot["pyr_level_kernel"]
    out = {"source": rep.split("/")[-1], "frames_per_launch": frames, "kernels": summary,
           "pyramid_blur_fast_dram_bytes_per_frame": (p[1] + p[2]) / frames}
    if len(sys.argv) > 5:                                 # a second report: one bf2nn_kernel launch (4096 queries x 1.25 M rows)
        o2 = subprocess.check_output(["ncu", "-i", sys.argv[5], "--page", "raw", "--csv"], text=True, stderr=subprocess.DEVNULL)
        r2 = list(csv.reader(o2.splitlines()))
        hdr, units = r2[0], r2[1]
        ix = {h: i for i, h in enumerate(hdr)}
        r = next(x for x in r2[2:] if "bf2nn_kernel" in x[ix["Kernel Name"]])
        out["bf2nn"] = {"source": sys.argv[5].split("/")[-1], "rows": 1250000, "queries": 4096,
                        "dram_bytes_per_launch": val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum"),
                        "time_us": val(r, "gpu__time_duration.sum"),
                        "alu_pipe_pct": float(r[ix['sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active']]),
                        "issue_active_pct": float(r[ix['smsp__issue_active.avg.pct_of_peak_sustained_active']])}
    json.dump(out, open(out_json, "w"), indent=1)
    for k, v in sorted(summary.items(), key=lambda kv: -kv[1]["time_us"]):
        print(f"{k:20s} x{v['launches']}  {v['time_us']:9.1f} us  {100 * v['share']:5.1f}%  rd {v['dram_read_bytes'] / 1e6:8.1f} MB  wr {v['dram_write_bytes'] / 1e6:8.1f} MB")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]), sys.argv[3], sys.argv[4])
