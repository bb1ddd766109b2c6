#!/usr/bin/env python
"""Summarise an `ncu --set full` capture: per kernel launch the duration, DRAM bytes and the pipe /
issue / shared-memory figures DESIGN.md quotes.  Writes a CSV summary and profiles/ncu_traffic.json
(DRAM bytes per pixel per kernel, which bench.py turns into `roofline.traffic`).

usage: scripts/ncu_traffic.py REPORT.ncu-rep PIXELS_PER_LAUNCH OUT_PREFIX
"""
import csv
import io
import json
import os
import subprocess
import sys

KEEP = [
    "Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__inst_issued.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "sm__maximum_warps_per_active_cycle_pct",
]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rep, pixels, prefix = sys.argv[1], float(sys.argv[2]), sys.argv[3]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    cols = [hdr.index(k) for k in KEEP if k in hdr]
    with open(prefix + "_summary.csv", "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([hdr[c] for c in cols])
        w.writerow([units[c] for c in cols])
        for r in body:
            w.writerow([r[c] for c in cols])
    out = {}
    ir, iw, it = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
    for r in body:
        name = r[hdr.index("Kernel Name")]
        key = name.split("::")[-1].split("<")[0].split("(")[0]
        if key in out:  # first launch of a kernel = the timed step's own (later ones are the other-stage variants)
            continue
        rd = float(r[ir]) * SCALE[units[ir]]
        wr = float(r[iw]) * SCALE[units[iw]]
        out[key] = {"kernel": name, "dram_bytes_read": rd, "dram_bytes_write": wr, "pixels_per_launch": pixels,
                    "dram_bytes_per_pixel": (rd + wr) / pixels, "duration_" + units[it]: float(r[it]),
                    "capture": os.path.basename(rep)}
    path = os.path.join(os.path.dirname(os.path.abspath(prefix)), "ncu_traffic.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
