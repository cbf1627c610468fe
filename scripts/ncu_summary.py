#!/usr/bin/env python3
"""Print the handful of ncu counters the round report quotes, per captured launch.
usage: ncu_summary.py <file.ncu-rep> [extra_metric_substring ...]"""
import csv
import io
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__inst_executed.avg.per_cycle_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.sum",
    "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum",
]
STALL = "smsp__average_warps_issue_stalled_"


def main():
    rep = sys.argv[1]
    extra = sys.argv[2:]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    if len(rows) < 3:
        print("no rows"); return
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        print("----", r[idx["Kernel Name"]][:70], "id", r[idx["ID"]])
        for w in WANT:
            if w in idx:
                print(f"  {w:74s} {r[idx[w]]:>18s} {units[idx[w]]}")
        stalls = [(float(r[i].replace(",", "") or 0), h) for h, i in idx.items() if h.startswith(STALL) and h.endswith("_per_issue_active.ratio")]
        for v, h in sorted(stalls, reverse=True)[:8]:
            print(f"  stall {h[len(STALL):-len('_per_issue_active.ratio')]:40s} {v:10.3f} warps/issue")
        for e in extra:
            for h, i in idx.items():
                if e in h and h not in WANT:
                    print(f"  {h:74s} {r[i]:>18s} {units[i]}")


if __name__ == "__main__":
    main()
