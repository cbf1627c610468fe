#!/usr/bin/env python3
"""Print the handful of ncu counters the round report quotes, per captured launch.
usage: ncu_summary.py <file.ncu-rep> [extra_metric_substring ...]
       ncu_summary.py --json <file.ncu-rep> <kernel name substring> > profiles/rNN_ncu_summary.json
The --json form averages the launches whose name contains the substring (duration-weighted for the ratios) into the
figures bench.py's roofline object quotes: issue-slot utilisation, active threads per instruction, DRAM and L2 bytes
per launch."""
import csv
import io
import os
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


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return 0.0


def summary_json(rep, needle):
    import json
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    sel = [r for r in rows[2:] if needle in r[idx["Kernel Name"]]]
    if not sel:
        sys.exit(f"no launch matches {needle!r}")
    def col(name):
        return [num(r[idx[name]]) for r in sel]
    dur = col("gpu__time_duration.sum")
    scale = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}.get(units[idx["gpu__time_duration.sum"]], 1.0)
    dur_us = [d * scale for d in dur]
    wsum = sum(dur_us)
    def wavg(name):
        return sum(v * w for v, w in zip(col(name), dur_us)) / wsum
    def bytes_of(name):
        u = units[idx[name]]
        k = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)
        return [v * k for v in col(name)]
    n = len(sel)
    dram = [a + b for a, b in zip(bytes_of("dram__bytes_read.sum"), bytes_of("dram__bytes_write.sum"))]
    stalls = {h[len(STALL):-len("_per_issue_active.ratio")]: round(wavg(h), 2) for h in idx
              if h.startswith(STALL) and h.endswith("_per_issue_active.ratio")}
    top = dict(sorted(stalls.items(), key=lambda kv: -kv[1])[:6])
    print(json.dumps({
        "source": f"ncu --set full --clock-control none, {n} launches of {needle} ({os.path.basename(rep)}), duration-weighted",
        "launches": n, "avg_launch_us": round(wsum / n, 1),
        "issue_slot_utilisation": round(wavg("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100.0, 4),
        "threads_per_instruction": round(wavg("smsp__thread_inst_executed_per_inst_executed.ratio"), 2),
        "warps_active_pct": round(wavg("sm__warps_active.avg.pct_of_peak_sustained_active"), 1),
        "l1tex_throughput_pct": round(wavg("l1tex__throughput.avg.pct_of_peak_sustained_elapsed"), 1),
        "lts_throughput_pct": round(wavg("lts__throughput.avg.pct_of_peak_sustained_elapsed"), 1),
        "dram_throughput_pct": round(wavg("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"), 1),
        "l1_sector_hit_rate_pct": round(wavg("l1tex__t_sector_hit_rate.pct"), 1),
        "dram_bytes_per_trace_launch": round(sum(dram) / n),
        # bytes between the SMs and the L2: what the L2 returned to the L1s plus what the SMs wrote through
        "lts_bytes_per_trace_launch": round((sum(bytes_of("l1tex__m_xbar2l1tex_read_bytes.sum")) + sum(bytes_of("l1tex__m_l1tex2xbar_write_bytes.sum"))) / n),
        "l1tex_bytes_per_trace_launch": round(32.0 * sum(col("SM_B.TriageCompute.l1tex__t_sectors.sum")) / n) if "SM_B.TriageCompute.l1tex__t_sectors.sum" in idx else None,
        "registers_per_thread": int(col("launch__registers_per_thread")[0]),
        "top_stalls_warps_per_issue": top,
    }, indent=1))


def main():
    if sys.argv[1] == "--json":
        summary_json(sys.argv[2], sys.argv[3])
        return
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
