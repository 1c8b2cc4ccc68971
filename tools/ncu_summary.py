#!/usr/bin/env python
"""Text summary of an .ncu-rep (key raw metrics + hottest source lines) for profiles/.
Usage: python tools/ncu_summary.py report.ncu-rep out.txt"""
import csv, subprocess, sys
from collections import defaultdict

KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__sass_average_branch_targets_threads_uniform.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_bytes.sum.per_second", "lts__t_bytes.sum.per_second",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "sm__sass_thread_inst_executed_op_fadd_pred_on.sum", "sm__sass_thread_inst_executed_op_fmul_pred_on.sum",
        "sm__sass_thread_inst_executed_op_ffma_pred_on.sum", "smsp__sass_thread_inst_executed_op_fp32_pred_on.sum",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]

def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    lines = []
    for krow in rows[2:]:
        lines.append(f"== kernel: {krow[hdr.index('Kernel Name')] if 'Kernel Name' in hdr else ''}")
        for i, h in enumerate(hdr):
            if h in KEYS or h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"):
                lines.append(f"{h:90s} {krow[i]:>18s} {units[i]}")
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
    cur, agg = None, []
    for r in csv.reader(src.splitlines()):
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif len(r) > 10 and r[0].isdigit():
            try:
                agg.append((cur, int(r[0]), r[1].strip()[:100], int(r[7]), int(r[8]), int(r[6])))
            except ValueError:
                pass
    ti, tt, ts = sum(a[3] for a in agg) or 1, sum(a[4] for a in agg), sum(a[5] for a in agg) or 1
    lines.append(f"\n== source correlation: warp instructions {ti}, thread instructions {tt}, avg active threads {tt/ti:.2f}")
    byf = defaultdict(lambda: [0, 0, 0])
    for a in agg:
        byf[a[0]][0] += a[3]; byf[a[0]][1] += a[4]; byf[a[0]][2] += a[5]
    for k, v in sorted(byf.items(), key=lambda kv: -kv[1][0]):
        lines.append(f"{k:28s} inst {100*v[0]/ti:5.1f}%  avg threads {v[1]/max(1,v[0]):5.1f}  stall samples {100*v[2]/ts:5.1f}%")
    lines.append("\n== hottest source lines (share of warp instructions, avg active threads, share of stall samples)")
    for a in sorted(agg, key=lambda a: -a[3])[:40]:
        lines.append(f"{a[0]:16s}:{a[1]:<4d} {100*a[3]/ti:5.2f}%  thr {a[4]/max(1,a[3]):4.1f}  smp {100*a[5]/ts:5.2f}% | {a[2]}")
    open(out, "w").write("\n".join(lines) + "\n")

if __name__ == "__main__":
    main()
