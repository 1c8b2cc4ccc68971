#!/usr/bin/env python
"""Text summary of an .ncu-rep (key raw metrics + hottest source lines) for profiles/.
Usage: python tools/ncu_summary.py report.ncu-rep out.txt [--commit HASH] [--samples N] [--json out.json --key w5]
--samples: camera paths per profiled launch -> executed FP32 flops per sample; --json merges {key: {...}} into a JSON file
(profiles/k_render_ncu.json, read by bench.py for roofline.executed and roofline.traffic)."""
import csv, json, os, subprocess, sys
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
        "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum", "smsp__sass_thread_inst_executed_op_fp32_pred_on.sum",
        "smsp__thread_inst_executed.sum",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]

FP32_PEAK_TFLOPS = 71.7      # measured FMA-chain throughput of a B200 on this pool (pt_measure_fp32_peak; nominal 74.4)


def main():
    rep, out = sys.argv[1], sys.argv[2]
    opt = dict(zip(sys.argv[3::2], sys.argv[4::2]))
    commit, n_samples = opt.get("--commit", "unknown"), float(opt.get("--samples", 0))
    record = {}
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    lines = [f"# profiled build: commit {commit}; report {os.path.basename(rep)}"]
    for krow in rows[2:3]:          # the first captured launch (the others are replays of the same frame)
        lines.append(f"== kernel: {krow[hdr.index('Kernel Name')] if 'Kernel Name' in hdr else ''}")
        for i, h in enumerate(hdr):
            if h in KEYS or h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"):
                lines.append(f"{h:90s} {krow[i]:>18s} {units[i]}")
        # derived: achieved bandwidths of the memory levels (sectors are 32 bytes) and where the L1 traffic comes from
        def val(name):
            for i, h in enumerate(hdr):                      # some columns carry a section prefix ("SM_B.TriageCompute.<metric>")
                if h == name or h.endswith("." + name):
                    try:
                        return float(krow[i].replace(",", ""))
                    except ValueError:
                        pass
            return None
        dur = val("gpu__time_duration.sum")
        if dur:
            dur_s = dur * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(units[hdr.index("gpu__time_duration.sum")], 1e-9)
            l1, l2 = val("l1tex__t_sectors.sum"), val("lts__t_sectors.sum")
            g_ld = val("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum")
            l_ld, l_st = val("l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum"), val("l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum")
            dr, dw = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
            lines.append("-- derived (32-byte sectors / kernel duration)")
            if l1: lines.append(f"L1 (l1tex) achieved                {l1 * 32 / dur_s / 1e9:10.1f} GB/s   ({l1 * 32 / 1e9:.2f} GB per launch)")
            if g_ld: lines.append(f"  global loads (BVH nodes, triangle records, materials)  {g_ld * 32 / dur_s / 1e9:10.1f} GB/s, all but {100 - (val('l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct') or 0):.2f} % served by L1")
            if l_ld is not None and l_st is not None: lines.append(f"  local loads / stores (traversal stack, per-lane sums, spills)  {l_ld * 32 / dur_s / 1e9:10.1f} / {l_st * 32 / dur_s / 1e9:.1f} GB/s (stores write through to L2)")
            if l2: lines.append(f"L2 (lts) achieved                  {l2 * 32 / dur_s / 1e9:10.1f} GB/s")
            if dr is not None and dw is not None:
                scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
                b = dr * scale.get(units[hdr.index("dram__bytes_read.sum")], 1.0) + dw * scale.get(units[hdr.index("dram__bytes_write.sum")], 1.0)
                lines.append(f"HBM (dram) achieved                {b / dur_s / 1e9:10.1f} GB/s   ({b / 1e6:.1f} MB per launch)")
            # executed FP32 work (report BOTH rooflines): thread-level FADD + FMUL + 2 x FFMA, predicated-on
            fa, fm, ff = (val("smsp__sass_thread_inst_executed_op_fadd_pred_on.sum"), val("smsp__sass_thread_inst_executed_op_fmul_pred_on.sum"),
                          val("smsp__sass_thread_inst_executed_op_ffma_pred_on.sum"))
            if fa is not None and fm is not None and ff is not None:
                flop = fa + fm + 2 * ff
                tf = flop / dur_s / 1e12
                lines.append(f"executed FP32 (FADD {fa:.3e} + FMUL {fm:.3e} + 2 x FFMA {ff:.3e}) = {flop:.4e} flop per launch = {tf:8.2f} TFLOP/s"
                             f" = {tf / FP32_PEAK_TFLOPS:.3f} of the measured FP32 peak ({FP32_PEAK_TFLOPS} TFLOP/s)"
                             + (f"; {flop / n_samples:.1f} executed flop per sample" if n_samples else ""))
                ti = val("smsp__thread_inst_executed.sum") or val("smsp__sass_thread_inst_executed.sum")
                record.update(executed_flop_per_launch=flop, executed_tflops_under_ncu=tf, fadd=fa, fmul=fm, ffma=ff)
                if n_samples:
                    record.update(samples_per_launch=n_samples, executed_flop_per_sample=flop / n_samples)
                if ti:
                    lines.append(f"FP32 share of thread instructions   {100 * (fa + fm + ff) / ti:10.1f} %")
            if dr is not None and dw is not None:
                record.update(dram_bytes_per_launch=b, dram_read=dr * scale.get(units[hdr.index("dram__bytes_read.sum")], 1.0),
                              dram_write=dw * scale.get(units[hdr.index("dram__bytes_write.sum")], 1.0))
            wi = val("smsp__inst_executed.sum")
            if wi and n_samples:
                lines.append(f"warp instructions per sample        {wi / n_samples:10.1f}")
                record.update(warp_inst_per_sample=wi / n_samples)
            record.update(kernel=krow[hdr.index('Kernel Name')] if 'Kernel Name' in hdr else '', kernel_ms_under_ncu=dur_s * 1e3, commit=commit,
                          threads_per_inst=val("smsp__thread_inst_executed_per_inst_executed.ratio"),
                          issue_active_pct=val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                          registers=val("launch__registers_per_thread"))
            wee = val("smsp__thread_inst_executed_per_inst_executed.ratio")
            if wee: lines.append(f"warp execution efficiency          {100 * wee / 32:10.1f} %   (active threads per instruction / 32)")
            be = val("smsp__sass_average_branch_targets_threads_uniform.pct")
            if be: lines.append(f"branch efficiency                  {be:10.1f} %   (branch targets uniform across the warp)")
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
    if "--json" in opt and record:
        path = opt["--json"]
        allr = json.load(open(path)) if os.path.exists(path) else {}
        allr[opt.get("--key", "k")] = record
        json.dump(allr, open(path, "w"), indent=1, sort_keys=True)

if __name__ == "__main__":
    main()
