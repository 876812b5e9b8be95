"""Summarise an .ncu-rep (read here, no GPU): one row per launch with the counters the design notes cite.
Usage: python tools/ncu_summary.py gpurun_out/prof_x.ncu-rep [out.csv]"""
import csv, subprocess, sys, io
KEEP = [
    ("Kernel Name", "kernel"), ("gpu__time_duration.sum", "ms"), ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64%"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu%"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
    ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
    ("lts__t_sector_hit_rate.pct", "l2hit%"), ("l1tex__t_sector_hit_rate.pct", "l1hit%"),
    ("smsp__inst_executed.sum", "inst"),
    ("smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "st_noinst"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "st_long"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "st_short"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "st_wait"),
    ("smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "st_branch"),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "st_math"),
    ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "st_lg"),
    ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "st_notsel"),
    ("smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio", "st_disp"),
    ("sass__inst_executed_local_loads", "lld"), ("sass__inst_executed_local_stores", "lst"),
]
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
cols = [(hdr.index(k), short) for k, short in KEEP if k in hdr]
out = [[short for _, short in cols]]
for r in data:
    line = []
    for i, short in cols:
        v = r[i]
        if short == "kernel":
            v = v.split("(")[0].replace("void ", "")
        else:
            try:
                x = float(v.replace(",", ""))
                if units[i] in ("Mbyte",): v = f"{x:.1f}MB"
                elif units[i] in ("Kbyte",): v = f"{x/1024:.2f}MB"
                elif units[i] in ("Gbyte",): v = f"{x*1024:.0f}MB"
                elif units[i] == "byte": v = f"{x/1e6:.2f}MB"
                elif short == "ms" and units[i] == "us": v = f"{x/1000:.4f}"
                elif short == "ms" and units[i] == "ns": v = f"{x/1e6:.4f}"
                elif short in ("inst", "lld", "lst", "grid", "regs"): v = f"{x:.0f}"
                else: v = f"{x:.2f}"
            except ValueError:
                pass
        line.append(v)
    out.append(line)
w = csv.writer(open(sys.argv[2], "w") if len(sys.argv) > 2 else sys.stdout)
w.writerows(out)
