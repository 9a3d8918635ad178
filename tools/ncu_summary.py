"""One line per kernel of an `ncu --page raw --csv` export: the metrics profiles/README.md quotes."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, vals = rows[0], rows[2:]
want = {
    "ms": "gpu__time_duration.sum", "inst": "smsp__inst_executed.sum", "issue%": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "fma%": "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "xu%": "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "alu%": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "lsu%": "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "warps%": "sm__warps_active.avg.pct_of_peak_sustained_active", "regs": "launch__registers_per_thread",
    "dram_rd": "dram__bytes_read.sum", "dram_wr": "dram__bytes_write.sum", "dram%": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l2%": "lts__t_sectors.avg.pct_of_peak_sustained_elapsed", "l2_hit%": "lts__t_sector_hit_rate.pct", "l1_hit%": "l1tex__t_sector_hit_rate.pct",
    "long_sb": "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "math_thr": "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "mio_thr": "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "not_sel": "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "wait": "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "branch": "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
}
units = rows[1]
ki = hdr.index("Kernel Name")
for v in vals:
    name = v[ki].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
    out = []
    for k, m in want.items():
        if m in hdr:
            i = hdr.index(m)
            out.append(f"{k}={v[i]}{units[i] if k in ('ms', 'dram_rd', 'dram_wr') else ''}")
    print(name, "|", " ".join(out))
