"""Condenses an .ncu-rep into the per-kernel metrics DESIGN.md and bench.py quote.
usage: python scripts/ncu_summary.py report.ncu-rep out.csv [kernel-substring ...]
One row per profiled launch: duration, DRAM bytes read+written, L2 bytes, issue-slot utilisation, threads per
instruction, occupancy, registers, local-memory traffic."""
import csv, io, subprocess, sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
    "sm__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__inst_issued.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    filt = sys.argv[3:]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, body = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(head)}
    keep = [m for m in METRICS if m in col]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["id", "kernel"] + [f"{m} [{units[col[m]]}]" for m in keep])
        for r in body:
            name = r[col["Kernel Name"]]
            if filt and not any(s in name for s in filt):
                continue
            w.writerow([r[col["ID"]], name[:90]] + [r[col[m]] for m in keep])
    print(open(out).read())


if __name__ == "__main__":
    main()
