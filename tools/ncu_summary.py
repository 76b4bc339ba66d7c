#!/usr/bin/env python
"""Reduces `ncu -i X.ncu-rep --page raw --csv` to the metrics the roofline
entries quote (one block per profiled launch).  usage: ncu_summary.py raw.csv [points]"""
import csv
import sys

KEYS = [
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "launch__shared_mem_per_block_static", "gpu__time_duration.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__t_sectors.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
    "l1tex__data_pipe_lsu_wavefronts.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed",
    "l1tex__f_wavefronts.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__cycles_active.avg", "sm__cycles_elapsed.avg.per_second",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    points = float(sys.argv[2]) if len(sys.argv) > 2 else None
    header = rows[0]
    units = rows[1]
    for r in rows[2:]:
        d = dict(zip(header, r))
        u = dict(zip(header, units))
        print("Kernel Name =", d.get("Kernel Name"))
        for k in KEYS:
            if k in d:
                print(f"{k} = {d[k]} {u.get(k, '')}")
        try:
            def val(k):
                v = float(d[k].replace(",", ""))
                unit = u.get(k, "")
                scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}.get(unit, 1.0)
                return v * scale
            traffic = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
            line = f"traffic (dram read + write) = {traffic / 1e9:.4f} GB"
            if points:
                line += f" = {traffic / points:.3f} B/point; warp instructions / point = {val('smsp__inst_executed.sum') / points:.2f}"
            print(line)
        except Exception as e:  # noqa: BLE001
            print("traffic: n/a", e)
        print("---")


if __name__ == "__main__":
    main()
