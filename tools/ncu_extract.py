"""Filter the raw page of an ncu report into the (kernel, metric, unit, value) rows kept under profiles/.

    ncu -i x.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_extract.py raw.csv [more.csv ...] > profiles/rNN_prof_*.csv
"""
from __future__ import annotations

import csv
import sys

KEEP = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "dram__bytes.sum.per_second",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "l1tex__t_sector_hit_rate.pct",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "launch__grid_size",
    "launch__block_size",
    "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
    "launch__waves_per_multiprocessor",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
]


def main():
    w = csv.writer(sys.stdout)
    w.writerow(["kernel", "metric", "unit", "value"])
    for path in sys.argv[1:]:
        rows = [r for r in csv.reader(open(path)) if r]
        hdr, units = rows[0], rows[1]
        ik = hdr.index("Kernel Name")
        for r in rows[2:]:
            name = r[ik].replace("bi::", "")[:70]
            for h, u, v in zip(hdr, units, r):
                if h in KEEP:
                    w.writerow([name, h, u, v])


if __name__ == "__main__":
    main()
