#!/usr/bin/env python
"""ncu_summary.py REPORT.ncu-rep -- the handful of counters we judge a kernel by (reads the report with `ncu -i`)."""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum ", "dram__bytes_write.sum ", "gpu__dram_throughput.avg.pct",
        "lts__throughput.avg.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct",
        "sm__warps_active.avg.pct", "launch__registers_per_thread ", "launch__occupancy_limit",
        "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum ", "smsp__issue_active.avg.pct",
        "smsp__average_warps_issue_stalled", "l1tex__t_sector_hit_rate", "lts__t_sector_hit_rate",
        "l1tex__data_bank_conflicts_pipe_lsu.sum ", "smsp__average_warp_latency_per_inst_issued",
        "lts__t_sectors_op_red.sum ", "lts__t_sectors_op_atom.sum ", "lts__t_sectors_srcunit_tex_op_read.sum ",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ", "launch__shared_mem_per_block_dynamic"]


def main():
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for vals in rows[2:]:
        name = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        print("==", name[:110])
        for h, u, v in zip(hdr, units, vals):
            hh = h + " "
            if any(w in hh for w in WANT):
                try:
                    if float(v) == 0.0 and "stalled" in h:
                        continue
                except ValueError:
                    pass
                print(f"  {h:90s} {v} {u}")


if __name__ == "__main__":
    main()
