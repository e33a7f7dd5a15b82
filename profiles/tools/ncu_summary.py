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


def traffic_json(rows, hdr, units, report, dest, key):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of every captured kernel -> dest (bench.py reads
    `key` from it for roofline.traffic instead of a constant typed into the source)."""
    import json
    import os
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    ir, iw, it = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
    launches = []
    for v in rows[2:]:
        launches.append({"kernel": v[hdr.index("Kernel Name")][:120],
                         "dram_bytes_read": float(v[ir]) * scale.get(units[ir], 1),
                         "dram_bytes_write": float(v[iw]) * scale.get(units[iw], 1),
                         "duration_us": float(v[it]) * {"ns": 1e-3, "us": 1, "ms": 1e3}.get(units[it], 1)})
    data = {}
    if os.path.exists(dest):
        data = json.load(open(dest))
    main_k = max(launches, key=lambda d: d["dram_bytes_read"] + d["dram_bytes_write"])
    data[key] = {"dram_bytes": main_k["dram_bytes_read"] + main_k["dram_bytes_write"], "kernel": main_k["kernel"],
                 "duration_us_under_ncu": main_k["duration_us"], "report": os.path.basename(report), "launches": launches,
                 "note": "ncu flushes the caches before every replay: cold-cache traffic of one launch"}
    json.dump(data, open(dest, "w"), indent=1)


def main():
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    if "--traffic-json" in sys.argv:  # ncu_summary.py REPORT --traffic-json profiles/r02_traffic.json sym7
        i = sys.argv.index("--traffic-json")
        return traffic_json(rows, hdr, units, sys.argv[1], sys.argv[i + 1], sys.argv[i + 2])
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
