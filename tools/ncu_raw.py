#!/usr/bin/env python
"""Key raw metrics of an ncu report.  usage: tools/ncu_raw.py report.ncu-rep"""
import csv, io, subprocess, sys
txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct",
        "lts__t_sector_hit_rate.pct", "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
        "sm__inst_executed.sum", "smsp__issue_active.avg.pct", "sm__warps_active.avg.pct", "launch__registers_per_thread",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active", "smsp__average_warps_issue_stalled_barrier_per",
        "smsp__average_warps_issue_stalled_short_scoreboard_per", "smsp__average_warps_issue_stalled_lg_throttle_per",
        "smsp__average_warps_issue_stalled_mio_throttle_per", "lts__t_sectors_srcunit_tex_lookup_hit.sum", "lts__t_sectors_srcunit_tex_lookup_miss.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "smsp__inst_executed.sum "]
for k in range(2, len(rows)):
    print("== launch", k - 2, rows[k][rows[0].index("Kernel Name")] if "Kernel Name" in rows[0] else "")
    for h, u, v in zip(rows[0], rows[1], rows[k]):
        if any(h.startswith(w) for w in want) and "per_second" not in h and "pct_of_peak" not in h.replace("avg.pct_of_peak", ""):
            print("  %-90s %-12s %s" % (h, u, v))
