#!/usr/bin/env python
"""Print selected raw metrics of every kernel in an ncu report:  python tools/ncu_raw.py report.ncu-rep [metric-regex]"""
import csv
import io
import re
import subprocess
import sys

DEFAULT = (r"Kernel Name|gpu__time_duration.sum|launch__(registers_per_thread|grid_size|block_size)|sm__warps_active.avg.pct|"
           r"dram__bytes_(read|write).sum$|smsp__issue_active.avg.pct|l1tex__throughput.avg.pct|lts__throughput.avg.pct|"
           r"gpu__dram_throughput.avg.pct|smsp__inst_executed.sum$|lts__t_sector_hit_rate.pct|l1tex__t_sector_hit_rate.pct|"
           r"smsp__average_warps_issue_stalled_.*_per_issue_active|lts__t_sectors_op_(read|write|atom|red).sum$|"
           r"l1tex__t_requests_pipe_lsu_mem_global_op_(ld|st).sum$|l1tex__t_sectors_pipe_lsu_mem_global_op_(ld|st).sum$")
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
pat = re.compile(sys.argv[2] if len(sys.argv) > 2 else DEFAULT)
for r in rows[2:]:
    print("----")
    for i, h in enumerate(hdr):
        if pat.search(h):
            v = r[i]
            try:
                if float(v.replace(",", "")) == 0 and "stalled" in h:
                    continue
            except ValueError:
                pass
            print(f"{h:78s} {v[:70]} {units[i]}")
