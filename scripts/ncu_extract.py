#!/usr/bin/env python
"""Transpose `ncu -i X.ncu-rep --page raw --csv` into one row per metric, one column per
kernel (first capture of each kernel name), keeping the metrics the profiles/ README quotes.

    ncu -i gpurun_out/r1c_prof.ncu-rep --page raw --csv > /tmp/raw.csv
    python scripts/ncu_extract.py /tmp/raw.csv profiles/r1c_ncu_full_metrics.csv
"""
import csv
import re
import sys

KEEP = re.compile(
    r"gpu__time_duration|dram__bytes_(read|write)\.sum$|dram__throughput|launch__registers|"
    r"launch__occupancy|launch__grid_size|launch__block_size|shared_mem|sm__warps_active|"
    r"achieved_occupancy|lts__t_sector_hit_rate|l1tex__t_sector_hit_rate|pipe_fp64|"
    r"smsp__inst_executed\.sum$|smsp__average_warp.*_per_issue|smsp__average_warps_issue_stalled|"
    r"sm__throughput|sm__inst_executed_pipe|lts__t_bytes\.sum$|l1tex__data_pipe_lsu_wavefronts|"
    r"smsp__cycles_active\.avg$|sm__cycles_elapsed\.max$|local_(load|store)|dram__cycles")


def main(src, dst):
    rows = list(csv.reader(open(src)))
    head = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    names, units = rows[head], rows[head + 1]
    kcol = names.index("Kernel Name")
    seen, cols = [], []
    for r in rows[head + 2:]:
        short = r[kcol].split("(")[0]
        if short not in seen:
            seen.append(short)
            cols.append(r)
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + seen)
        for j, n in enumerate(names):
            if n == "Kernel Name" or KEEP.search(n):
                w.writerow([n, units[j]] + [c[j] for c in cols])


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
