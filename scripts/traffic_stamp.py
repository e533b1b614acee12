#!/usr/bin/env python
"""profiles/r2_traffic.json from the `ncu --set full` capture of scripts/ncu_round2.sh: DRAM bytes
(read + written) per launch of the three hot kernels, stamped with the hash of the kernel sources
they were measured on (bench.py prints `traffic: null` once the sources differ).

    ncu -i gpurun_out/r2_prof.ncu-rep --page raw --csv > /tmp/raw.csv
    python scripts/ncu_extract.py /tmp/raw.csv profiles/r2_ncu_full_metrics.csv
    python scripts/traffic_stamp.py profiles/r2_ncu_full_metrics.csv profiles/r2_traffic.json
"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main(src, dst):
    from bench import kernel_sources_stamp
    rows = list(csv.reader(open(src)))
    kernels = rows[0][2:]
    total = [0.0] * len(kernels)
    for r in rows[1:]:
        if r[0].endswith("dram__bytes_read.sum") or r[0].endswith("dram__bytes_write.sum"):
            for k in range(len(kernels)):
                total[k] += float(r[2 + k].replace(",", "")) * SCALE[r[1]]
    out = {
        "source": "profiles/r2_ncu_full_metrics.csv (ncu --set full --clock-control none, one launch each, "
                  "scripts/ncu_round2.sh: 64 taxa x 1M sites, 122 candidates, 105 directional CLVs; "
                  "compact 13-double sumtables)",
        "kernel_sources_sha16": kernel_sources_stamp(),
        "stamp_of": "sha256 over pt_b200/csrc/kernels.cuh + blo.cuh + engine.cu, first 16 hex digits; bench.py "
                    "prints traffic: null when the sources no longer match",
    }
    for name, b in zip(kernels, total):
        short = "partials" if "partials" in name else "nni_prepare" if "nni_prepare" in name else "newton"
        out[short] = {"dram_bytes_per_launch": b}
    with open(dst, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
