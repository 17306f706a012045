#!/usr/bin/env python3
"""Markdown summary of an `ncu --set full` report: one row per kernel (first captured launch of each name).

usage: ncu_summary.py report.ncu-rep [title] [traffic.json]  > profiles/xyz.md
Prints the table and a JSON line with DRAM bytes (read + write) per launch.  With a third argument the same numbers are
written as the traffic file bench.py reads (profiles/r02_traffic.json), stamped with the digest of the CUDA sources in
this tree (bench.source_digest): bench.py refuses the file when the sources have changed since the capture.
"""
import csv
import io
import json
import subprocess
import sys

METRICS = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "gpu__time_duration.sum",
           "launch__registers_per_thread", "lts__t_sector_hit_rate.pct", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum"]
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rep = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 else rep
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units, body = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(head)}
    kcol = col["Kernel Name"]
    seen, traffic = {}, {}
    for r in body:
        name = r[kcol].split("(")[0].split("::")[-1]
        if name in seen:
            continue
        seen[name] = r
    print(f"# {title}")
    print("| kernel | " + " | ".join(METRICS) + " |")
    print("|---|" + "---|" * len(METRICS))
    for name, r in seen.items():
        cells = []
        for m in METRICS:
            i = col.get(m)
            cells.append("n/a" if i is None else f"{r[i]} {units[i]}".strip())
        print(f"| {name} | " + " | ".join(cells) + " |")
        tot = 0.0
        for m in METRICS[:2]:
            i = col[m]
            tot += float(r[i].replace(",", "")) * SCALE.get(units[i], 1)
        traffic[name] = int(tot)
    print()
    print("DRAM traffic per launch (read + write), bytes: " + json.dumps(traffic))
    if len(sys.argv) > 3:
        import os
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        import bench
        doc = {"source": "%s (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, first captured launch of each kernel)" % title,
               "source_digest": bench.source_digest(), "kernels": {k: {"dram_bytes_per_launch": v} for k, v in traffic.items()}}
        json.dump(doc, open(sys.argv[3], "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
