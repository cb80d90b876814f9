#!/usr/bin/env python
"""Summarise an `ncu --set full` report into the per-kernel JSON that bench.py reads for `roofline.traffic`.

    ncu -i gpurun_out/prof.ncu-rep --page raw --csv > raw.csv     (done here through subprocess)
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/r1_ncu_full_summary.json

Per kernel (template arguments shortened): launches profiled, mean duration, DRAM bytes read / written per
launch, registers, grid/block, achieved warps, the register occupancy limit."""
import csv
import io
import json
import re
import subprocess
import sys

COLS = {
    "gpu__time_duration.sum": "dur",
    "dram__bytes_read.sum": "dram_rd",
    "dram__bytes_write.sum": "dram_wr",
    "launch__registers_per_thread": "regs",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct",
    "launch__occupancy_limit_registers": "occupancy_limit_regs",
    "smsp__inst_executed.sum": "warp_instructions",
}
SCALE = {"nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "ns": 1e-3, "us": 1.0, "ms": 1e3,
         "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def short(name):
    name = name.split("(")[0] if not name.startswith("void (") else name
    m = re.search(r"(k_[A-Za-z0-9_]+(<.*>)?)", name)
    return (m.group(1) if m else name).strip()


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, body = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(head)}
    acc = {}
    for r in body:
        k = short(r[idx["Kernel Name"]])
        a = acc.setdefault(k, {"n": 0})
        a["n"] += 1
        for col, key in COLS.items():
            if col not in idx or r[idx[col]] in ("", "n/a"):
                continue
            v = float(r[idx[col]].replace(",", "")) * SCALE.get(units[idx[col]], 1.0)
            a[key] = a.get(key, 0.0) + v
    summary = {}
    for k, a in acc.items():
        n = a.pop("n")
        e = {key: a[key] / n for key in a}
        if "dur" in e:
            e["dur_us"] = e.pop("dur")
        e["launches_profiled"] = n
        e["dram_traffic_bytes_per_launch"] = e.get("dram_rd", 0.0) + e.get("dram_wr", 0.0)
        summary[k] = e
    with open(out, "w") as f:
        json.dump(summary, f, indent=1)
    for k, e in summary.items():
        print("%-60s n=%d  %.2f us  dram %.2f MB" % (k, e["launches_profiled"], e.get("dur_us", 0), e["dram_traffic_bytes_per_launch"] / 1e6))


if __name__ == "__main__":
    main()
