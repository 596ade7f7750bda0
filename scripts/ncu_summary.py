#!/usr/bin/env python
"""Turn ncu outputs brought back in gpurun_out/ into the small text summaries kept under profiles/.

    python scripts/ncu_summary.py launches gpurun_out/launches_TAG.csv  > profiles/TAG_launches.md
    python scripts/ncu_summary.py full     gpurun_out/prof_TAG.ncu-rep > profiles/TAG_full.md

`launches`: per-kernel count, total and mean device time and SHARE of the listed window
(cold-cache, serialised: shares are meaningful, absolutes are not).
`full`: per captured launch the DRAM traffic, duration, DRAM throughput, registers, occupancy,
L1/L2 hit rates (read with `ncu -i ... --page raw --csv`, no GPU needed).
"""
from __future__ import annotations

import csv
import io
import subprocess
import sys
from collections import OrderedDict


def launches(path: str) -> None:
    rows = [r for r in csv.reader(open(path)) if r and r[0].isdigit()]
    agg: "OrderedDict[str, list[float]]" = OrderedDict()
    for r in rows:
        name, val = r[4], float(r[-1])
        agg.setdefault(name, []).append(val)
    total = sum(sum(v) for v in agg.values())
    print(f"source: {path} ({len(rows)} launches, {total / 1e6:.3f} ms of device time in the window)\n")
    print("| kernel | launches | total ms | mean us | share |")
    print("|---|---:|---:|---:|---:|")
    for name, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"| `{name}` | {len(v)} | {sum(v) / 1e6:.3f} | {sum(v) / len(v) / 1e3:.2f} | {100 * sum(v) / total:.1f} % |")


WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("smsp__inst_executed.sum", "warp instructions"),
]


def full(path: str) -> None:
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    cols = [(hdr.index(k), label) for k, label in WANT if k in hdr]
    kname = hdr.index("Kernel Name")
    print(f"source: {path} (ncu --set full --clock-control none; per launch)\n")
    print("| kernel | " + " | ".join(label for _, label in cols) + " | dram total | achieved dram GB/s |")
    print("|---|" + "---:|" * (len(cols) + 2))
    for r in rows[2:]:
        vals = []
        d = {}
        for i, label in cols:
            vals.append(f"{r[i]} {units[i]}".strip())
            d[label] = (float(r[i].replace(",", "")), units[i])
        def to_bytes(v):
            x, u = v
            return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        def to_sec(v):
            x, u = v
            return x * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1}.get(u, 1e-9)
        tot = to_bytes(d["dram read"]) + to_bytes(d["dram write"])
        gbs = tot / to_sec(d["duration"]) / 1e9
        print(f"| `{r[kname]}` | " + " | ".join(vals) + f" | {tot / 1e6:.1f} MB | {gbs:.0f} |")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
