#!/usr/bin/env python
"""Summarise an `ncu --set full` report into profiles/ncu_summary.json (read by bench.py for roofline.traffic), offline.

usage: python scripts/ncu_summary.py <report.ncu-rep> [out.json]

Per kernel (averaged over the captured launches): duration, DRAM bytes read/written per launch, registers, achieved
occupancy, issue-active, fp64-pipe utilisation, executed warp instructions, active lanes per instruction. The keys "knn",
"rows_reduce", "front" are aliases bench.py looks up for the dominant stage.
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6}
WANT = {
    "gpu__time_duration.sum": "time_us",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "lts__t_bytes.sum": "l2_bytes",
    "lts__t_sectors.sum": "l2_sectors",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
    "launch__registers_per_thread": "regs",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "occ",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active": "fp64",
    "smsp__inst_executed.sum": "inst",
    "smsp__thread_inst_executed_per_inst_executed.ratio": "lanes",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
}


def main():
    rep = sys.argv[1]
    out_path = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "profiles", "ncu_summary.json")
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    H, U = rows[0], rows[1]
    ki = H.index("Kernel Name")
    acc = {}
    for r in rows[2:]:
        name = re.sub(r"^void\s+|ml::", "", r[ki]).split("(")[0]
        d = acc.setdefault(name, {"launches": 0})
        d["launches"] += 1
        for m, k in WANT.items():
            if m in H:
                i = H.index(m)
                try:
                    v = float(r[i].replace(",", "")) * UNIT.get(U[i], 1.0)
                except ValueError:
                    continue
                d[k] = d.get(k, 0.0) + v
    for d in acc.values():
        n = d["launches"]
        for k in list(d):
            if k != "launches":
                d[k] /= n
        d["dram_bytes_per_launch"] = d.get("dram_read", 0.0) + d.get("dram_write", 0.0)
        if "l2_bytes" not in d and "l2_sectors" in d:  # `--set full` carries the sector count, not lts__t_bytes: 32 B per sector
            d["l2_bytes"] = 32.0 * d["l2_sectors"]
    out = {"all": acc, "source": os.path.basename(rep)}
    for alias, pat in (("knn", "k_knn<5>"), ("knn", "k_front_knn<5"), ("rows_reduce", "k_rows_reduce<5"), ("scan_tile", "k_scan_tile<5"), ("front", "k_range_filter_bin")):
        for name, d in acc.items():
            if name.startswith(pat):
                out[alias] = d
    json.dump(out, open(out_path, "w"), indent=1)
    for name, d in acc.items():
        print(f"{name:40s} x{d['launches']} {d.get('time_us', 0):8.1f} us  dram {d['dram_bytes_per_launch'] / 1e6:7.2f} MB  l2 {d.get('l2_bytes', 0) / 1e6:7.2f} MB  regs {d.get('regs', 0):4.0f} "
              f"occ {d.get('occ', 0):5.1f}% issue {d.get('issue', 0):5.1f}% fp64 {d.get('fp64', 0):5.1f}% lanes {d.get('lanes', 0):5.1f}")


if __name__ == "__main__":
    main()
