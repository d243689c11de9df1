#!/usr/bin/env python
"""Attribute ncu warp-stall samples and executed instructions to CUDA source lines, offline (no GPU).

usage: python scripts/ncu_lines.py <report.ncu-rep> <kernel regex> [top N]

`ncu --page source --csv` prints SASS rows (address, samples, instructions executed); the CUDA line of each SASS
instruction comes from `nvdisasm -g` on the cubin extracted from mins_b200/libmins_lidar.so (compiled with -lineinfo).
Instructions are matched by their offset inside the function.
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def sass_rows(rep, kernel):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kernel}"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    kname, H, data = None, None, []
    for r in rows:
        if r and r[0] == "Kernel Name":
            if kname is not None:
                break  # first launch only
            kname = r[1]
        elif r and r[0] == "Address":
            H = r
        elif H and len(r) == len(H):
            data.append(r)
    return kname, H, data


def line_table(mangled_hint):
    so = os.environ.get("MINS_LIDAR_LIB") or os.path.join(ROOT, "mins_b200", "libmins_lidar.so")
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=d, capture_output=True)
        cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
    table, cur_fn, cur_line = {}, None, None
    for ln in txt.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m:
            cur_fn, cur_line = m.group(1), None
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur_line = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m and cur_fn and mangled_hint in cur_fn:
            table.setdefault(cur_fn, {})[int(m.group(1), 16)] = (cur_line, m.group(2).strip())
    return table


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    kname, H, data = sass_rows(rep, kernel)
    print("kernel:", kname)
    hint = re.search(r"(k_[a-z_0-9]+)", kname).group(1)
    tables = line_table(hint)
    ia, isamp, iinst = H.index("Address"), H.index("# Samples"), H.index("Instructions Executed")
    base = min(int(r[ia], 16) for r in data)
    # pick the function variant whose instruction count matches
    fn = min(tables, key=lambda f: abs(len(tables[f]) - len(data)))
    tab = tables[fn]
    per_line, tot_s, tot_i = {}, 0, 0
    for r in data:
        off = int(r[ia], 16) - base
        s, i = int(r[isamp] or 0), int(r[iinst] or 0)
        line = tab.get(off, (None, ""))[0]
        a = per_line.setdefault(line, [0, 0])
        a[0] += s
        a[1] += i
        tot_s += s
        tot_i += i
    src_cache = {}

    def src(line):
        if not line:
            return ""
        f, n = line
        for dp, _, fs in os.walk(os.path.join(ROOT, "mins_b200", "csrc")):
            if f in fs:
                if f not in src_cache:
                    src_cache[f] = open(os.path.join(dp, f)).read().splitlines()
                return src_cache[f][n - 1].strip()[:100] if n - 1 < len(src_cache[f]) else ""
        return ""

    print(f"function {fn[:60]}  sass {len(data)} instrs  samples {tot_s}  warp-instructions {tot_i}")
    print(f"{'samples':>8} {'%':>6} {'instr':>9} {'%':>6}  line")
    for line, (s, i) in sorted(per_line.items(), key=lambda x: -x[1][0])[:top]:
        print(f"{s:8d} {100 * s / max(tot_s, 1):6.1f} {i:9d} {100 * i / max(tot_i, 1):6.1f}  {line}  {src(line)}")


if __name__ == "__main__":
    main()
