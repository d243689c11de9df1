#!/usr/bin/env python
"""Instruction mix of one kernel from cuobjdump -sass (static counts): python scripts/sass_stats.py lib.so <substring> [--dump]"""
import collections
import re
import subprocess
import sys

so, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
fn, blocks = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1)
        blocks[fn] = []
    elif fn and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        blocks[fn].append(line)
for fn, lines in blocks.items():
    if pat not in fn:
        continue
    ops = collections.Counter()
    for ln in lines:
        t = ln.split("*/", 1)[1].strip().rstrip(";").split()
        op = t[1] if t[0].startswith("@") else t[0]
        ops[op.split(".")[0]] += 1
    print(fn[:110], "instructions:", len(lines))
    print("  " + ", ".join(f"{k}:{v}" for k, v in ops.most_common(28)))
    if "--dump" in sys.argv:
        print("\n".join(lines))
