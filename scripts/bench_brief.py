#!/usr/bin/env python
"""print the few numbers of a bench.py JSON line that matter while tuning: python scripts/bench_brief.py file.json [label]"""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
st = d.get("stage_ms_per_step", {})
print((sys.argv[2] if len(sys.argv) > 2 else ""), "value %.4g pts/s  dev %.1f us  e2e %.1f us  launches %s | " % (
    d["value"], d["ms_per_step"] * 1e3, d["e2e"]["ms_per_step"] * 1e3, d.get("gpu_launches")) +
    "  ".join(f"{k} {v * 1e3:.1f}" for k, v in st.items()), "| roofline frac %.4f (%s)" % (d["roofline"]["frac"], d["roofline"]["kernel"][:14]))
