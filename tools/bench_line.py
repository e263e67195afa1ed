#!/usr/bin/env python
"""stdin: bench.py output; stdout: one compact JSON line (label, us per launch, roofline fraction, step/e2e ms)."""
import json
import sys

label = sys.argv[1]
for ln in sys.stdin:
    if ln.startswith("{"):
        d = json.loads(ln)
        r = d["roofline"]
        print(json.dumps({"label": label, "splits": d["config"].get("num_kv_splits"), "us_per_launch": round(r["us_per_launch"], 2),
                          "achieved": round(r["achieved"], 1), "unit": r["unit"], "frac": round(r["frac"], 3),
                          "step_ms": round(d["ms_per_step"], 4), "e2e_ms": round(d["e2e"]["ms_per_step"], 4),
                          "kernels": d.get("kernels"), "clocks": d.get("clocks", {}).get("sm_mhz")}))
