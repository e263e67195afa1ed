#!/usr/bin/env python
"""torch.profiler table of the reference engine + drop-in on a bench engine workload (where a decode step goes).

    python tools/engine_profile.py --workload cfg3e --new-tokens 24 --out gpurun_out/engine_profile.txt
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import bench
    from harness import ref_engine
    from oracle import reference_engine as R

    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg3e")
    ap.add_argument("--new-tokens", type=int, default=24)
    ap.add_argument("--eager", action="store_true")
    ap.add_argument("--no-executor", action="store_true")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "engine_profile.txt"))
    a = ap.parse_args()
    wl = dict(bench.WORKLOADS[a.workload], new_tokens=a.new_tokens)
    warm, real = bench.engine_prompts(wl, R.PRESETS[wl["preset"]]["vocab_size"])
    info = ref_engine.run(wl["preset"], device="cuda", dtype="bfloat16", dropin=True, executor=not a.no_executor,
                          enforce_eager=a.eager, fast_init=True, calls=[warm, real], max_tokens=[1, a.new_tokens],
                          profile=a.out, **bench.engine_config(wl))
    c = info["calls"][1]
    print(json.dumps({"decode_ms_per_step": c["decode_s"] / max(c["decode_steps"], 1) * 1e3, "prefill_s": c["prefill_s"],
                      "graphs": info["graphs_captured"]}))


if __name__ == "__main__":
    main()
