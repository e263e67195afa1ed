#!/bin/bash
# round-1 run 5: new reduce/PDL/heuristic, fused qk-norm+rope+store, harness fused elementwise
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_gpu.txt
timeout 300 python bench.py --workload cfg4 --steps 20 --no-cpu-baseline > gpurun_out/bench_cfg4.json 2> gpurun_out/bench_cfg4.err; echo "cfg4 exit $?"
MVB200_NO_PDL=1 timeout 300 python bench.py --workload cfg4 --steps 20 --no-cpu-baseline > gpurun_out/bench_cfg4_nopdl.json 2> gpurun_out/bench_cfg4_nopdl.err; echo "cfg4 nopdl exit $?"
timeout 300 python bench.py --workload cfg4 --steps 20 --no-cpu-baseline --splits 32 > gpurun_out/bench_cfg4_s32.json 2> gpurun_out/bench_cfg4_s32.err
timeout 300 python bench.py --workload cfg4 --steps 20 --no-cpu-baseline --splits 74 > gpurun_out/bench_cfg4_s74.json 2> gpurun_out/bench_cfg4_s74.err
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"
timeout 900 python tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 --profile gpurun_out/e2e_cfg3_profile.txt > gpurun_out/e2e_cfg3_tp1.json 2> gpurun_out/e2e_cfg3_tp1.err; echo "e2e cfg3 exit $?"
MVB200_HARNESS_UNFUSED=1 timeout 900 python tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 > gpurun_out/e2e_cfg3_tp1_unfused.json 2> gpurun_out/e2e_cfg3_tp1_unfused.err; echo "e2e cfg3 unfused exit $?"
python - <<'PY'
import json
for f in ("bench_default","bench_cfg4","bench_cfg4_nopdl","bench_cfg4_s32","bench_cfg4_s74","e2e_cfg3_tp1","e2e_cfg3_tp1_unfused"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json"))
        print(f, round(d["value"],1), d["unit"], d.get("roofline",{}).get("frac"), d.get("roofline",{}).get("us_per_launch"), d.get("config",{}).get("num_kv_splits"), d.get("decode_ms_per_step"), d.get("prefill_s"))
    except Exception as e:
        print(f, "ERR", e)
PY
head -45 gpurun_out/e2e_cfg3_profile.txt | cut -c1-95,180-260
