#!/bin/bash
# full GPU pass: smoke, all GPU tests, default bench (both arms), prefill benches, e2e engine numbers
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; cat gpurun_out/smoke.log | tail -2
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.txt
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"
timeout 600 python bench.py --workload cfg3p --steps 10 > gpurun_out/bench_cfg3p.json 2> gpurun_out/bench_cfg3p.err; echo "cfg3p exit $?"
timeout 600 python bench.py --workload cfg3p1 --steps 10 --no-cpu-baseline > gpurun_out/bench_cfg3p1.json 2> gpurun_out/bench_cfg3p1.err; echo "cfg3p1 exit $?"
timeout 600 python bench.py --workload cfg4 --steps 20 --no-cpu-baseline > gpurun_out/bench_cfg4.json 2> gpurun_out/bench_cfg4.err; echo "cfg4 exit $?"
timeout 900 python tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 > gpurun_out/e2e_cfg3_tp1.json 2> gpurun_out/e2e_cfg3_tp1.err; echo "e2e cfg3 exit $?"
python - <<'PY'
import json
for f in ("bench_default","bench_cfg3p","bench_cfg3p1","bench_cfg4","e2e_cfg3_tp1"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json"))
        print(f, round(d["value"],1), d["unit"], d.get("roofline",{}).get("frac"), d.get("decode_ms_per_step"), d.get("prefill_s"))
    except Exception as e:
        print(f, "ERR", e)
PY
