#!/bin/bash
# round-1 run 6: device-side decode bookkeeping (decode_advance + greedy_argmax), A/B against the host loop
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_gpu.txt
timeout 600 python tools/e2e_bench.py --model qwen3-0.6b --seqs 256 --prompt 2048 --cached 0 --new-tokens 64 > gpurun_out/e2e_cfg2_loop.json 2> gpurun_out/e2e_cfg2_loop.err; echo "e2e cfg2 loop exit $?"
MVB200_HARNESS_HOST_LOOP=1 timeout 600 python tools/e2e_bench.py --model qwen3-0.6b --seqs 256 --prompt 2048 --cached 0 --new-tokens 64 > gpurun_out/e2e_cfg2_host.json 2> gpurun_out/e2e_cfg2_host.err; echo "e2e cfg2 host exit $?"
timeout 600 python tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 > gpurun_out/e2e_cfg3_loop.json 2> gpurun_out/e2e_cfg3_loop.err; echo "e2e cfg3 loop exit $?"
MVB200_HARNESS_HOST_LOOP=1 timeout 600 python tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 > gpurun_out/e2e_cfg3_host.json 2> gpurun_out/e2e_cfg3_host.err; echo "e2e cfg3 host exit $?"
python - <<'PY'
import json
for f in ("e2e_cfg2_loop","e2e_cfg2_host","e2e_cfg3_loop","e2e_cfg3_host"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json"))
        print(f, round(d["value"],1), d["unit"], d.get("decode_ms_per_step"), d.get("prefill_s"))
    except Exception as e:
        print(f, "ERR", e); print(open(f"gpurun_out/{f}.err").read()[-1500:])
PY
