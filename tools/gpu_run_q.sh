#!/bin/bash
# after the block-pool rollback fix: cfg2 / cfg3 / cfg5-rank e2e on one GPU
set -u
mkdir -p gpurun_out
timeout 600 python tools/e2e_bench.py --model qwen3-0.6b --seqs 256 --prompt 2048 --cached 0 --new-tokens 64 > gpurun_out/e2e_cfg2_tp1.json 2> gpurun_out/e2e_cfg2_tp1.err; echo "cfg2 exit $?"
timeout 600 python tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 > gpurun_out/e2e_cfg3_tp1.json 2> gpurun_out/e2e_cfg3_tp1.err; echo "cfg3 exit $?"
timeout 900 python tools/e2e_bench.py --model qwen2-72b-rank8 --seqs 512 --prompt 1024 --cached 0 --new-tokens 128 --repeat 1 --watchdog 800 > gpurun_out/e2e_cfg5_rank.json 2> gpurun_out/e2e_cfg5_rank.err; echo "cfg5 rank exit $?"
timeout 600 python tools/e2e_bench.py --model qwen2-7b --seqs 1 --prompt 32768 --cached 0 --new-tokens 64 --max-model-len 33024 --repeat 1 > gpurun_out/e2e_cfg4_tp1.json 2> gpurun_out/e2e_cfg4_tp1.err; echo "cfg4 exit $?"
for f in e2e_cfg2_tp1 e2e_cfg3_tp1 e2e_cfg5_rank e2e_cfg4_tp1; do echo "== $f"; grep -h -o '"value": [0-9.]*\|"decode_ms_per_step": [0-9.]*\|"prefill_tokens_per_s": [0-9.]*\|"prefix_hit_tokens": [0-9]*' gpurun_out/$f.json | tr '\n' ' '; echo; tail -3 gpurun_out/$f.err; done
