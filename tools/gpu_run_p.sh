#!/bin/bash
# cfg5 per-rank emulation on one GPU: Qwen2-72B TP8 rank shapes, 512 seqs x 1024 prompt, 128 new tokens
set -u
mkdir -p gpurun_out
timeout 900 python tools/e2e_bench.py --model qwen2-72b-rank8 --seqs 512 --prompt 1024 --cached 0 --new-tokens 128 --repeat 1 --watchdog 800 > gpurun_out/e2e_cfg5_rank.json 2> gpurun_out/e2e_cfg5_rank.err; echo "e2e cfg5 rank exit $?"
tail -c 1200 gpurun_out/e2e_cfg5_rank.json; tail -5 gpurun_out/e2e_cfg5_rank.err
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
