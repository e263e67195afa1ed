#!/bin/bash
# cfg5: Qwen2-72B dims at TP=8, max_num_seqs=512, max_num_batched_tokens=16384, 512 prompts x 1024 tokens, 128 new tokens
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29621 tools/e2e_bench.py --model qwen2-72b --seqs 512 --prompt 1024 --cached 0 --new-tokens 128 --repeat 1 --watchdog 800 > gpurun_out/e2e_cfg5_tp8.json 2> gpurun_out/e2e_cfg5_tp8.err; echo "cfg5 tp8 exit $?"
grep -h -o '"value": [0-9.]*\|"decode_ms_per_step": [0-9.]*\|"prefill_tokens_per_s": [0-9.]*\|"prefix_hit_tokens": [0-9]*' gpurun_out/e2e_cfg5_tp8.json | tr '\n' ' '; echo; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/e2e_cfg5_tp8.err | tail -5
