#!/bin/bash
# 8 GPUs: one-shot / two-shot / NCCL latency table, cfg5 (Qwen2-72B TP8) and cfg2 TP8 e2e with the size-selected peer kernels
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29611 tests/mp_peer_allreduce_check.py > gpurun_out/peer_check_tp8.txt 2> gpurun_out/peer_check_tp8.err; echo "peer check exit $?"; cat gpurun_out/peer_check_tp8.txt; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/peer_check_tp8.err | tail -5
timeout 600 $TR --master-port 29621 tools/e2e_bench.py --model qwen2-72b --seqs 512 --prompt 1024 --cached 0 --new-tokens 128 --repeat 1 --watchdog 500 > gpurun_out/e2e_cfg5_tp8.json 2> gpurun_out/e2e_cfg5_tp8.err; echo "cfg5 tp8 exit $?"
timeout 600 $TR --master-port 29614 tools/e2e_bench.py --model qwen3-0.6b --seqs 256 --prompt 2048 --cached 0 --new-tokens 64 --repeat 1 --watchdog 500 > gpurun_out/e2e_cfg2_tp8_peer.json 2> gpurun_out/e2e_cfg2_tp8_peer.err; echo "e2e cfg2 exit $?"
for f in e2e_cfg5_tp8 e2e_cfg2_tp8_peer; do echo "== $f"; grep -h -o '"value": [0-9.]*\|"decode_ms_per_step": [0-9.]*\|"prefill_tokens_per_s": [0-9.]*' gpurun_out/$f.json | tr '\n' ' '; echo; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/$f.err | tail -4; done
