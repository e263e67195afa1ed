#!/bin/bash
# round-1 run 9 (2 GPUs): peer all-reduce after the epoch-ordering fix: tests, parity + latency table vs NCCL, e2e cfg2 TP2
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_peer_allreduce.py -m gpu -x -q > gpurun_out/pytest_peer.txt 2>&1; echo "peer pytest exit $?"; tail -8 gpurun_out/pytest_peer.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29611 tests/mp_peer_allreduce_check.py > gpurun_out/peer_check_tp2.txt 2> gpurun_out/peer_check_tp2.err; echo "peer check exit $?"; cat gpurun_out/peer_check_tp2.txt; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/peer_check_tp2.err | tail -5
timeout 600 $TR --master-port 29614 tools/e2e_bench.py --model qwen3-0.6b --seqs 256 --prompt 2048 --cached 0 --new-tokens 64 --watchdog 500 > gpurun_out/e2e_cfg2_tp2_peer.json 2> gpurun_out/e2e_cfg2_tp2_peer.err; echo "e2e cfg2 tp2 peer exit $?"
timeout 600 $TR --master-port 29612 tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 --watchdog 500 > gpurun_out/e2e_cfg3_tp2_peer.json 2> gpurun_out/e2e_cfg3_tp2_peer.err; echo "e2e cfg3 tp2 peer exit $?"
grep -h -o '"metric": "[^"]*"\|"value": [0-9.]*\|"decode_ms_per_step": [0-9.]*' gpurun_out/e2e_cfg2_tp2_peer.json gpurun_out/e2e_cfg3_tp2_peer.json | tr '\n' ' '
