#!/bin/bash
# final multi-GPU regression of the round's end state at TP4: peer check + e2e cfg3/cfg2
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29611 tests/mp_peer_allreduce_check.py > gpurun_out/peer_check_tp4.txt 2> gpurun_out/peer_check_tp4.err; echo "peer check exit $?"; cut -c1-330 gpurun_out/peer_check_tp4.txt
timeout 600 $TR --master-port 29612 tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 --watchdog 500 > gpurun_out/e2e_cfg3_tp4_peer.json 2> gpurun_out/e2e_cfg3_tp4_peer.err; echo "e2e cfg3 exit $?"
timeout 600 $TR --master-port 29614 tools/e2e_bench.py --model qwen3-0.6b --seqs 256 --prompt 2048 --cached 0 --new-tokens 64 --repeat 1 --watchdog 500 > gpurun_out/e2e_cfg2_tp4_peer.json 2> gpurun_out/e2e_cfg2_tp4_peer.err; echo "e2e cfg2 exit $?"
for f in e2e_cfg3_tp4_peer e2e_cfg2_tp4_peer; do echo "== $f"; grep -h -o '"value": [0-9.]*\|"decode_ms_per_step": [0-9.]*\|"prefill_tokens_per_s": [0-9.]*' gpurun_out/$f.json | tr '\n' ' '; echo; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/$f.err | tail -3; done
