#!/bin/bash
# multi-GPU evidence at TP = N: peer all-reduce parity/latency, e2e with peer vs NCCL, bench.py --gpus N
set -u
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29611 tests/mp_peer_allreduce_check.py > gpurun_out/peer_check_tp$N.txt 2> gpurun_out/peer_check_tp$N.err; echo "peer check exit $?"; cat gpurun_out/peer_check_tp$N.txt; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/peer_check_tp$N.err | tail -5
timeout 600 $TR --master-port 29612 tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 --watchdog 500 > gpurun_out/e2e_cfg3_tp${N}_peer.json 2> gpurun_out/e2e_cfg3_tp${N}_peer.err; echo "e2e cfg3 peer exit $?"
MVB200_HARNESS_NCCL=1 timeout 600 $TR --master-port 29613 tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 --watchdog 500 > gpurun_out/e2e_cfg3_tp${N}_nccl.json 2> gpurun_out/e2e_cfg3_tp${N}_nccl.err; echo "e2e cfg3 nccl exit $?"
timeout 600 $TR --master-port 29614 tools/e2e_bench.py --model qwen3-0.6b --seqs 256 --prompt 2048 --cached 0 --new-tokens 64 --repeat 1 --watchdog 500 > gpurun_out/e2e_cfg2_tp${N}_peer.json 2> gpurun_out/e2e_cfg2_tp${N}_peer.err; echo "e2e cfg2 peer exit $?"
timeout 600 $TR --master-port 29615 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_default_tp$N.json 2> gpurun_out/bench_default_tp$N.err; echo "bench tp$N exit $?"
for f in e2e_cfg3_tp${N}_peer e2e_cfg3_tp${N}_nccl e2e_cfg2_tp${N}_peer bench_default_tp$N; do echo "== $f"; grep -h -o '"metric": "[^"]*"\|"value": [0-9.]*\|"decode_ms_per_step": [0-9.]*\|"us_per_launch": [0-9.]*' gpurun_out/$f.json | tr '\n' ' '; echo; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/$f.err | tail -3; done
