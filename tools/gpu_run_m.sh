#!/bin/bash
# round-1 run 8 (2 GPUs): peer all-reduce over CUDA IPC / NVLink: parity + latency vs NCCL, e2e TP2 A/B
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_peer_allreduce.py -m gpu -x -q > gpurun_out/pytest_peer.txt 2>&1; echo "peer pytest exit $?"; tail -8 gpurun_out/pytest_peer.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29611 tests/mp_peer_allreduce_check.py > gpurun_out/peer_check_tp2.txt 2> gpurun_out/peer_check_tp2.err; echo "peer check exit $?"; cat gpurun_out/peer_check_tp2.txt; tail -5 gpurun_out/peer_check_tp2.err
timeout 600 $TR --master-port 29612 tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 --watchdog 500 > gpurun_out/e2e_cfg3_tp2_peer.json 2> gpurun_out/e2e_cfg3_tp2_peer.err; echo "e2e cfg3 tp2 peer exit $?"
MVB200_HARNESS_NCCL=1 timeout 600 $TR --master-port 29613 tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 --watchdog 500 > gpurun_out/e2e_cfg3_tp2_nccl.json 2> gpurun_out/e2e_cfg3_tp2_nccl.err; echo "e2e cfg3 tp2 nccl exit $?"
timeout 600 $TR --master-port 29614 tools/e2e_bench.py --model qwen3-0.6b --seqs 256 --prompt 2048 --cached 0 --new-tokens 64 --watchdog 500 > gpurun_out/e2e_cfg2_tp2_peer.json 2> gpurun_out/e2e_cfg2_tp2_peer.err; echo "e2e cfg2 tp2 peer exit $?"
MVB200_HARNESS_NCCL=1 timeout 600 $TR --master-port 29615 tools/e2e_bench.py --model qwen3-0.6b --seqs 256 --prompt 2048 --cached 0 --new-tokens 64 --watchdog 500 > gpurun_out/e2e_cfg2_tp2_nccl.json 2> gpurun_out/e2e_cfg2_tp2_nccl.err; echo "e2e cfg2 tp2 nccl exit $?"
python - <<'PY'
import json
for f in ("e2e_cfg3_tp2_peer","e2e_cfg3_tp2_nccl","e2e_cfg2_tp2_peer","e2e_cfg2_tp2_nccl"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json"))
        print(f, round(d["value"],1), d["unit"], d.get("decode_ms_per_step"), d.get("prefill_s"))
    except Exception as e:
        print(f, "ERR", e); print(open(f"gpurun_out/{f}.err").read()[-1500:])
PY
