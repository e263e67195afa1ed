#!/bin/bash
# ncu launch list of the end-to-end engine (prefill + a few eager decode steps), one GPU.
# Per-launch times under ncu are serialised and cold-cache: use the SHARES, not the absolute values.
set -u
mkdir -p gpurun_out
CMD="python tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 4 --repeat 1 --eager"
$CMD > gpurun_out/e2e_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv \
    --log-file gpurun_out/launches_e2e.csv $CMD > gpurun_out/ncu_e2e.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/e2e_plain.log | cut -c1-300; wc -l gpurun_out/launches_e2e.csv
