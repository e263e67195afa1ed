#!/bin/bash
set -u
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 140 $TR tools/e2e_bench.py --model qwen3-0.6b --seqs 4 --prompt 256 --cached 0 --new-tokens 8 --repeat 1 --eager --watchdog 90 > gpurun_out/dbg_eager.json 2> gpurun_out/dbg_eager.err; echo "eager exit $?"
grep -v "^\*\|OMP_NUM" gpurun_out/dbg_eager.err | tail -40 | cut -c1-200; cat gpurun_out/dbg_eager.json | cut -c1-300
timeout 140 $TR tools/e2e_bench.py --model qwen3-0.6b --seqs 4 --prompt 256 --cached 0 --new-tokens 8 --repeat 1 --watchdog 90 > gpurun_out/dbg_graph.json 2> gpurun_out/dbg_graph.err; echo "graph exit $?"
grep -v "^\*\|OMP_NUM" gpurun_out/dbg_graph.err | tail -40 | cut -c1-200; cat gpurun_out/dbg_graph.json | cut -c1-300
