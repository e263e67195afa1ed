#!/bin/bash
# multi-GPU evidence: bench.py and the e2e engine under torchrun at N GPUs of one node (TP = N)
set -u
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_default_tp$N.json 2> gpurun_out/bench_default_tp$N.err; echo "bench tp$N exit $?"
timeout 600 $TR bench.py --gpus $N --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference_tp$N.json 2> gpurun_out/bench_reference_tp$N.err; echo "reference tp$N exit $?"
timeout 600 $TR bench.py --gpus $N --workload cfg3p --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cfg3p_tp$N.json 2> gpurun_out/bench_cfg3p_tp$N.err; echo "cfg3p tp$N exit $?"
timeout 600 $TR bench.py --gpus $N --workload cfg4 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cfg4_tp$N.json 2> gpurun_out/bench_cfg4_tp$N.err; echo "cfg4 tp$N exit $?"
timeout 900 $TR tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 > gpurun_out/e2e_cfg3_tp$N.json 2> gpurun_out/e2e_cfg3_tp$N.err; echo "e2e cfg3 tp$N exit $?"
timeout 900 $TR tools/e2e_bench.py --model qwen3-0.6b --seqs 256 --prompt 2048 --cached 0 --new-tokens 64 --repeat 1 > gpurun_out/e2e_cfg2_tp$N.json 2> gpurun_out/e2e_cfg2_tp$N.err; echo "e2e cfg2 tp$N exit $?"
for f in bench_default bench_reference bench_cfg3p bench_cfg4 e2e_cfg3 e2e_cfg2; do echo "== $f"; tail -c 600 gpurun_out/${f}_tp$N.json; echo; tail -3 gpurun_out/${f}_tp$N.err; done
