#!/bin/bash
# round-1 GPU pass A: parity tests, bench (both arms), e2e engine tokens/s, ncu of the prefill kernel
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.txt
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref exit $?"
timeout 600 python bench.py --workload cfg3p --steps 10 > gpurun_out/bench_cfg3p.json 2> gpurun_out/bench_cfg3p.err; echo "cfg3p exit $?"
timeout 900 python tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 > gpurun_out/e2e_cfg3_tp1.json 2> gpurun_out/e2e_cfg3_tp1.err; echo "e2e cfg3 exit $?"
timeout 900 python tools/e2e_bench.py --model qwen3-0.6b --seqs 256 --prompt 2048 --cached 0 --new-tokens 64 --repeat 1 > gpurun_out/e2e_cfg2_tp1.json 2> gpurun_out/e2e_cfg2_tp1.err; echo "e2e cfg2 exit $?"
CMD="python bench.py --workload cfg3p --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:prefill_tcgen05 -s 40 -c 2 \
    -f -o gpurun_out/prof_prefill $CMD > gpurun_out/ncu_prefill_full.log 2>&1; echo "ncu exit $?"
