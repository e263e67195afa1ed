#!/bin/bash
# round-1 run 10 (1 GPU): full GPU suite with the optimised peer kernel, default bench + reference arm, prefill bench
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/pytest_gpu.txt
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference exit $?"
timeout 600 python bench.py --workload cfg3p --steps 10 --no-cpu-baseline > gpurun_out/bench_cfg3p.json 2> gpurun_out/bench_cfg3p.err; echo "cfg3p exit $?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.txt 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.txt
for f in bench_default bench_reference bench_cfg3p; do echo "== $f"; tail -c 900 gpurun_out/$f.json; echo; tail -2 gpurun_out/$f.err; done
