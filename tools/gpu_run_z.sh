#!/bin/bash
# end state of round 1 (1 GPU): full GPU suite + smoke + default bench (both arms)
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.txt 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.txt
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"; grep -o '"value": [0-9.]*\|"frac": [0-9.]*\|"us_per_launch": [0-9.]*' gpurun_out/bench_default.json | tr '\n' ' '; echo
