#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_decode_meta.py tests/test_harness.py -m gpu -x -q > gpurun_out/pytest_meta.txt 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_meta.txt
timeout 300 python tools/bench_aux_kernels.py > gpurun_out/aux_kernels.txt 2> gpurun_out/aux_kernels.err; echo "aux exit $?"; grep "argmax" gpurun_out/aux_kernels.txt | cut -c1-160
