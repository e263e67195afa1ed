#!/bin/bash
# round-1 run 7 (1 GPU): peer all-reduce kernel with in-process ranks + full GPU suite
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_peer_allreduce.py -m gpu -x -q > gpurun_out/pytest_peer.txt 2>&1; echo "peer pytest exit $?"; tail -15 gpurun_out/pytest_peer.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.txt
