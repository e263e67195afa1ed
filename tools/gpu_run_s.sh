#!/bin/bash
# 1 GPU: two-shot peer all-reduce with in-process ranks + the whole GPU suite
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_peer_allreduce.py -m gpu -x -q > gpurun_out/pytest_peer.txt 2>&1; echo "peer pytest exit $?"; tail -12 gpurun_out/pytest_peer.txt
