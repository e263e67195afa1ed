#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 300 python tools/prefill_trace.py > gpurun_out/prefill_trace.log 2>&1; echo "trace exit $?"
timeout 300 probes/umma_pacing 10 > gpurun_out/umma_pacing2.log 2>&1; echo "pacing exit $?"; grep -E "LDTM|ex2" gpurun_out/umma_pacing2.log
