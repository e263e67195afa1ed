#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "prefill or layer or smoke" > gpurun_out/pytest_prefill.txt 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_prefill.txt
for pc in ${PCS:-1 2 4}; do for emu in ${EMUS:-0 32}; do
  MVB200_PREFILL_PCHUNKS=$pc MVB200_PREFILL_EMU=$emu timeout 300 python bench.py --workload cfg3p --steps 10 --no-cpu-baseline > gpurun_out/cfg3p_pc${pc}_emu$emu.json 2> gpurun_out/cfg3p_pc${pc}_emu$emu.err
  python -c "import json;d=json.load(open('gpurun_out/cfg3p_pc${pc}_emu$emu.json'));print('pc',$pc,'emu',$emu,d['roofline']['kernel'],d['roofline']['achieved'],d['roofline']['us_per_launch'],d['clocks']['sm_mhz'])"
done; done
MVB200_PREFILL_PCHUNKS=${TRACE_PC:-4} MVB200_PREFILL_EMU=${TRACE_EMU:-32} timeout 120 python tools/prefill_trace.py > gpurun_out/prefill2_trace.log 2>&1; echo "trace exit $?"
