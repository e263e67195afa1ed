#!/bin/bash
set -u
mkdir -p gpurun_out
for pc in ${PCS:-1 4}; do for emu in ${EMUS:-32}; do
  MVB200_PREFILL_PCHUNKS=$pc MVB200_PREFILL_EMU=$emu timeout 300 python bench.py --workload cfg3p --steps 10 --no-cpu-baseline > gpurun_out/cfg3p_pc${pc}_emu$emu.json 2> gpurun_out/cfg3p_pc${pc}_emu$emu.err
  python -c "import json;d=json.load(open('gpurun_out/cfg3p_pc${pc}_emu$emu.json'));print('pc',$pc,'emu',$emu,d['roofline']['kernel'],d['roofline']['achieved'],d['roofline']['us_per_launch'],d['clocks']['sm_mhz'])"
done; done
