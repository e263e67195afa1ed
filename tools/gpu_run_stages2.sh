#!/bin/bash
# cfg4 (1 x 32k) decode: 2-stage ring (64 KiB smem, 3 CTAs/SM could co-reside) against the default, two split counts
mkdir -p gpurun_out
for cfg in "2 37" "2 74" "3 74" "2 111"; do
  set -- $cfg
  echo "stages=$1 splits=$2" >> gpurun_out/stages2_cfg4.txt
  MVB200_DECODE_STAGES=$1 timeout 60 python bench.py --workload cfg4 --splits $2 --no-cpu-baseline --steps 50 --warmup 5 2>/dev/null | tail -1 >> gpurun_out/stages2_cfg4.txt
done
python - <<'P'
import json
for ln in open('gpurun_out/stages2_cfg4.txt'):
    ln=ln.strip()
    if ln.startswith('stages'): print(ln, end=' ')
    elif ln.startswith('{'):
        d=json.loads(ln); print(round(d['roofline']['us_per_launch'],2),'us/launch', round(d['ms_per_step'],4),'ms/step', round(d['value']),'GB/s')
P
