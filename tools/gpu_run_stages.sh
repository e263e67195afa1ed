#!/bin/bash
# cfg4 (1 x 32k) decode: TMA ring depth sweep (MVB200_DECODE_STAGES), one CTA per SM
mkdir -p gpurun_out
for s in 3 4 5 6; do
  echo "stages=$s" >> gpurun_out/stages_cfg4.txt
  MVB200_DECODE_STAGES=$s timeout 60 python bench.py --workload cfg4 --no-cpu-baseline --steps 50 --warmup 5 2>/dev/null | tail -1 >> gpurun_out/stages_cfg4.txt
done
python - <<'P'
import json
for ln in open('gpurun_out/stages_cfg4.txt'):
    ln=ln.strip()
    if ln.startswith('stages'): print(ln, end=' ')
    elif ln.startswith('{'):
        d=json.loads(ln); print(round(d['roofline']['us_per_launch'],2),'us/launch', round(d['ms_per_step'],4),'ms/step', round(d['value']),'GB/s')
P
