#!/bin/bash
# cfg4 split-count sweep, e2e per-kernel profile (Qwen3-4B decode B=8), cfg2 e2e
set -u
mkdir -p gpurun_out
for s in ${SPLITS:-16 32 37 64 74 96 111 128}; do
  timeout 200 python bench.py --workload cfg4 --steps 20 --no-cpu-baseline --splits $s > gpurun_out/cfg4_s$s.json 2> gpurun_out/cfg4_s$s.err
  python -c "import json;d=json.load(open('gpurun_out/cfg4_s$s.json'));print('splits',$s,round(d['roofline']['us_per_launch'],2),'us',round(d['roofline']['frac'],3))"
done
timeout 600 python tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 64 --repeat 1 --profile gpurun_out/e2e_cfg3_profile.txt > gpurun_out/e2e_cfg3_short.json 2> gpurun_out/e2e_cfg3_short.err; echo "e2e exit $?"
head -60 gpurun_out/e2e_cfg3_profile.txt | cut -c1-95,180-260
