#!/bin/bash
# ncu evidence for the decode kernel (run on the B200 box under gpurun; one GPU).
# Recipe: /opt/skills/guides/B200_PROFILING.md.  Outputs land in gpurun_out/.
set -u
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/launches_decode.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/ncu_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:paged_decode_kernel -s 60 -c 3 \
    -f -o gpurun_out/prof_decode $CMD > gpurun_out/ncu_full.log 2>&1
echo "profile_decode done"
