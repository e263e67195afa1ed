#!/bin/bash
# ncu evidence for the prefill kernel (run on the B200 box under gpurun; one GPU).
# Recipe: /opt/skills/guides/B200_PROFILING.md.  Outputs land in gpurun_out/.
set -u
mkdir -p gpurun_out
CMD="python bench.py --workload cfg3p --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/ncu_prefill_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/launches_prefill.csv $CMD > gpurun_out/ncu_prefill_launches.log 2>&1
$CMD > gpurun_out/ncu_prefill_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:prefill_fa2q -s 40 -c 2 \
    -f -o gpurun_out/prof_prefill $CMD > gpurun_out/ncu_prefill_full.log 2>&1
echo "profile_prefill done"
