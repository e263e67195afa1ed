#!/bin/bash
# ncu --set full of the 128-bit qk_norm_rope_store kernel and greedy_argmax at BASELINE sizes (one GPU)
set -u
mkdir -p gpurun_out
CMD="python tools/bench_aux_kernels.py"
$CMD > gpurun_out/aux_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:qk_rope_store_vec_kernel|greedy_argmax_kernel" -c 8 \
    -f -o gpurun_out/prof_aux $CMD > gpurun_out/ncu_aux.log 2>&1
echo "ncu exit $?"
ncu -i gpurun_out/prof_aux.ncu-rep --page raw --csv > gpurun_out/prof_aux_raw.csv 2>/dev/null; wc -c gpurun_out/prof_aux_raw.csv
