#!/bin/bash
set -u
mkdir -p gpurun_out
OUT=gpurun_out/${1:-decode_sweep3}.jsonl
: > $OUT
run() { label=$1; shift
  python bench.py --no-cpu-baseline --steps 20 --warmup 3 "$@" 2>gpurun_out/decode_sweep.err | python -c "
import sys, json
for ln in sys.stdin:
    if ln.startswith('{'):
        d = json.loads(ln); r = d['roofline']
        print(json.dumps({'label': '$label', 'splits': d['config']['num_kv_splits'], 'us_per_launch': round(r['us_per_launch'], 2),
                          'GBps': round(r['achieved'], 1), 'frac': round(r['frac'], 3), 'step_ms': round(d['ms_per_step'], 4)}))
" >> $OUT || echo "{\"label\": \"$label\", \"error\": true}" >> $OUT
}
export MVB200_DECODE_REDUCE=kernel
run cfg4_s37_static --workload cfg4 --splits 37
run cfg4_s74_static --workload cfg4 --splits 74
run cfg4_s37_ctx2368_static --workload cfg4 --splits 37 --ctx 2368
MVB200_DECODE_DYNAMIC=1 run cfg4_s37_dynamic --workload cfg4 --splits 37
run cfg3d_s4 --workload cfg3d --splits 4
run cfg3d_s9 --workload cfg3d --splits 9
run cfg3d_s12 --workload cfg3d --splits 12
MVB200_DECODE_STAGES=4 MVB200_DECODE_CTAS_PER_SM=1 run cfg4_s37_4st_1cta --workload cfg4 --splits 37
MVB200_DECODE_STAGES=5 MVB200_DECODE_CTAS_PER_SM=1 run cfg4_s37_5st_1cta --workload cfg4 --splits 37
MVB200_DECODE_STAGES=2 run cfg4_s74_2st --workload cfg4 --splits 74
cat $OUT
