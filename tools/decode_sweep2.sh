#!/bin/bash
# fixed-overhead probe of the split path: same unit counts, tiny contexts
set -u
mkdir -p gpurun_out
OUT=gpurun_out/${1:-decode_sweep2}.jsonl
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
run cfg4 --workload cfg4
run cfg4_s37 --workload cfg4 --splits 37
run cfg4_s37_ctx2368 --workload cfg4 --splits 37 --ctx 2368
run cfg4_s1_ctx64 --workload cfg4 --splits 1 --ctx 64
run cfg4_s1_ctx2368 --workload cfg4 --splits 1 --ctx 2368
run cfg4_s2_ctx2368 --workload cfg4 --splits 2 --ctx 2368
run cfg3d --workload cfg3d
run cfg3d_s9 --workload cfg3d --splits 9
run cfg3d_s1 --workload cfg3d --splits 1
run cfg3d_s1_ctx64 --workload cfg3d --splits 1 --ctx 64
run cfg2tp8 --workload cfg2tp8
run cfg5d --workload cfg5d
MVB200_NO_PDL=1 run cfg4_s37_nopdl --workload cfg4 --splits 37
cat $OUT
