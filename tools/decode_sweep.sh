#!/bin/bash
# Decode-kernel shapes of VERDICT r1 item 4 through bench.py (one GPU): per-launch time of the decode-only
# graph for every BASELINE decode shape, plus split-count / residency sweeps on the small-work shapes.
set -u
mkdir -p gpurun_out
OUT=gpurun_out/${1:-decode_sweep}.jsonl
: > $OUT
run() { # label, args...
  label=$1; shift
  python bench.py --no-cpu-baseline --steps 20 --warmup 3 "$@" 2>gpurun_out/decode_sweep.err | python -c "
import sys, json
for ln in sys.stdin:
    if ln.startswith('{'):
        d = json.loads(ln); r = d['roofline']
        print(json.dumps({'label': '$label', 'splits': d['config']['num_kv_splits'], 'us_per_launch': round(r['us_per_launch'], 2),
                          'GBps': round(r['achieved'], 1), 'frac': round(r['frac'], 3), 'step_ms': round(d['ms_per_step'], 4),
                          'e2e_ms': round(d['e2e']['ms_per_step'], 4), 'kernels': d.get('kernels')}))
" >> $OUT || echo "{\"label\": \"$label\", \"error\": true}" >> $OUT
}
for wl in cfg2 cfg2r cfg2tp8 cfg3d cfg4 cfg5d; do run $wl --workload $wl; done
for s in 37 74; do run cfg4_s$s --workload cfg4 --splits $s; done
for s in 1 2; do run cfg2tp8_s$s --workload cfg2tp8 --splits $s; done
for s in 5 9; do run cfg3d_s$s --workload cfg3d --splits $s; done
for s in 1 2; do run cfg5d_s$s --workload cfg5d --splits $s; done
MVB200_DECODE_CTAS_PER_SM=1 run cfg4_1cta --workload cfg4
MVB200_DECODE_CTAS_PER_SM=1 run cfg3d_1cta --workload cfg3d
MVB200_DECODE_STAGES=2 run cfg2_2stages --workload cfg2
cat $OUT
