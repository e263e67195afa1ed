#!/bin/bash
# e2e leg of bench.py: copies on side streams (default) against everything on one stream (MVB200_E2E_SERIAL=1)
mkdir -p gpurun_out
timeout 100 python bench.py > gpurun_out/bench_run25_default.json 2> gpurun_out/bench_run25_default.err
MVB200_E2E_SERIAL=1 timeout 60 python bench.py --no-cpu-baseline > gpurun_out/bench_run25_default_serial.json 2>/dev/null
timeout 60 python bench.py --workload cfg3p --no-cpu-baseline > gpurun_out/bench_run25_cfg3p.json 2> gpurun_out/bench_run25_cfg3p.err
python - <<'P'
import json
for f in ('default','default_serial','cfg3p'):
    try:
        d=json.loads(open(f'gpurun_out/bench_run25_{f}.json').read().strip().splitlines()[-1])
        print(f, 'value', round(d['value'],1), d['unit'], 'ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'e2e ms', round(d['e2e']['ms_per_step'],3))
    except Exception as e:
        print(f, 'FAILED', e)
P
tail -n 3 gpurun_out/bench_run25_default.err; tail -n 3 gpurun_out/bench_run25_cfg3p.err
