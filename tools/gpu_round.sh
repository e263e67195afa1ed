#!/bin/bash
# One evidence run on the B200 box (under gpurun, one GPU): GPU tests, smoke, both bench arms, helper-kernel
# table, per-shape decode lines, ncu launch list + full captures.  Everything lands in gpurun_out/<tag>_*.
# usage: tools/gpu_round.sh <tag> [stages...]
#   stages: tests smoke bench ref aux shapes decode_sweep ncu_decode ncu_shapes ncu_store ncu_prefill
set -u
TAG=${1:-run}; shift || true
STAGES=${*:-tests smoke bench ref aux shapes ncu_decode ncu_shapes ncu_store}
O=gpurun_out
mkdir -p $O
has() { [[ " $STAGES " == *" $1 "* ]]; }
if has tests; then
  timeout 1500 python -m pytest tests -m gpu -q -x > $O/${TAG}_pytest_gpu.txt 2>&1; echo "pytest rc $?" >> $O/${TAG}_pytest_gpu.txt
  tail -3 $O/${TAG}_pytest_gpu.txt
fi
if has smoke; then
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.txt 2>&1; echo "smoke rc $?" >> $O/${TAG}_smoke.txt
  tail -2 $O/${TAG}_smoke.txt
fi
if has bench; then
  timeout 900 python bench.py > $O/${TAG}_bench_default.json 2> $O/${TAG}_bench_default.err; echo "bench rc $?"
  tail -c 600 $O/${TAG}_bench_default.err
fi
if has ref; then
  timeout 600 python bench.py --impl reference > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; echo "ref rc $?"
fi
if has aux; then
  timeout 300 python tools/bench_aux_kernels.py > $O/${TAG}_aux_kernels.txt 2>&1; echo "aux rc $?"
fi
if has shapes; then
  : > $O/${TAG}_shapes.jsonl
  for wl in cfg2r cfg2tp8 cfg3d cfg4 cfg5d; do
    timeout 300 python bench.py --no-cpu-baseline --steps 20 --warmup 3 --workload $wl 2>>$O/${TAG}_shapes.err | python tools/bench_line.py $wl >> $O/${TAG}_shapes.jsonl
  done
  for wl in cfg3p cfg3p1; do
    timeout 300 python bench.py --no-cpu-baseline --steps 8 --warmup 3 --workload $wl 2>>$O/${TAG}_shapes.err | python tools/bench_line.py $wl >> $O/${TAG}_shapes.jsonl
  done
  cat $O/${TAG}_shapes.jsonl
fi
if has decode_sweep; then
  : > $O/${TAG}_decode_sweep.jsonl
  sw() { label=$1; wl=$2; shift 2
    # the MVB200_DECODE_* knobs exist in the tuning build only (python minivllm_b200/build.py --trace)
    env "$@" MVB200_LIB=minivllm_b200/libmvb200_trace.so timeout 300 python bench.py --no-cpu-baseline --steps 20 --warmup 3 --workload $wl ${EXTRA:-} 2>>$O/${TAG}_sweep.err | python tools/bench_line.py $label >> $O/${TAG}_decode_sweep.jsonl
  }
  sw cfg4_default cfg4 A=1
  sw cfg4_noearly cfg4 MVB200_DECODE_NO_EARLY=1
  EXTRA="--splits 37" sw cfg4_s37 cfg4 A=1
  EXTRA="--splits 37" sw cfg4_s37_1cta cfg4 MVB200_DECODE_CTAS_PER_SM=1
  EXTRA="--splits 37" sw cfg4_s37_1cta_4st cfg4 MVB200_DECODE_CTAS_PER_SM=1 MVB200_DECODE_STAGES=4
  sw cfg4_kernel_reduce cfg4 MVB200_DECODE_REDUCE=kernel
  sw cfg3d_default cfg3d A=1
  sw cfg3d_noearly cfg3d MVB200_DECODE_NO_EARLY=1
  EXTRA="--splits 2" sw cfg3d_s2 cfg3d A=1
  EXTRA="--splits 2" sw cfg3d_s2_1cta cfg3d MVB200_DECODE_CTAS_PER_SM=1
  EXTRA="--splits 1" sw cfg3d_s1 cfg3d A=1
  sw cfg2tp8_default cfg2tp8 A=1
  sw cfg2tp8_noearly cfg2tp8 MVB200_DECODE_NO_EARLY=1
  sw cfg5d_default cfg5d A=1
  sw cfg2_default cfg2 A=1
  cat $O/${TAG}_decode_sweep.jsonl
fi
NCU="ncu --clock-control none"
if has ncu_decode; then
  CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-prefill --no-engine"
  $CMD > $O/${TAG}_ncu_plain.log 2>&1 &&
  $NCU --metrics gpu__time_duration.sum -c 700 --csv --log-file $O/${TAG}_decode_launches.csv $CMD > $O/${TAG}_ncu_launches.log 2>&1
  $NCU --set full --import-source on -k regex:paged_decode_kernel -s 60 -c 2 -f -o $O/${TAG}_prof_decode_cfg2 $CMD > $O/${TAG}_ncu_full.log 2>&1
  ncu -i $O/${TAG}_prof_decode_cfg2.ncu-rep --page raw --csv > $O/${TAG}_decode_cfg2_ncu_full.csv 2>/dev/null
fi
if has ncu_shapes; then
  for wl in cfg4 cfg2tp8 cfg3d; do
    CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload $wl"
    $NCU --set full --import-source on -k regex:paged_decode_kernel -s 60 -c 2 -f -o $O/${TAG}_prof_decode_$wl $CMD > $O/${TAG}_ncu_$wl.log 2>&1
    ncu -i $O/${TAG}_prof_decode_$wl.ncu-rep --page raw --csv > $O/${TAG}_decode_${wl}_ncu_full.csv 2>/dev/null
  done
fi
if has ncu_store; then
  CMD="python tools/bench_aux_kernels.py"
  $NCU --set full --import-source on -k regex:store_kvcache -c 12 -f -o $O/${TAG}_prof_store $CMD > $O/${TAG}_ncu_store.log 2>&1
  ncu -i $O/${TAG}_prof_store.ncu-rep --page raw --csv > $O/${TAG}_store_ncu_full.csv 2>/dev/null
fi
if has ncu_prefill; then
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --workload cfg3p"
  $CMD > $O/${TAG}_ncu_prefill_plain.log 2>&1 &&
  $NCU --set full --import-source on -k regex:prefill_ -s 8 -c 2 -f -o $O/${TAG}_prof_prefill $CMD > $O/${TAG}_ncu_prefill.log 2>&1
  ncu -i $O/${TAG}_prof_prefill.ncu-rep --page raw --csv > $O/${TAG}_prefill_ncu_full.csv 2>/dev/null
fi
ls -la $O | tail -40
echo "gpu_round $TAG done"
