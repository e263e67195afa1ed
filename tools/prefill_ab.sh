#!/bin/bash
# A/B of the prefill kernel variants on the B200 box (tuning build: python minivllm_b200/build.py --trace, done in the
# build container; libmvb200_trace.so travels).  usage: tools/prefill_ab.sh <tag>
set -u
TAG=${1:-prefill_ab}
O=gpurun_out; mkdir -p $O
# parity of the product library first
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_flash_attn_comparator.py -m gpu -q -x -k "prefill or layer_flow" > $O/${TAG}_pytest.txt 2>&1; echo "pytest rc $?" >> $O/${TAG}_pytest.txt; tail -3 $O/${TAG}_pytest.txt
: > $O/${TAG}.jsonl
run() { label=$1; shift
  env "$@" MVB200_LIB=minivllm_b200/libmvb200_trace.so timeout 300 python bench.py --no-cpu-baseline --steps 8 --warmup 3 --workload ${WL:-cfg3p} 2>>$O/${TAG}.err | python tools/bench_line.py $label >> $O/${TAG}.jsonl
}
run h1_e32 MVB200_PREFILL_HALVES=1 MVB200_PREFILL_EMU=32
run h2_e32 MVB200_PREFILL_HALVES=2 MVB200_PREFILL_EMU=32
run h2_e0 MVB200_PREFILL_HALVES=2 MVB200_PREFILL_EMU=0
run h2_e16 MVB200_PREFILL_HALVES=2 MVB200_PREFILL_EMU=16
run h2_e48 MVB200_PREFILL_HALVES=2 MVB200_PREFILL_EMU=48
run h2_e64 MVB200_PREFILL_HALVES=2 MVB200_PREFILL_EMU=64
WL=cfg3p1 run h1_e32_p1 MVB200_PREFILL_HALVES=1 MVB200_PREFILL_EMU=32
WL=cfg3p1 run h2_e32_p1 MVB200_PREFILL_HALVES=2 MVB200_PREFILL_EMU=32
cat $O/${TAG}.jsonl
MVB200_PREFILL_HALVES=2 timeout 300 python tools/prefill_trace.py > $O/${TAG}_trace_h2.log 2>&1
MVB200_PREFILL_HALVES=1 timeout 300 python tools/prefill_trace.py > $O/${TAG}_trace_h1.log 2>&1
head -50 $O/${TAG}_trace_h2.log
echo "prefill_ab done"
