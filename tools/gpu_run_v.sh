#!/bin/bash
# 1 GPU: vectorised qk_norm_rope_store path: parity tests, engine token streams, bandwidth table, cfg3 e2e
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_qk_rope.py tests/test_harness.py tests/test_decode_meta.py -m gpu -x -q > gpurun_out/pytest_rope.txt 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_rope.txt
timeout 300 python tools/bench_aux_kernels.py > gpurun_out/aux_kernels.txt 2> gpurun_out/aux_kernels.err; echo "aux exit $?"; grep qk_norm gpurun_out/aux_kernels.txt | cut -c1-200
MVB200_ROPE_PAIR=1 timeout 300 python tools/bench_aux_kernels.py 2>/dev/null | grep qk_norm | cut -c1-200
timeout 600 python tools/e2e_bench.py --model qwen3-4b --seqs 8 --prompt 2048 --cached 1024 --new-tokens 256 > gpurun_out/e2e_cfg3_tp1.json 2> gpurun_out/e2e_cfg3_tp1.err; echo "cfg3 exit $?"
grep -h -o '"value": [0-9.]*\|"decode_ms_per_step": [0-9.]*\|"prefill_tokens_per_s": [0-9.]*' gpurun_out/e2e_cfg3_tp1.json | tr '\n' ' '; echo
