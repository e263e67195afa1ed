"""Host-side logic of bench.py that needs no GPU."""

import json
import os

import bench


def test_split_rounds_times_exactly_the_requested_steps():
    for steps in (1, 2, 3, 4, 7, 20, 21):
        for rounds in (1, 2, 4):
            per = bench.split_rounds(steps, rounds)
            assert sum(per) == steps and all(n >= 1 for n in per) and len(per) == min(rounds, steps)


def test_ragged_context_lengths_are_seeded_and_in_range():
    wl = bench.WORKLOADS["cfg2r"]
    a, b = bench.context_lengths(wl), bench.context_lengths(wl)
    assert a == b and len(a) == wl["batch"]
    assert min(a) >= 1024 and max(a) <= 3072 and len(set(a)) > 100
    assert bench.context_lengths(bench.WORKLOADS["cfg2"]) == [2048] * 256


def test_kv_bytes_match_survey_figures():
    # cfg2 2.147 GB per layer, cfg4 67.1 MB per layer
    assert bench.kv_bytes_per_layer(256 * 2048, 8, 128) == 2147483648
    assert bench.kv_bytes_per_layer(32768, 4, 128) == 67108864
    assert bench.prefill_flops_per_layer(bench.WORKLOADS["cfg3p"], 32) == 206225539072  # 2.062e11


def test_traffic_is_reported_only_for_the_captured_shape():
    rows = json.load(open(os.path.join(bench.ROOT, "profiles", "decode_traffic.json")))
    rows = rows if isinstance(rows, list) else [rows]
    r = rows[0]
    sh = r["shape"]
    assert bench.matching_traffic(r["workload_key"], sh["kv_heads_per_rank"], sh["batch"], sh["ctx_sum"]) == r["dram_bytes_per_launch"]
    # one rank of `--gpus 8 --workload cfg2` launches the shape captured as cfg2tp8: matched by shape, not by name
    tp8 = [x for x in rows if x["workload_key"] == "cfg2tp8"]
    if tp8:
        assert bench.matching_traffic("cfg2", 1, 256, 524288, 2) == tp8[0]["dram_bytes_per_launch"]
    assert bench.matching_traffic(r["workload_key"], 3, sh["batch"], sh["ctx_sum"]) is None     # a shape never captured
    assert bench.matching_traffic(r["workload_key"], sh["kv_heads_per_rank"], sh["batch"], sh["ctx_sum"], 999) is None
    assert bench.matching_traffic("cfg4", 4, 1, 32768) is None or any(x["workload_key"] == "cfg4" for x in rows)


def test_prefill_traffic_is_matched_on_the_captured_shape():
    wl = bench.WORKLOADS["cfg3p"]
    t = bench.matching_prefill_traffic(wl["seqs"], wl["cached"], wl["new"], wl["H"], wl["KVH"], wl["D"])
    assert t is not None and 1.0 < t / (8 * 1024 * 32 * 128 * 2 * 2 + 8 * 2048 * 8 * 128 * 2 * 2) < 2.0  # 1.5x algorithmic (K/V re-streamed per head)
    assert bench.matching_prefill_traffic(wl["seqs"], wl["cached"], wl["new"], wl["H"] // 2, wl["KVH"] // 2, wl["D"]) is None


def test_engine_prompts_share_the_cached_prefix():
    wl = bench.WORKLOADS["cfg3e"]
    warm, real = bench.engine_prompts(wl, 151936)
    assert len(warm) == len(real) == 8
    for w, r in zip(warm, real):
        assert len(r) == 2048 and len(w) == 1025 and w[:1024] == r[:1024]
    cfg = bench.engine_config(wl)
    assert cfg["max_num_batched_tokens"] >= 8 * 1024 and cfg["max_model_len"] >= 2048 + 256
    assert cfg["num_kvcache_blocks"] >= 8 * ((2048 + 256) // 64 + 1)
