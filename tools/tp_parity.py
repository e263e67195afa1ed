"""Token-stream parity of the reference engine + drop-in at tensor-parallel size N, on N GPUs (verdict r2, missing 1).

Launch:  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
             tools/tp_parity.py --tag qwen3-0.6b-untied.bfloat16 --out gpurun_out/tp_parity.json

Every rank builds the reference's own engine (``oracle/_ref``, unmodified ``LLM`` / scheduler / model / sampler) with
``minivllm_b200.install_into_reference(executor=True, fuse_qk=True, peer_allreduce=True)``; weights come from the
TP-aware ``oracle.reference_engine.seeded_load`` (a TP = N engine holds exactly the TP = 1 weights, sharded as
models/layers/linear.py:229-270 and models/qwen_base.py:95-108 prescribe).  Rank 0 then compares the 32-token greedy
streams of golden calls A and B (B hits the prefix cache) with ``tests/golden/engine_named.npz`` — the streams the
reference recorded on CPU at TP = 1 — under the rule of ``tests/test_reference_engine._compare`` (equal, or a
proven tie of the reference's own logits) and prints / writes one JSON record.

What this covers that the single-GPU tests do not: per-rank KV-head sharding of the cache (kv heads / N per rank), the
attention kernels on the per-rank head geometry inside the engine, graph capture on every rank, and the decode-size
all-reduce through the one-shot peer kernel, all inside ``LLM.generate``.
"""

from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = {
    # tag: (preset, dtype, config overrides, tie gap) — the rows of tests/test_reference_engine.CASES
    "qwen3-0.6b.float32": ("qwen3-0.6b", "float32", {}, 2e-4),
    "qwen3-0.6b-untied.float32": ("qwen3-0.6b", "float32", {"tie_word_embeddings": False}, 2e-4),
    "qwen3-0.6b-untied.bfloat16": ("qwen3-0.6b", "bfloat16", {"tie_word_embeddings": False}, 0.0625),
}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--tag", default="qwen3-0.6b-untied.bfloat16", choices=sorted(CASES))
    ap.add_argument("--out", default="")
    ap.add_argument("--eager", action="store_true", help="no CUDA graphs (layer + executor rebinding only)")
    ap.add_argument("--nccl", action="store_true", help="leave the all-reduces on NCCL")
    a = ap.parse_args()
    preset, dtype, overrides, tie_gap = CASES[a.tag]

    import numpy as np
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    from harness import ref_engine

    z = np.load(os.path.join(ROOT, "tests", "golden", "engine_named.npz"))
    calls = []
    for call in ("a", "b"):
        n = int(z[f"{a.tag}.{call}.n"])
        calls.append([z[f"{a.tag}.{call}.prompt{i}"].tolist() for i in range(n)])
    info = ref_engine.run(
        preset, device="cuda", dtype=dtype, dropin=True, executor=True, fuse_qk=True, enforce_eager=a.eager,
        peer_allreduce=world > 1 and not a.nccl, calls=calls, max_tokens=32, record_logits=True,
        config_overrides=overrides, tensor_parallel_size=world, under_torchrun=world > 1,
        max_num_seqs=8, max_model_len=1024, num_kvcache_blocks=64)
    if not info.get("worker"):
        from tests.test_reference_engine import _compare

        rec = {"tag": a.tag, "tp": world, "graphs_captured": info["graphs_captured"],
               "kv_heads_per_rank": info["kv_heads_per_rank"], "attention_classes": info["attention_classes"],
               "kernels_seen": info["kernels_seen"], "allreduce": info.get("allreduce"),
               "cache_hits": [c["cache_hits"] for c in info["calls"]]}
        try:
            rec["parity"] = _compare(a.tag, info, tie_gap)
            rec["ok"] = True
        except AssertionError as e:
            rec["ok"] = False
            rec["error"] = str(e)[:2000]
        line = json.dumps(rec)
        print("TP_PARITY " + line, flush=True)
        if a.out:
            with open(a.out, "w") as f:
                f.write(line + "\n")
    sys.stdout.flush()
    sys.stderr.flush()
    try:
        torch.cuda.synchronize()
    except Exception:  # noqa: BLE001
        pass
    os._exit(0)  # captured graphs holding collectives make destroy_process_group() block (see bench.finish)


if __name__ == "__main__":
    main()
