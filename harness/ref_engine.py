"""Run the reference's own engine (``minivllm.LLM.generate``) — optionally with the B200 layer swapped in.

This is the drop-in boundary exercised for real: the reference's
``LLM`` / ``LLMEngine`` / ``Scheduler`` / ``BlockManager`` / ``ModelRunner`` /
``InferenceExecutor`` / model definitions / sampler all run unmodified (imported from
``/root/reference`` or its copy under ``oracle/_ref``); ``minivllm_b200.install_into_reference()``
replaces the ``Attention`` layer (and, with ``executor=True``, the executor's metadata
preparation / cache sizing / CUDA-Graph capture, the three subsystems north_star says change).

Arms:
  device=cpu                      the reference's eager path on the host cores (the oracle / CPU baseline)
  device=cuda, dropin=True        reference engine + sm_100a kernels (the product as a user would run it)
  device=cuda, dropin=False       the reference's own CUDA path (flash-attn / Triton) if the wheels run on sm_100

Harness code: imports ``oracle.reference_engine`` for the loader/tokenizer/config patches.
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def make_prompts(vocab: int, lens: list[int], seed: int = 600) -> list[list[int]]:
    import torch

    g = torch.Generator().manual_seed(seed)
    return [torch.randint(10, vocab - 10, (n,), generator=g).tolist() for n in lens]


def run(
    preset: str,
    *,
    device: str = "cpu",
    dtype: str = "float32",
    dropin: bool = False,
    executor: bool = False,
    calls: list[list[list[int]]],
    max_tokens: int | list[int] = 32,
    enforce_eager: bool = True,
    fast_init: bool = False,
    record_logits: bool = False,
    hide_flash_attn: bool | None = None,
    threads: int | None = None,
    config_overrides: dict | None = None,
    tensor_parallel_size: int = 1,
    under_torchrun: bool = False,
    isolate_cuda: bool = True,
    profile: str = "",
    fuse_qk: bool = False,
    peer_allreduce: bool = False,
    **config_kw,
) -> dict:
    """Build one reference ``LLM`` and run ``generate`` once per entry of ``calls`` (later calls can hit the
    prefix cache of earlier ones, block_manager.py:252-339).  Returns tokens, per-step timings and counters."""
    from oracle import reference_engine as R

    if hide_flash_attn is None:
        hide_flash_attn = device == "cpu" or dropin
    R.import_reference(device, hide_flash_attn=hide_flash_attn, isolate_cuda=isolate_cuda)
    import torch

    if threads:
        torch.set_num_threads(threads)
    info: dict = {"preset": preset, "device": device, "dtype": dtype, "dropin": dropin, "executor": executor,
                  "enforce_eager": enforce_eager, "reference_root": R.locate_reference()}
    if dropin:
        import minivllm_b200

        minivllm_b200.install_into_reference(executor=executor, fuse_qk=fuse_qk, peer_allreduce=peer_allreduce)
        info["native_lib"] = minivllm_b200._native.LIB_PATH
    elif executor:  # executor subsystems only, the reference's own attention layer (CPU-testable)
        from minivllm_b200 import reference_patch

        reference_patch.install_executor()
    local = int(os.environ.get("LOCAL_RANK", "0")) if under_torchrun else 0
    rank = int(os.environ.get("RANK", "0")) if under_torchrun else 0
    loader = R.fast_device_load(f"cuda:{local}" if device == "cuda" else "cpu", dtype, seed=1234 + rank) if fast_init else R.seeded_load
    t0 = time.perf_counter()
    if tensor_parallel_size > 1:
        # One process per GPU already exists (torchrun): the reference's LLMEngine must not spawn its own workers
        # (llm_engine.py:82-100) and its ModelRunner must not tear down a process group it did not create.
        config_kw["tensor_parallel_size"] = tensor_parallel_size
        if not under_torchrun:
            raise ValueError("tensor_parallel_size > 1 is driven from torchrun ranks (under_torchrun=True)")
        worker = _tp_setup(rank)
        if worker:
            from minivllm.config import Config
            from minivllm.engine.model_runner import ModelRunner

            path = R.write_model_dir(preset, dtype, **(config_overrides or {}))
            with R.engine_patches(loader):
                # blocks in ModelRunner._worker_loop (model_runner.py:124-171) until rank 0 broadcasts "exit"
                ModelRunner(Config(model=path, dtype=dtype, enforce_eager=enforce_eager, **config_kw), rank, _NoSpawn().Event())
            return {"worker": True, "rank": rank}
        from unittest import mock

        import minivllm.engine.llm_engine as llm_engine

        with mock.patch.object(llm_engine.mp, "get_context", lambda *_a, **_k: _NoSpawn()):
            llm = R.make_llm(preset, dtype=dtype, loader=loader, enforce_eager=enforce_eager,
                             config_overrides=config_overrides, **config_kw)
    else:
        llm = R.make_llm(preset, dtype=dtype, loader=loader, enforce_eager=enforce_eager,
                         config_overrides=config_overrides, **config_kw)
    info["init_s"] = time.perf_counter() - t0
    attn_cls = {type(m).__module__ + "." + type(m).__name__ for m in llm.model_runner.model_manager.model.modules()
                if hasattr(m, "k_cache")}
    info["attention_classes"] = sorted(attn_cls)
    sync = (lambda: torch.cuda.synchronize()) if device == "cuda" else (lambda: None)

    steps: list[tuple[int, float]] = []
    inner_step = llm.step

    def timed_step():
        sync()
        t = time.perf_counter()
        out = inner_step()
        sync()
        steps.append((out[1], time.perf_counter() - t))
        return out

    llm.step = timed_step
    # top-2 logits of every sampled row, in batch order (stronger than token ids: a near-tie is visible)
    top_log: list = []
    if record_logits:
        ex0 = llm.model_runner.inference_executor
        inner_sample = ex0._sample_tokens

        def sampling(logits, sequences):
            v, i = logits.float().topk(2, dim=-1)
            top_log.append(([s.seq_id for s in sequences], v.cpu().tolist(), i.cpu().tolist()))
            return inner_sample(logits, sequences)

        ex0._sample_tokens = sampling
    results = []
    for ci, prompts in enumerate(calls):
        sp = R.greedy_params(max_tokens[ci] if isinstance(max_tokens, (list, tuple)) else max_tokens)
        steps.clear()
        top_log.clear()
        hits0 = llm.scheduler.block_manager.stats.get("cache_hits", 0)
        t0 = time.perf_counter()
        if profile and ci == len(calls) - 1:
            from torch.profiler import ProfilerActivity, profile as torch_profile

            with torch_profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
                res = llm.generate(prompts, sp, use_tqdm=False)
            with open(profile, "w") as f:
                f.write(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=80))
                f.write("\n\n" + prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=30, max_name_column_width=80))
        else:
            res = llm.generate(prompts, sp, use_tqdm=False)
        sync()
        wall = time.perf_counter() - t0
        pre = [(n, t) for n, t in steps if n > 0]
        dec = [(-n, t) for n, t in steps if n < 0]
        results.append({
            "tokens": [r["token_ids"] for r in res],
            "wall_s": wall,
            "cache_hits": llm.scheduler.block_manager.stats.get("cache_hits", 0) - hits0,
            "prefill_tokens": sum(n for n, _ in pre), "prefill_s": sum(t for _, t in pre),
            "decode_tokens": sum(n for n, _ in dec), "decode_s": sum(t for _, t in dec),
            "decode_steps": len(dec),
        })
        if record_logits:
            # regroup per sequence in request order: top2[p][t] = (v1, v2, id1, id2)
            ids = sorted({sid for sids, _, _ in top_log for sid in sids})
            per = {sid: [] for sid in ids}
            for sids, vals, idx in top_log:
                for sid, v, i in zip(sids, vals, idx):
                    per[sid].append((v[0], v[1], i[0], i[1]))
            results[-1]["top2"] = [per[sid] for sid in ids]
    info["calls"] = results
    if dropin:
        import minivllm_b200

        info["last_kernel"] = minivllm_b200.ops.last_kernel()
        info["kernels_seen"] = sorted(minivllm_b200.ops.kernels_seen())
        if peer_allreduce:
            from minivllm_b200 import reference_patch

            info["allreduce"] = reference_patch.peer_allreduce_status()
    ex = llm.model_runner.inference_executor
    info["graphs_captured"] = sorted(getattr(ex, "graphs", {}).keys())
    info["kv_heads_per_rank"] = int(ex.kv_cache.size(4)) if ex.kv_cache is not None else None
    try:
        llm.exit()
    except Exception as e:  # the reference's cleanup is best-effort
        info["exit_error"] = repr(e)
    return info


class _NoSpawn:
    """Stands in for ``mp.get_context("spawn")`` in LLMEngine.__init__ when torchrun already started one
    process per GPU: no worker is spawned, ``exit()`` finds nothing to join."""

    class _Proc:
        def start(self):
            pass

        def join(self, timeout=None):
            pass

        def is_alive(self):
            return False

        def terminate(self):
            pass

        kill = terminate

    def Event(self):
        import threading

        return threading.Event()

    def Process(self, target=None, args=(), kwargs=None):
        return self._Proc()


def _tp_setup(rank: int) -> bool:
    """Harness-level fixes that let the reference's TP run inside torchrun ranks.  Returns True on workers."""
    import torch.distributed as dist

    import minivllm.engine.distributed_manager as dm

    if not dist.is_initialized():
        raise RuntimeError("under_torchrun: init_process_group first (the model must see the group, SURVEY 5 defect 1)")

    def cleanup(self):  # the launcher owns the process group; destroying it under live NCCL graphs blocks
        if self.is_distributed and dist.is_initialized():
            dist.barrier()
        self._initialized = False

    dm.DistributedManager.cleanup = cleanup
    return rank != 0


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="qwen3-0.6b")
    ap.add_argument("--device", default="cpu")
    ap.add_argument("--dtype", default="float32")
    ap.add_argument("--dropin", action="store_true")
    ap.add_argument("--executor", action="store_true")
    ap.add_argument("--graphs", action="store_true")
    ap.add_argument("--fast-init", action="store_true")
    ap.add_argument("--lens", default="165,13,203,64,250")
    ap.add_argument("--repeat-call", type=int, default=1, help="issue the same prompts this many times (2nd+ hit the prefix cache)")
    ap.add_argument("--max-tokens", type=int, default=32)
    ap.add_argument("--max-num-seqs", type=int, default=8)
    ap.add_argument("--max-model-len", type=int, default=1024)
    ap.add_argument("--max-num-batched-tokens", type=int, default=16384)
    ap.add_argument("--num-kvcache-blocks", type=int, default=64)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--golden", default="", help="npz with <preset>.<dtype>.prompt{i}/tokens{i}: compare instead of free-running")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    from oracle import reference_engine as R

    vocab = R.PRESETS[a.preset]["vocab_size"]
    prompts = make_prompts(vocab, [int(x) for x in a.lens.split(",")])
    info = run(a.preset, device=a.device, dtype=a.dtype, dropin=a.dropin, executor=a.executor,
               calls=[prompts] * a.repeat_call, max_tokens=a.max_tokens, enforce_eager=not a.graphs,
               fast_init=a.fast_init, threads=a.threads or None, max_num_seqs=a.max_num_seqs,
               max_model_len=a.max_model_len, max_num_batched_tokens=a.max_num_batched_tokens,
               num_kvcache_blocks=a.num_kvcache_blocks)
    slim = {k: v for k, v in info.items() if k != "calls"}
    slim["calls"] = [{k: v for k, v in c.items() if k != "tokens"} | {"first_tokens": [t[:6] for t in c["tokens"]]}
                     for c in info["calls"]]
    print(json.dumps(slim))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(info, f)


if __name__ == "__main__":
    main()
