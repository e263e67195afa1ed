#!/usr/bin/env python
"""bench.py — the attention hot path on BASELINE.json's headline configuration.

BASELINE.json's metric has three parts; the default run reports all three in ONE JSON line:

1. ``value`` / ``roofline`` / ``e2e`` — cfg2 (``configs[1]``): Qwen3-0.6B decode, 256 sequences x 2048-token
   context, bf16, CUDA Graph.  One *step* = one decode step's attention hot path over the whole batch: for each
   of the 28 layers ``store_kvcache`` of the 256 new K/V rows, then paged split-KV GQA decode attention over
   that layer's own paged cache (62 GB of KV in HBM at N=1: every step streams inputs far larger than L2).
     value    = KV-read GB/s, device-timed, inputs resident in HBM, the 28-layer step replayed as one CUDA graph
     e2e      = same metric through ``Attention.forward`` with HOST (pinned) q/k/v per layer: the H2D copies of
                inputs + metadata and the D2H copy of every layer's output are memcpy nodes of ONE captured
                graph (copy-in of layer l+1 / copy-out of layer l-1 on side branches), all inside the timed region
     roofline = ``paged_decode_kernel`` timed alone (decode-only graph, interleaved with the full-step graph)
                against the measured HBM peak
2. ``prefill`` — cfg3's prefill (Qwen3-4B: 8 sequences x (1024 cached + 1024 new tokens), 50 % prefix hits):
   TFLOP/s of store + prefix-aware varlen prefill attention, tensor roofline, e2e with host buffers.
3. ``qwen3_4b_tok_s`` — cfg3 end to end through the REFERENCE's own engine (``minivllm.LLM.generate`` from
   ``oracle/_ref``: its scheduler, block manager, model runner, Qwen3 model, sampler) with
   ``minivllm_b200.install_into_reference(executor=True)``: 8 x 2048-token prompts, 50 % prefix hits, 256 decode
   tokens, CUDA Graph, TP = N.

N > 1 = the reference's tensor-parallel split: KV/Q heads shard over ranks, no collective in the attention path
(NCCL only for the barrier and the max-over-ranks of the time; the engine leg adds the reference's all-reduce).

``--impl reference`` times the reference's own CPU implementation (``oracle/_ref``: ``PageAttention.__call__``
page_attention.py:35-133 + ``StandardAttentionBackend.store_kv_cache`` attention_backend.py:237-266) on the
host cores, on a bounded sample of the same workload.  ``--workload`` selects a single leg / another config.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: layers, H, KVH, D, batch, ctx
    "cfg2": dict(name="Qwen3-0.6B decode 256 seqs x 2048 ctx, bf16, CUDA Graph", layers=28, H=16, KVH=8, D=128, batch=256, ctx=2048),
    # the ragged variant of cfg2, context lengths ~ U[1024, 3072] (seeded), same mean
    "cfg2r": dict(name="Qwen3-0.6B decode 256 seqs x U[1024,3072] ctx (ragged), bf16, CUDA Graph", layers=28, H=16, KVH=8, D=128,
                  batch=256, ctx=2048, ragged=(1024, 3072)),
    "cfg4": dict(name="Qwen2-7B decode 1 seq x 32768 ctx (split-KV stress), bf16, CUDA Graph", layers=28, H=28, KVH=4, D=128, batch=1, ctx=32768),
    "cfg5d": dict(name="Qwen2-72B TP8-rank decode 512 seqs x 1024 ctx, bf16, CUDA Graph", layers=80, H=8, KVH=1, D=128, batch=512, ctx=1024),
    # per-rank shapes of the TP configurations, runnable on one GPU (what one of N ranks computes)
    "cfg2tp8": dict(name="Qwen3-0.6B decode 256 seqs x 2048 ctx, ONE rank of TP8 (KVH 1), bf16, CUDA Graph", layers=28, H=2, KVH=1, D=128, batch=256, ctx=2048),
    "cfg3d": dict(name="Qwen3-4B decode 8 seqs x 2304 ctx, bf16, CUDA Graph", layers=36, H=32, KVH=8, D=128, batch=8, ctx=2304),
    "tiny": dict(name="debug", layers=2, H=16, KVH=8, D=128, batch=8, ctx=512),
    # prefill workloads (metric: prefill attention TFLOP/s)
    "cfg3p": dict(name="Qwen3-4B prefill 8 seqs x (1024 cached + 1024 new), bf16, 50% prefix hits", kind="prefill",
                  layers=36, H=32, KVH=8, D=128, seqs=8, cached=1024, new=1024),
    "cfg3p1": dict(name="Qwen3-4B prefill 1 seq x (4096 cached + 4096 new), bf16, 50% prefix hits", kind="prefill",
                   layers=36, H=32, KVH=8, D=128, seqs=1, cached=4096, new=4096),
    "tinyp": dict(name="debug prefill", kind="prefill", layers=2, H=8, KVH=2, D=128, seqs=2, cached=128, new=256),
    # engine workloads (metric: end-to-end decode tokens/s through the reference's LLM.generate)
    "cfg3e": dict(name="Qwen3-4B engine: 8 prompts x 2048 tokens (1024 prefix-cache hits each), 256 new tokens, CUDA Graph",
                  kind="engine", preset="qwen3-4b", seqs=8, prompt=2048, cached=1024, new_tokens=256),
    "tinye": dict(name="debug engine", kind="engine", preset="tiny-qwen3", seqs=4, prompt=256, cached=128, new_tokens=16),
}
BLOCK = 64
METRIC = "paged-attn decode KV-read GB/s"


def kv_bytes_per_layer(ctx_sum, kvh, d, elem=2):
    """sum_b ctx_b * KVH * D * 2 (K and V) * elem."""
    return ctx_sum * kvh * d * 2 * elem


def context_lengths(wl, seed=0):
    """Per-sequence context lengths of a decode workload (uniform, or seeded U[lo, hi] for the ragged variant)."""
    import torch

    if "ragged" not in wl:
        return [wl["ctx"]] * wl["batch"]
    lo, hi = wl["ragged"]
    g = torch.Generator().manual_seed(seed)
    return torch.randint(lo, hi + 1, (wl["batch"],), generator=g).tolist()


# ---------------------------------------------------------------------------
# process environment (one process per GPU)
# ---------------------------------------------------------------------------


class Env:
    def __init__(self):
        import torch
        import torch.distributed as dist

        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device (no CPU fallback for the product path)")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1 and not dist.is_initialized():
            dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, ms: float) -> float:
        if self.world > 1:
            t = self.torch.tensor([ms], device=self.dev)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def timed(self, fn, steps, warmup, before_each=None):
        """EXACTLY ``steps`` timed calls of ``fn`` bracketed by barrier + synchronize, CUDA events on the launch
        stream, max over ranks.  ``before_each`` (e.g. an L2 flush) runs untimed before every call."""
        torch = self.torch
        for _ in range(warmup):
            fn()
        self.barrier()
        if before_each is None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
        else:
            ms = 0.0
            for _ in range(steps):
                before_each()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                torch.cuda.synchronize()
                ms += e0.elapsed_time(e1)
        ms = self.max_over_ranks(ms)
        self.barrier()
        return ms / steps

    def timed_interleaved(self, fns: dict, steps, warmup, rounds=4):
        """Time several step functions in alternating rounds (A B A B ...) so that a drifting power cap hits
        all of them alike; every function is timed for exactly ``steps`` calls in total."""
        rounds = max(1, min(rounds, steps))
        per = [steps // rounds + (1 if r < steps % rounds else 0) for r in range(rounds)]
        total = {k: 0.0 for k in fns}
        for r, n in enumerate(per):
            for k, fn in fns.items():
                total[k] += self.timed(fn, n, warmup if r == 0 else 1) * n
        return {k: v / steps for k, v in total.items()}


def split_rounds(steps, rounds):
    rounds = max(1, min(rounds, steps))
    return [steps // rounds + (1 if r < steps % rounds else 0) for r in range(rounds)]


# ---------------------------------------------------------------------------
# the e2e leg: host buffers through Attention.forward, the whole step as ONE CUDA graph
# ---------------------------------------------------------------------------


class HostPipeline:
    """Per-layer loop over HOST buffers: while layer ``l`` computes on the main branch, layer ``l+1``'s pinned-host
    q/k/v are copied in on a second stream (two device buffer sets) and layer ``l``'s result goes back to pinned
    host memory on a third.  ``capture()`` records that schedule — memcpy nodes included — into one CUDA graph
    (fork/join of the side streams inside the capture become graph edges), so a step costs the host one
    ``cudaGraphLaunch`` and the N >= 2 figure is no longer bound by eager per-layer Python.  Every copy of every
    step executes inside the timed region."""

    def __init__(self, torch, like):
        self.torch = torch
        self.h2d, self.d2h = torch.cuda.Stream(), torch.cuda.Stream()
        self.bufs = [tuple(torch.empty_like(t) for t in like) for _ in range(2)]
        self.ready = [torch.cuda.Event() for _ in range(2)]
        self.consumed = [torch.cuda.Event() for _ in range(2)]
        self.graph = None

    def _copy_in(self, l, host_in):
        i = l & 1
        with self.torch.cuda.stream(self.h2d):
            if l >= 2:
                self.h2d.wait_event(self.consumed[i])  # layer l-2 has read this buffer set
            for dst, src in zip(self.bufs[i], host_in(l)):
                dst.copy_(src, non_blocking=True)
            self.ready[i].record(self.h2d)

    def run(self, layers, host_in, host_out, prologue=None):
        torch = self.torch
        cur = torch.cuda.current_stream()
        if prologue is not None:
            prologue()  # metadata H2D copies, on the main branch
        self.h2d.wait_stream(cur)  # nothing of this step starts before the step (and its start event) does
        self.d2h.wait_stream(cur)
        self._copy_in(0, host_in)
        for l, a in enumerate(layers):
            i = l & 1
            if l + 1 < len(layers):
                self._copy_in(l + 1, host_in)
            cur.wait_event(self.ready[i])
            o = a(*self.bufs[i])
            self.consumed[i].record(cur)
            with torch.cuda.stream(self.d2h):
                self.d2h.wait_event(self.consumed[i])
                host_out(l).copy_(o, non_blocking=True)
            o.record_stream(self.d2h)
        cur.wait_stream(self.h2d)  # the step ends (and its end event fires) after the last copy out
        cur.wait_stream(self.d2h)

    def capture(self, layers, host_in, host_out, prologue=None):
        torch = self.torch
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            self.run(layers, host_in, host_out, prologue)  # warm-up, eager
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.run(layers, host_in, host_out, prologue)
        return self.graph


E2E_API = ("minivllm_b200.Attention.forward per layer, pinned host q/k/v in, pinned host out; the whole step "
           "(memcpy nodes + kernels, copy-in of layer l+1 and copy-out of layer l-1 on side branches) is one CUDA graph")


class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.samples = []
        self._stop = threading.Event()
        self._t = None

    def _run(self):
        while not self._stop.is_set():
            try:
                r = subprocess.run(
                    ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits"],
                    capture_output=True, text=True, timeout=5,
                )
                if r.returncode == 0 and r.stdout.strip():
                    self.samples.append([x.strip() for x in r.stdout.strip().splitlines()[0].split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[0]))
                mx = max(mx, float(s[1]))
            except (ValueError, IndexError):
                continue
            for nm, val in zip(names, s[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def matching_traffic(wl_key, kvh_rank, batch, ctx_sum, q_heads_rank=None):
    """``roofline.traffic``: the ncu ``dram__bytes_read + dram__bytes_write`` per launch — only when a capture of
    exactly this PER-RANK launch shape is on file (profiles/decode_traffic.json), else None.  The workload name is not
    part of the match: one rank of ``--gpus 8 --workload cfg2`` launches the shape captured as ``cfg2tp8`` (verdict
    r2, 5c); ``wl_key`` only breaks ties between captures of the same shape."""
    try:
        rows = json.load(open(os.path.join(ROOT, "profiles", "decode_traffic.json")))
    except Exception:
        return None
    rows = rows if isinstance(rows, list) else [rows]
    hits = []
    for r in rows:
        shape = r.get("shape", {})
        if shape.get("kv_heads_per_rank") == kvh_rank and shape.get("batch") == batch and shape.get("ctx_sum") == ctx_sum \
                and (q_heads_rank is None or shape.get("q_heads_per_rank") in (None, q_heads_rank)):
            hits.append(r)
    hits.sort(key=lambda r: r.get("workload_key") != wl_key)
    return hits[0].get("dram_bytes_per_launch") if hits else None


def matching_prefill_traffic(seqs, cached, new, q_heads_rank, kvh_rank, head_dim):
    """The prefill leg's ``roofline.traffic``: ncu DRAM bytes per launch of the prefill kernel when a capture of
    exactly this per-rank shape is on file (profiles/prefill_traffic.json), else None."""
    try:
        rows = json.load(open(os.path.join(ROOT, "profiles", "prefill_traffic.json")))
    except Exception:
        return None
    want = {"sequences": seqs, "cached_tokens": cached, "new_tokens": new, "q_heads_per_rank": q_heads_rank,
            "kv_heads_per_rank": kvh_rank, "head_dim": head_dim}
    for r in rows if isinstance(rows, list) else [rows]:
        if r.get("shape") == want:
            return r.get("dram_bytes_per_launch")
    return None


# ---------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's own CPU eager path on a bounded sample
# ---------------------------------------------------------------------------


def reference_cpu_ops():
    """(store_fn, decode_fn, kind): the reference's own classes from oracle/_ref (or /root/reference), else the
    torch-CPU port (``oracle/eager_port.py``) with ``kind = 'port'``."""
    from oracle import reference_engine as R

    if R.locate_reference() is not None:
        try:
            R.import_reference("cpu", hide_flash_attn=True, isolate_cuda=False)
            from minivllm.models.layers.attention_backend import StandardAttentionBackend
            from minivllm.models.layers.page_attention import PageAttention

            backend, page = StandardAttentionBackend(), PageAttention()
            return backend.store_kv_cache, page, "reference", R.locate_reference()
        except Exception as e:  # noqa: BLE001
            sys.stderr.write(f"bench: reference import failed ({e!r}); timing the port instead\n")
    from oracle import eager_port as port

    return port.store_kvcache, port.paged_decode, "port", "oracle/eager_port.py"


def cpu_sample_setup(wl, sample_batch, sample_layers, seed=0):
    import torch

    g = torch.Generator().manual_seed(seed)
    kvh, h, d = wl["KVH"], wl["H"], wl["D"]
    lens = context_lengths(wl)[:sample_batch]
    nblk = max(lens) // BLOCK + 1
    nb = sample_batch * nblk
    caches = []
    for _ in range(sample_layers):
        kc = torch.randn(nb, BLOCK, kvh, d, generator=g).to(torch.bfloat16)
        vc = torch.randn(nb, BLOCK, kvh, d, generator=g).to(torch.bfloat16)
        caches.append((kc, vc))
    perm = torch.randperm(nb, generator=g).view(sample_batch, nblk).to(torch.int32)
    ctx_lens = torch.tensor(lens, dtype=torch.int32)
    last = (ctx_lens.long() - 1)
    slots = (perm[torch.arange(sample_batch), last // BLOCK].long() * BLOCK + last % BLOCK).to(torch.int32)
    q = torch.randn(sample_batch, h, d, generator=g).to(torch.bfloat16)
    k = torch.randn(sample_batch, kvh, d, generator=g).to(torch.bfloat16)
    v = torch.randn(sample_batch, kvh, d, generator=g).to(torch.bfloat16)
    return caches, perm, ctx_lens, slots, q, k, v, sum(lens)


def run_cpu_baseline(wl, steps, warmup, sample_batch=64, sample_layers=4):
    import torch

    store, decode, kind, where = reference_cpu_ops()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sample_batch = min(sample_batch, wl["batch"])
    sample_layers = min(sample_layers, wl["layers"])
    caches, perm, ctx_lens, slots, q, k, v, ctx_sum = cpu_sample_setup(wl, sample_batch, sample_layers)
    scale = wl["D"] ** -0.5

    def step():
        out = None
        with torch.inference_mode():
            for kc, vc in caches:
                store(k, v, kc, vc, slots)
                out = decode(q, kc, vc, perm, ctx_lens, scale)
        return out

    for _ in range(max(1, warmup)):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    nbytes = kv_bytes_per_layer(ctx_sum, wl["KVH"], wl["D"]) * sample_layers
    what = ("the reference's own StandardAttentionBackend.store_kv_cache + PageAttention.__call__ (page_attention.py:35-133), "
            f"unmodified, imported from {os.path.relpath(where, ROOT) if where.startswith(ROOT) else where}") if kind == "reference" \
        else "oracle/eager_port.py = torch-CPU port of page_attention.py:35-133"
    return {
        "value": nbytes / dt / 1e9, "unit": "GB/s", "cores": torch.get_num_threads(), "kind": kind,
        "sample": f"{sample_batch} of {wl['batch']} sequences, {sample_layers} of {wl['layers']} layers, bf16, "
                  f"store + paged decode per layer, {steps} steps; {what}",
        "ms_per_step": dt * 1e3, "seconds_timed": dt * steps,
    }


def run_reference_arm(args, wl):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    if wl.get("kind") == "prefill":
        base, metric, unit = run_prefill_cpu_baseline(wl, args.steps, args.warmup), "prefill attention TFLOP/s", "TFLOP/s"
    elif wl.get("kind") == "engine":
        base, metric, unit = run_engine_cpu_baseline(wl), "end-to-end decode tokens/s", "tokens/s"
    else:
        base, metric, unit = run_cpu_baseline(wl, args.steps, args.warmup), METRIC, "GB/s"
    line = {
        "impl": "reference", "metric": metric, "value": base["value"], "unit": unit, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": base["ms_per_step"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": wl["name"], "sample": base["sample"]},
        "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": base["value"], "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ---------------------------------------------------------------------------
# leg 1: paged decode (cfg2 and friends)
# ---------------------------------------------------------------------------


def run_decode(env: Env, args, wl, wl_key) -> dict:
    torch = env.torch
    import minivllm_b200 as m

    tp, rank, dev = env.world, env.rank, env.dev
    if wl["H"] % tp or wl["KVH"] % tp:
        raise SystemExit(f"TP={tp} does not divide H={wl['H']} / KVH={wl['KVH']} (reference rule qwen_base.py:95-104)")
    L, H, KVH, D = wl["layers"], wl["H"] // tp, wl["KVH"] // tp, wl["D"]
    B = wl["batch"]
    lens = context_lengths(wl)
    ctx_sum, ctx_max = sum(lens), max(lens)
    scale = D**-0.5
    nblk = ctx_max // BLOCK + 1
    NB = B * nblk
    torch.manual_seed(1234 + rank)

    # ---- paged cache [2, L, NB, 64, KVH/tp, D] (inference_executor.py:209-221), random contents
    kv_cache = torch.empty(2, L, NB, BLOCK, KVH, D, dtype=torch.bfloat16, device=dev)
    flat = kv_cache.view(-1)
    chunk = 1 << 28
    for i in range(0, flat.numel(), chunk):
        flat[i : i + chunk].normal_()
    g = torch.Generator().manual_seed(0)
    block_tables = torch.randperm(NB, generator=g).view(B, nblk).to(torch.int32).to(dev)  # non-contiguous pages
    context_lens = torch.tensor(lens, dtype=torch.int32, device=dev)
    last = context_lens.long() - 1
    slot_mapping = (block_tables[torch.arange(B, device=dev), last // BLOCK].long() * BLOCK + last % BLOCK).to(torch.int32)

    layers = []
    for l in range(L):
        a = m.Attention(H, D, scale, KVH)
        a.k_cache, a.v_cache = kv_cache[0, l], kv_cache[1, l]
        layers.append(a)

    q = torch.randn(B, H, D, device=dev).to(torch.bfloat16)
    k = torch.randn(B, KVH, D, device=dev).to(torch.bfloat16)
    v = torch.randn(B, KVH, D, device=dev).to(torch.bfloat16)
    splits = args.splits or m.ops.decode_num_splits(B, KVH, nblk * BLOCK, dev)
    ws_n = m.ops.decode_workspace_numel(B, H, D, splits)
    workspace = m.ops.decode_workspace(ws_n, dev)
    outs = torch.empty(L, B, H, D, dtype=torch.bfloat16, device=dev)

    def set_ctx(sl=slot_mapping, cl=context_lens, bt=block_tables):
        m.set_context(False, slot_mapping=sl, context_lens=cl, block_tables=bt,
                      decode_num_splits=splits, decode_workspace=workspace)

    def full_step():  # the hot path of one decode step: store + paged decode, every layer (one launch each)
        o = None
        for a in layers:
            o = a(q, k, v)  # the layer's own output tensor, as the model consumes it (no copy kernel in between)
        return o

    def decode_only():
        for l, a in enumerate(layers):
            m.paged_decode_attention(q, a.k_cache, a.v_cache, block_tables, context_lens, scale,
                                     num_splits=splits, workspace=workspace, out=outs[l], kv_prefix_stable=True)

    def capture(fn):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            fn()
        return gr

    set_ctx()
    seen0 = set(m.ops.kernels_seen())
    g_full = capture(full_step)
    g_dec = capture(decode_only)
    kernels = sorted(set(m.ops.kernels_seen()) - seen0) or sorted(m.ops.kernels_seen())

    # ---- e2e: host buffers through Attention.forward, every copy a node of the captured step
    hq = [torch.randn(B, H, D).to(torch.bfloat16).pin_memory() for _ in range(L)]
    hk = [torch.randn(B, KVH, D).to(torch.bfloat16).pin_memory() for _ in range(L)]
    hv = [torch.randn(B, KVH, D).to(torch.bfloat16).pin_memory() for _ in range(L)]
    hout = [torch.empty(B, H, D, dtype=torch.bfloat16).pin_memory() for _ in range(L)]
    h_slots, h_ctx, h_bt = slot_mapping.cpu().pin_memory(), context_lens.cpu().pin_memory(), block_tables.cpu().pin_memory()
    d_slots, d_ctx, d_bt = torch.empty_like(slot_mapping), torch.empty_like(context_lens), torch.empty_like(block_tables)
    pipe = HostPipeline(torch, (q, k, v))

    def meta_in():
        d_slots.copy_(h_slots, non_blocking=True)
        d_ctx.copy_(h_ctx, non_blocking=True)
        d_bt.copy_(h_bt, non_blocking=True)

    set_ctx(d_slots, d_ctx, d_bt)
    g_e2e = pipe.capture(layers, lambda l: (hq[l], hk[l], hv[l]), lambda l: hout[l], prologue=meta_in)
    m.reset_context()

    with ClockSampler(env.local_rank) as clocks:
        # the decode-only graph (roofline), the full step (value) and the host-buffer step (e2e) in alternating
        # rounds: the three figures see the same clocks
        ms = env.timed_interleaved({"step": g_full.replay, "decode": g_dec.replay, "e2e": g_e2e.replay},
                                   args.steps, args.warmup)
        # nvidia-smi answers in ~100 ms: keep the same load running until a few readings exist
        t_end = time.time() + 3.0
        while len(clocks.samples) < 5 and time.time() < t_end:
            for _ in range(10):
                g_full.replay()
            torch.cuda.synchronize()
    ms_step, ms_dec, ms_e2e = ms["step"], ms["decode"], ms["e2e"]
    h2d = L * (hq[0].numel() + hk[0].numel() + hv[0].numel()) * 2 + (h_slots.numel() + h_ctx.numel() + h_bt.numel()) * 4
    d2h = L * hout[0].numel() * 2
    del kv_cache, layers, g_full, g_dec, g_e2e, pipe, outs, hq, hk, hv, hout
    torch.cuda.empty_cache()

    bytes_layer_rank = kv_bytes_per_layer(ctx_sum, KVH, D)
    bytes_step_job = bytes_layer_rank * L * tp
    launches_per_step = L  # one paged_decode launch per layer: the store of the new tokens and the split reduction are fused into it
    peaks = load_peaks()
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = bytes_layer_rank / (ms_dec / L * 1e-3) / 1e9
    return {
        "metric": METRIC, "value": bytes_step_job / (ms_step * 1e-3) / 1e9, "unit": "GB/s",
        "n_gpus": tp, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {
            "workload": wl["name"], "layers": L, "batch": B, "context": wl["ctx"] if "ragged" not in wl else f"U{list(wl['ragged'])}, sum {ctx_sum}",
            "q_heads_per_rank": H, "kv_heads_per_rank": KVH, "head_dim": D, "kvcache_block_size": BLOCK, "parallelism": f"tp{tp}",
            "num_kv_splits": splits, "cuda_graph": True,
            "timing": "full-step, decode-only and host-buffer graphs timed in alternating rounds, exactly `steps` replays each",
            "l2": f"inputs ({bytes_layer_rank * L / 1e9:.1f} GB of KV per rank per step) are >> 126 MB L2; no flush needed",
        },
        "tokens_per_s_attention_only": B / (ms_step * 1e-3),
        "roofline": {
            "bound": "hbm", "kernel": "paged_decode_kernel<bf16,128>", "achieved": achieved, "peak": peak,
            "unit": "GB/s", "frac": achieved / peak, "traffic": matching_traffic(wl_key, KVH, B, ctx_sum, H),
            "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy, read+write)" if peaks else "fallback 6650 GB/s",
            "frac_of_8TBps_spec": achieved / 8000.0,
            "bytes_per_launch": bytes_layer_rank, "us_per_launch": ms_dec / L * 1e3,
            "step_ms": ms_step, "dominant_kernel_ms_per_step": ms_dec,
        },
        "e2e": {"value": bytes_step_job / (ms_e2e * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "api": E2E_API},
        "gpu_launches": launches_per_step * args.steps,
        "kernels": kernels,
        "clocks": clocks.summary(),
    }


# ---------------------------------------------------------------------------
# leg 2: prefill workloads
# ---------------------------------------------------------------------------


def prefill_flops_per_layer(wl, heads):
    """4 * H * D * sum_s (Lq*C + Lq(Lq+1)/2) — unmasked (q,k) pairs only."""
    lq, c = wl["new"], wl["cached"]
    pairs = wl["seqs"] * (lq * c + lq * (lq + 1) // 2)
    return 4 * heads * wl["D"] * pairs


def run_prefill_cpu_baseline(wl, steps, warmup, sample_heads=4):
    import torch

    from oracle import eager_port as port

    torch.set_num_threads(os.cpu_count() or 1)
    g = torch.Generator().manual_seed(0)
    d, group = wl["D"], wl["H"] // wl["KVH"]
    h, kvh = sample_heads, max(1, sample_heads // group)
    lk, lq = wl["cached"] + wl["new"], wl["new"]
    nblk = lk // BLOCK
    kc = torch.randn(nblk, BLOCK, kvh, d, generator=g).to(torch.bfloat16)
    vc = torch.randn(nblk, BLOCK, kvh, d, generator=g).to(torch.bfloat16)
    bt = torch.randperm(nblk, generator=g).view(1, nblk).to(torch.int32)
    q = torch.randn(lq, h, d, generator=g).to(torch.bfloat16)
    cu_q, cu_k = torch.tensor([0, lq]), torch.tensor([0, lk])
    scale = d ** -0.5
    for _ in range(max(1, warmup)):
        port.prefill_paged(q, kc, vc, bt, cu_q, cu_k, scale)
    t0 = time.perf_counter()
    for _ in range(steps):
        port.prefill_paged(q, kc, vc, bt, cu_q, cu_k, scale)
    dt = (time.perf_counter() - t0) / steps
    one = dict(wl, seqs=1)
    return {
        "value": prefill_flops_per_layer(one, h) / dt / 1e12, "unit": "TFLOP/s", "cores": torch.get_num_threads(), "kind": "port",
        "sample": f"1 of {wl['seqs']} sequences ({wl['cached']} cached + {wl['new']} new), {h} of {wl['H']} heads, 1 of {wl['layers']} layers, bf16 "
                  "(oracle/eager_port.prefill_paged: gather through block table + the reference's eager matmul/softmax/matmul, attention.py:554-597; "
                  "a port because the reference's own eager prefill mis-slices K/V under prefix hits, SURVEY 8(a) P2)",
        "ms_per_step": dt * 1e3,
    }


def run_prefill(env: Env, args, wl, steps=None) -> dict:
    torch = env.torch
    import minivllm_b200 as m

    steps = steps or args.steps
    tp, rank, dev = env.world, env.rank, env.dev
    L, H, KVH, D = wl["layers"], wl["H"] // tp, wl["KVH"] // tp, wl["D"]
    S, C, NQ = wl["seqs"], wl["cached"], wl["new"]
    lk = C + NQ
    nblk = (lk + BLOCK - 1) // BLOCK
    NB = S * nblk + 4
    scale = D ** -0.5
    torch.manual_seed(1234 + rank)
    kv_cache = torch.empty(2, L, NB, BLOCK, KVH, D, dtype=torch.bfloat16, device=dev).normal_()
    g = torch.Generator().manual_seed(0)
    perm = torch.randperm(NB, generator=g)[: S * nblk].view(S, nblk).to(torch.int32)
    block_tables = perm.to(dev)
    pos = torch.arange(C, lk)
    slot_mapping = torch.cat([perm[s, pos // BLOCK].long() * BLOCK + pos % BLOCK for s in range(S)]).to(torch.int32).to(dev)
    cu_q = torch.arange(0, (S + 1) * NQ, NQ, dtype=torch.int32, device=dev)
    cu_k = torch.arange(0, (S + 1) * lk, lk, dtype=torch.int32, device=dev)
    N = S * NQ
    layers = []
    for l in range(L):
        a = m.Attention(H, D, scale, KVH)
        a.k_cache, a.v_cache = kv_cache[0, l], kv_cache[1, l]
        layers.append(a)
    q = torch.randn(N, H, D, device=dev).to(torch.bfloat16)
    k = torch.randn(N, KVH, D, device=dev).to(torch.bfloat16)
    v = torch.randn(N, KVH, D, device=dev).to(torch.bfloat16)
    outs = torch.empty(2, N, H, D, dtype=torch.bfloat16, device=dev)
    flush = torch.empty(192 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def set_ctx(sl, bt, cq, ck):
        m.set_context(True, NQ, lk, cq, ck, sl, None, bt)

    def full_step():  # store + prefix-aware prefill attention, every layer (eager launches, like the engine)
        for l, a in enumerate(layers):
            outs[l & 1] = a(q, k, v)

    def attn_only():
        for l, a in enumerate(layers):
            m.prefill_varlen_attention(q, None, None, cu_q, cu_k, NQ, lk, scale, k_cache=a.k_cache, v_cache=a.v_cache,
                                       block_tables=block_tables, out=outs[l & 1])

    def flush_l2():
        flush.fill_(1)

    set_ctx(slot_mapping, block_tables, cu_q, cu_k)
    full_step()
    kernel = m.ops.last_kernel()
    # e2e: host q/k/v per layer in, result out; prefill is eager in the engine (never graph-captured,
    # inference_executor.py:706-716) so this leg launches eagerly, copies on side streams
    hq, hk, hv = q.cpu().pin_memory(), k.cpu().pin_memory(), v.cpu().pin_memory()
    hout = torch.empty(N, H, D, dtype=torch.bfloat16).pin_memory()
    h_meta = [t.cpu().pin_memory() for t in (slot_mapping, block_tables, cu_q, cu_k)]
    d_meta = [torch.empty_like(t) for t in (slot_mapping, block_tables, cu_q, cu_k)]
    pipe = HostPipeline(torch, (q, k, v))

    def e2e_step():
        def meta_in():
            for dst, src in zip(d_meta, h_meta):
                dst.copy_(src, non_blocking=True)
        set_ctx(*d_meta)
        pipe.run(layers, lambda l: (hq, hk, hv), lambda l: hout, prologue=meta_in)

    with ClockSampler(env.local_rank) as clocks:
        per = split_rounds(steps, 2)
        tot = {"step": 0.0, "attn": 0.0}
        for r, n in enumerate(per):
            set_ctx(slot_mapping, block_tables, cu_q, cu_k)
            tot["step"] += env.timed(full_step, n, args.warmup if r == 0 else 1, before_each=flush_l2) * n
            tot["attn"] += env.timed(attn_only, n, args.warmup if r == 0 else 1, before_each=flush_l2) * n
        t_end = time.time() + 2.0
        while len(clocks.samples) < 5 and time.time() < t_end:
            full_step()
            torch.cuda.synchronize()
    ms_step, ms_attn = tot["step"] / steps, tot["attn"] / steps
    ms_e2e = env.timed(e2e_step, max(2, steps // 4), 1, before_each=flush_l2)
    m.reset_context()
    h2d = L * (hq.numel() + hk.numel() + hv.numel()) * 2 + sum(t.numel() for t in h_meta) * 4
    d2h = L * hout.numel() * 2
    del kv_cache, layers, pipe, hq, hk, hv, hout, flush, outs
    torch.cuda.empty_cache()
    fl_layer = prefill_flops_per_layer(wl, H)
    fl_job = fl_layer * L * tp
    peaks = load_peaks()
    peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
    achieved = fl_layer / (ms_attn / L * 1e-3) / 1e12
    return {
        "metric": "prefill attention TFLOP/s", "value": fl_job / (ms_step * 1e-3) / 1e12, "unit": "TFLOP/s",
        "n_gpus": tp, "steps": steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": wl["name"], "layers": L, "sequences": S, "cached_tokens": C, "new_tokens": NQ,
                   "q_heads_per_rank": H, "kv_heads_per_rank": KVH, "head_dim": D, "kvcache_block_size": BLOCK,
                   "parallelism": f"tp{tp}", "cuda_graph": False, "l2": "192 MB write between timed steps flushes the 126 MB L2"},
        "tokens_per_s_attention_only": N / (ms_step * 1e-3),
        "roofline": {"bound": "tensor", "kernel": kernel, "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak, "traffic": matching_prefill_traffic(S, C, NQ, H, KVH, D),
                     "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained (kernel timed inside a 36-layer step)" if peaks else "fallback 1400 TFLOP/s",
                     "frac_of_burst_peak": achieved / float(peaks.get("bf16_tflops", 1590.0)),
                     "flops_per_launch": fl_layer, "us_per_launch": ms_attn / L * 1e3},
        "e2e": {"value": fl_job / (ms_e2e * 1e-3) / 1e12, "unit": "TFLOP/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "api": "minivllm_b200.Attention.forward per layer (eager), pinned host q/k/v in, pinned host out, copies on side streams"},
        "gpu_launches": L * 2 * steps,
        "clocks": clocks.summary(),
    }


# ---------------------------------------------------------------------------
# N >= 2: one untimed parity check of the cross-process collective that follows the path
# ---------------------------------------------------------------------------


def run_collective_check(env: Env) -> dict:
    """The fused peer all-reduce + residual add + RMSNorm (``minivllm_b200.comm``, the collective after the
    row-parallel o_proj / down_proj, linear.py:314-319) across the launcher's ranks, against the CPU oracle (used
    as the checker, outside every timed region).  Every rank regenerates every rank's input from a shared seed."""
    torch, dist = env.torch, env.dist
    from minivllm_b200.comm import NvlsAllReduce, PeerAllReduce
    from oracle import allreduce_oracle as O  # checker only

    world, rank, dev = env.world, env.rank, env.dev
    gloo = dist.new_group(backend="gloo")
    shapes = [(8, 2560), (256, 1024), (512, 8192)]  # cfg3 decode message, cfg2 decode message, cfg5 decode message
    try:
        comm = NvlsAllReduce(rank, world, max(r * h for r, h in shapes) * 2, group=gloo, device=dev)
        algos = (1, 2, 3)
    except Exception:  # noqa: BLE001  (no multicast mapping on this fabric)
        comm = PeerAllReduce(rank, world, max(r * h for r, h in shapes) * 2, group=gloo, device=dev)
        algos = (1, 2)
    worst, exact, nvls_ulps = 0.0, True, 0.0
    for rows, hidden in shapes:
        g = torch.Generator().manual_seed(1000 + rows + hidden)
        xs = [torch.randn(rows, hidden, generator=g).to(torch.bfloat16) for _ in range(world)]
        res0 = torch.randn(rows, hidden, generator=g).to(torch.bfloat16)
        w = (1.0 + 0.1 * torch.randn(hidden, generator=g)).to(torch.bfloat16)
        want_y, want_res = O.allreduce_add_rms_norm(xs, w, 1e-6, res0)
        for algo in algos:
            res = res0.to(dev)
            y, new_res = comm.allreduce_add_rms_norm(xs[rank].to(dev), w.to(dev), 1e-6, res, algo=algo)
            torch.cuda.synchronize()
            worst = max(worst, (y.cpu().float() - want_y.float()).abs().max().item())
            if algo == 3:  # the switch sums and rounds its own way (like NCCL's NVLS path): one ulp of the largest addend
                a, b = new_res.cpu().float(), want_res.float()
                big = torch.stack([t_.float().abs() for t_ in xs] + [res0.float().abs()]).max(dim=0).values
                big = torch.maximum(big, b.abs())  # ... or of the sum itself, whichever is larger
                nvls_ulps = max(nvls_ulps, ((a - b).abs() / (big.clamp_min(2.0 ** -120) * 2.0 ** -7)).max().item())
            else:
                exact = exact and bool(torch.equal(new_res.cpu(), want_res))
    t = torch.tensor([worst, 0.0 if exact else 1.0, nvls_ulps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.barrier()
    comm.close()
    return {"kernel": "peer all-reduce + residual add + RMSNorm (one-shot and two-shot peer stores over NVLink"
                      + ("; NVLS multimem.ld_reduce / multimem.st" if 3 in algos else "") + ")",
            "algos": list(algos), "world": world, "shapes": shapes, "max_abs_vs_oracle": float(t[0].item()), "residual_exact": bool(t[1].item() == 0.0),
            "nvls_residual_max_ulps": float(t[2].item()) if 3 in algos else None,
            "bar": "residual stream bit-exact for world 2 (<= 1 ulp per hop beyond); normalised output within 0.0625 (bf16)"}


# ---------------------------------------------------------------------------
# leg 3: end-to-end decode tokens/s through the reference's own LLM.generate with the layer swapped in
# ---------------------------------------------------------------------------


def engine_prompts(wl, vocab, seed=0):
    import torch

    g = torch.Generator().manual_seed(seed)  # same prompts on every rank
    lo, hi = 10, vocab - 10
    prefixes = [torch.randint(lo, hi, (wl["cached"],), generator=g).tolist() for _ in range(wl["seqs"])]
    warm = [p + [lo + i] for i, p in enumerate(prefixes)]  # hash + retain the prefix blocks (block_manager.py:252-339)
    real = [p + torch.randint(lo, hi, (wl["prompt"] - wl["cached"],), generator=g).tolist() for p in prefixes]
    return warm, real


def engine_config(wl):
    blocks_per_seq = (wl["prompt"] + wl["new_tokens"] + BLOCK - 1) // BLOCK + 1
    return dict(max_num_seqs=wl["seqs"], max_model_len=max(4096, wl["prompt"] + wl["new_tokens"] + BLOCK),
                max_num_batched_tokens=16384,
                num_kvcache_blocks=wl["seqs"] * (blocks_per_seq + wl["cached"] // BLOCK + 2) + 16)


def run_engine(env: Env | None, args, wl) -> dict:
    """cfg3 through ``minivllm.LLM.generate`` (reference engine, ``oracle/_ref``) + ``install_into_reference``."""
    from harness import ref_engine
    from oracle import reference_engine as R

    if R.locate_reference() is None:
        return {"unavailable": "oracle/_ref missing (run __graft_entry__.build() where /root/reference exists)"}
    world = env.world if env is not None else 1
    vocab = R.PRESETS[wl["preset"]]["vocab_size"]
    warm, real = engine_prompts(wl, vocab)
    info = ref_engine.run(
        wl["preset"], device="cuda", dtype="bfloat16", dropin=True, executor=True, fuse_qk=True, enforce_eager=False, fast_init=True,
        peer_allreduce=world > 1 and os.environ.get("MVB200_ENGINE_NCCL") is None,
        calls=[warm, real], max_tokens=[1, wl["new_tokens"]], tensor_parallel_size=world, under_torchrun=world > 1,
        **engine_config(wl))
    if info.get("worker"):
        return {"worker": True}
    c = info["calls"][1]
    return {
        "metric": "end-to-end decode tokens/s", "value": c["decode_tokens"] / c["decode_s"], "unit": "tokens/s",
        "n_gpus": world, "parallelism": f"tp{world}", "dtype": "bf16", "data": "synthetic prompts, random-init weights",
        "engine": "the reference's own minivllm.LLM.generate (oracle/_ref, unmodified) with minivllm_b200.install_into_reference(executor=True, fuse_qk=True"
                  + (", peer_allreduce=True)" if world > 1 and os.environ.get("MVB200_ENGINE_NCCL") is None else ")"),
        "config": {"workload": wl["name"], "model": wl["preset"], "sequences": wl["seqs"], "prompt_tokens": wl["prompt"],
                   "prefix_cached_tokens": wl["cached"], "new_tokens": wl["new_tokens"], "cuda_graph": True},
        "decode_ms_per_step": c["decode_s"] / max(c["decode_steps"], 1) * 1e3, "decode_steps": c["decode_steps"],
        "prefill_s": c["prefill_s"], "prefill_tokens_per_s": c["prefill_tokens"] / max(c["prefill_s"], 1e-9),
        "prefix_cache_hit_blocks": c["cache_hits"], "graphs_captured": info["graphs_captured"],
        "attention_classes": info["attention_classes"], "kernels_seen": info.get("kernels_seen"),
        "engine_init_s": info["init_s"],
        # which collective served the row-parallel all-reduces (None at TP 1: there are none)
        "allreduce": info.get("allreduce"),
    }


def run_engine_cpu_baseline(wl, seqs=2, prompt=256, new_tokens=8):
    """The reference's eager engine on the host cores, reduced batch (stated): a baseline, not a target."""
    from harness import ref_engine
    from oracle import reference_engine as R

    if R.locate_reference() is None:
        raise SystemExit("reference not available")
    small = dict(wl, seqs=seqs, prompt=prompt, cached=0, new_tokens=new_tokens)
    vocab = R.PRESETS[wl["preset"]]["vocab_size"]
    _, real = engine_prompts(small, vocab)
    info = ref_engine.run(wl["preset"], device="cpu", dtype="bfloat16", calls=[real], max_tokens=new_tokens, fast_init=True,
                          isolate_cuda=False, threads=os.cpu_count(), **engine_config(small))
    c = info["calls"][0]
    return {"value": c["decode_tokens"] / c["decode_s"], "unit": "tokens/s", "cores": os.cpu_count(), "kind": "reference",
            "sample": f"{seqs} prompts x {prompt} tokens, {new_tokens} new tokens, bf16, enforce_eager (the reference's CPU eager engine, unmodified)",
            "ms_per_step": c["decode_s"] / max(c["decode_steps"], 1) * 1e3}


# ---------------------------------------------------------------------------


def finish(env: Env | None, hard_exit: bool):
    sys.stdout.flush()
    sys.stderr.flush()
    if env is not None and env.world > 1:
        if hard_exit:
            # captured graphs that hold NCCL kernels make destroy_process_group() block (torch 2.11 / NCCL 2.28):
            # drain and leave without tearing the communicator down
            try:
                env.torch.cuda.synchronize()
            except Exception:
                pass
            os._exit(0)
        if env.dist.is_initialized():
            env.dist.destroy_process_group()


_STDOUT = sys.stdout


def emit(line: dict) -> None:
    """The ONE JSON line of a run, on the process's real stdout (everything else — the reference's loggers included —
    goes to stderr, see main)."""
    _STDOUT.write(json.dumps(line) + "\n")
    _STDOUT.flush()


def main():
    # the reference configures loggers that write to sys.stdout when it is imported: keep stdout for the JSON line
    sys.stdout = sys.stderr
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="default", choices=["default"] + sorted(WORKLOADS))
    ap.add_argument("--splits", type=int, default=0, help="override the KV split count (0 = library heuristic)")
    ap.add_argument("--ctx", type=int, default=0, help="exploration: override the context length of a decode workload")
    ap.add_argument("--batch", type=int, default=0, help="exploration: override the batch of a decode workload")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-prefill", action="store_true", help="default workload: skip the cfg3 prefill leg")
    ap.add_argument("--no-engine", action="store_true", help="default workload: skip the Qwen3-4B engine leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    default = args.workload == "default"
    wl_key = "cfg2" if default else args.workload
    wl = dict(WORKLOADS[wl_key])
    if args.ctx or args.batch:
        wl.update(ctx=args.ctx or wl["ctx"], batch=args.batch or wl["batch"], name=wl["name"] + f" [override ctx={args.ctx} batch={args.batch}]")
    if args.impl == "reference":
        run_reference_arm(args, wl)
        return
    kind = wl.get("kind", "decode")
    if kind == "engine":
        env = Env() if int(os.environ.get("WORLD_SIZE", "1")) > 1 else None
        line = run_engine(env, args, wl)
        if not line.get("worker") and int(os.environ.get("RANK", "0")) == 0:
            emit(line)
        finish(env, hard_exit=True)
        return
    env = Env()
    if kind == "prefill":
        line = run_prefill(env, args, wl)
        if env.rank == 0 and env.world == 1 and not args.no_cpu_baseline:
            base = run_prefill_cpu_baseline(wl, steps=3, warmup=1)
            line["cpu_baseline"] = {k_: base[k_] for k_ in ("value", "unit", "cores", "kind", "sample")}
        if env.rank == 0:
            emit(line)
        finish(env, hard_exit=False)
        return
    line = run_decode(env, args, wl, wl_key)
    hard = False
    if default:
        if not args.no_prefill:
            try:
                p = run_prefill(env, args, WORKLOADS["cfg3p"], steps=max(4, args.steps // 2))
                line["prefill"] = {k_: p[k_] for k_ in ("metric", "value", "unit", "ms_per_step", "steps", "config", "roofline", "e2e", "gpu_launches")}
                line["gpu_launches"] += p["gpu_launches"]
            except Exception as e:  # noqa: BLE001
                line["prefill"] = {"error": repr(e)[:300]}
        if env.world > 1:
            try:
                line["collective_check"] = run_collective_check(env)
            except Exception as e:  # noqa: BLE001
                line["collective_check"] = {"error": repr(e)[:300]}
        if not args.no_engine:
            # the reference engine's own shutdown tears NCCL state down: last GPU leg
            hard = True
            try:
                e = run_engine(env, args, WORKLOADS["cfg3e"])
                if e.get("worker"):
                    finish(env, hard_exit=True)
                    return
                line["qwen3_4b_tok_s"] = e
            except Exception as ex:  # noqa: BLE001
                line["qwen3_4b_tok_s"] = {"error": repr(ex)[:400]}
        if env.rank == 0 and env.world == 1 and not args.no_cpu_baseline:
            base = run_cpu_baseline(wl, steps=10, warmup=1)
            line["cpu_baseline"] = {k_: base[k_] for k_ in ("value", "unit", "cores", "kind", "sample")}
    elif env.rank == 0 and env.world == 1 and not args.no_cpu_baseline:
        base = run_cpu_baseline(wl, steps=10, warmup=1)
        line["cpu_baseline"] = {k_: base[k_] for k_ in ("value", "unit", "cores", "kind", "sample")}
    if env.rank == 0:
        emit(line)
    finish(env, hard_exit=hard)


if __name__ == "__main__":
    main()
