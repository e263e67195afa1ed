"""Minimal prefill-then-decode engine with the reference's block/slot layout (harness, not product).

``Seq`` and ``BlockPool`` restate what the hot path depends on from the reference's
``Sequence`` (engine/sequence.py:31) and ``BlockManager`` (engine/block_manager.py:72-450): 64-token
blocks, FIFO free list, chained 64-bit hashes of full blocks for prefix hits (a hit shares the physical
block and adds 64 to ``num_cached_tokens``), a new block when ``len % 64 == 1``.  Scheduling is the
simplest thing that exercises the path: all prompts are prefilled (in batches bounded by
``max_num_batched_tokens`` / ``max_num_seqs``), then decoded together; sampling is argmax.
"""

from __future__ import annotations

import hashlib
import os
import struct
import time
from collections import deque

import torch

import minivllm_b200 as m
from minivllm_b200.executor import HotPathExecutor

from .model import ModelConfig, QwenLM

try:
    import xxhash

    def _hash64(data: bytes) -> int:
        return xxhash.xxh64(data).intdigest()
except ImportError:  # pragma: no cover

    def _hash64(data: bytes) -> int:
        return int.from_bytes(hashlib.blake2b(data, digest_size=8).digest(), "little")


class Seq:
    def __init__(self, token_ids, max_tokens, block_size=64):
        self.token_ids = list(token_ids)
        self.num_prompt_tokens = len(self.token_ids)
        self.num_cached_tokens = 0
        self.block_table: list[int] = []
        self.block_size = block_size
        self.max_tokens = max_tokens

    def __len__(self):
        return len(self.token_ids)

    @property
    def num_blocks(self):
        return (len(self.token_ids) + self.block_size - 1) // self.block_size

    def block(self, i):
        return self.token_ids[i * self.block_size:(i + 1) * self.block_size]

    @property
    def completion(self):
        return self.token_ids[self.num_prompt_tokens:]

    @property
    def done(self):
        return len(self.token_ids) - self.num_prompt_tokens >= self.max_tokens


class BlockPool:
    def __init__(self, num_blocks, block_size=64):
        self.block_size = block_size
        self.free = deque(range(num_blocks))
        self.ref = [0] * num_blocks
        self.hash_of = [-1] * num_blocks
        self.tokens_of: list[list[int] | None] = [None] * num_blocks
        self.by_hash: dict[int, int] = {}
        self.hits = self.misses = 0

    @staticmethod
    def compute_hash(token_ids, prefix=-1):
        data = b"" if prefix == -1 else prefix.to_bytes(8, "little")
        return _hash64(data + struct.pack(f"<{len(token_ids)}i", *token_ids))

    def _take(self, block_id=None):
        if block_id is None:
            block_id = self.free.popleft()
        else:
            self.free.remove(block_id)
        self.ref[block_id] = 1
        return block_id

    def allocate(self, seq: Seq):
        if len(self.free) < seq.num_blocks:
            raise ValueError(f"No free blocks available for allocation: need {seq.num_blocks}, have {len(self.free)}")
        h = -1
        seq._fresh = []  # blocks whose hash this call registered (their KV is not computed yet)
        hits0, misses0 = self.hits, self.misses
        for i in range(seq.num_blocks):
            toks = seq.block(i)
            h = self.compute_hash(toks, h) if len(toks) == self.block_size else -1
            bid = self.by_hash.get(h, -1) if h != -1 else -1
            if bid != -1 and self.tokens_of[bid] == toks:
                self.hits += 1
                seq.num_cached_tokens += self.block_size
                if self.ref[bid] > 0:
                    self.ref[bid] += 1
                else:
                    self._take(bid)
            else:
                self.misses += 1
                bid = self._take()
                self.hash_of[bid], self.tokens_of[bid] = h, (toks if h != -1 else None)
                if h != -1:
                    self.by_hash[h] = bid
                    seq._fresh.append(bid)
            seq.block_table.append(bid)
        seq._counted = (self.hits - hits0, self.misses - misses0)

    def may_append(self, seq: Seq):
        n, bs = len(seq), self.block_size
        if n % bs == 1:
            bid = self._take()
            self.hash_of[bid], self.tokens_of[bid] = -1, None
            seq.block_table.append(bid)
            return bid
        elif n % bs == 0:
            last = seq.block_table[-1]
            prefix = self.hash_of[seq.block_table[-2]] if len(seq.block_table) > 1 else -1
            toks = seq.block(seq.num_blocks - 1)
            h = self.compute_hash(toks, prefix)
            self.hash_of[last], self.tokens_of[last] = h, toks
            self.by_hash[h] = last
        return None

    def deallocate(self, seq: Seq, computed: bool = True):
        """``computed=False``: the sequence is put back before its prefill ran, so the hashes ``allocate``
        registered for it must not be hit by anybody (their blocks hold no KV yet)."""
        if not computed:
            dh, dm = getattr(seq, "_counted", (0, 0))
            self.hits, self.misses = self.hits - dh, self.misses - dm
            for bid in getattr(seq, "_fresh", []):
                h = self.hash_of[bid]
                if self.by_hash.get(h) == bid:
                    del self.by_hash[h]
                self.hash_of[bid], self.tokens_of[bid] = -1, None
        for bid in reversed(seq.block_table):
            self.ref[bid] -= 1
            if self.ref[bid] == 0:
                self.free.append(bid)  # contents and hash stay: a later prompt can still hit them
        seq.block_table = []
        seq.num_cached_tokens = 0


class Engine:
    def __init__(self, cfg: ModelConfig, *, device, dtype=torch.bfloat16, tp=1, rank=0, max_num_seqs=256,
                 max_model_len=4096, max_num_batched_tokens=16384, num_blocks=None, enforce_eager=False,
                 init="fast", attention_cls=None, seed=1234, device_loop=None):
        self.cfg, self.device, self.dtype, self.tp, self.rank = cfg, torch.device(device), dtype, tp, rank
        self.max_num_seqs, self.max_model_len = max_num_seqs, max_model_len
        self.max_num_batched_tokens = max_num_batched_tokens
        prev = torch.get_default_dtype()
        torch.set_default_dtype(dtype)
        try:
            with torch.device(self.device):
                self.model = QwenLM(cfg, attention_cls or m.Attention, tp, rank)
        finally:
            torch.set_default_dtype(prev)
        torch.manual_seed(seed + rank)
        self.model.init_weights(init, seed)
        self.model.eval()
        if tp > 1 and self.device.type == "cuda" and os.environ.get("MVB200_HARNESS_NCCL", "0") in ("", "0"):
            import torch.distributed as dist

            self.model.enable_peer_allreduce(max_num_seqs, group=dist.new_group(backend="gloo"))
        self.exe = HotPathExecutor(
            self.model, num_layers=cfg.num_layers, num_kv_heads=cfg.num_kv_heads // tp, num_heads=cfg.num_heads // tp,
            head_dim=cfg.head_dim, hidden_size=cfg.hidden_size, dtype=dtype, device=self.device,
            max_num_seqs=max_num_seqs, max_model_len=max_model_len,
        )
        if num_blocks is None:
            num_blocks = max_num_seqs * self.exe.max_blocks_per_seq
        self.exe.allocate_kv_cache(num_blocks)
        self.pool = BlockPool(num_blocks)
        self.enforce_eager = enforce_eager
        # device-resident decode bookkeeping (decode_advance + greedy_argmax inside the graph); the host
        # loop of the reference (prepare_decode every step) stays available for A/B: MVB200_HARNESS_HOST_LOOP=1
        if device_loop is None:
            device_loop = os.environ.get("MVB200_HARNESS_HOST_LOOP", "0") in ("", "0")
        self.device_loop = bool(device_loop) and self.device.type == "cuda"
        if self.device.type == "cuda" and (self.device_loop or not enforce_eager):
            self.exe.capture_graphs(token_fn=self.model.pick_tokens if self.device_loop else None,
                                    capture=not enforce_eager)
        self.stats: dict[str, float] = {}

    def _sync(self):
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)

    @torch.inference_mode()
    def _prefill(self, seqs, forced=None):
        hidden = self.exe.run(seqs, True)
        cu = [0]
        for s in seqs:
            cu.append(cu[-1] + len(s) - s.num_cached_tokens)
        last = torch.tensor([c - 1 for c in cu[1:]], device=hidden.device)  # embed_head.py:212-214
        return hidden[last]

    @torch.inference_mode()
    def generate(self, prompts, max_tokens, forced=None, return_logits=False):
        """Greedy generation.  ``forced[i][t]``, when given, is appended instead of the argmax (teacher
        forcing for parity tests); the model's own choice and logits are still returned."""
        per_seq = isinstance(max_tokens, (list, tuple))
        seqs = [Seq(p, max_tokens[i] if per_seq else max_tokens) for i, p in enumerate(prompts)]
        picked = [[] for _ in seqs]
        logits_log = [[] for _ in seqs]
        order = {id(s): i for i, s in enumerate(seqs)}

        def emit(batch, hidden, step_of, toks=None):
            if toks is None:
                toks = self.model.greedy_tokens(hidden).tolist()
            lg = self.model.compute_logits(hidden).float().cpu() if return_logits else None
            for j, s in enumerate(batch):
                i = order[id(s)]
                picked[i].append(toks[j])
                if lg is not None:
                    logits_log[i].append(lg[j])
                t = len(s) - s.num_prompt_tokens
                s.token_ids.append(forced[i][t] if forced is not None else toks[j])

        self._sync()
        t0 = time.perf_counter()
        waiting = deque(seqs)
        running: list[Seq] = []
        n_prefill_tokens = 0
        while waiting:
            batch, ntok = [], 0
            while waiting and len(batch) < self.max_num_seqs:
                s = waiting[0]
                self.pool.allocate(s)
                need = len(s) - s.num_cached_tokens
                if batch and ntok + need > self.max_num_batched_tokens:
                    self.pool.deallocate(s, computed=False)
                    break
                waiting.popleft()
                batch.append(s)
                ntok += need
            n_prefill_tokens += ntok
            emit(batch, self._prefill(batch), 0)
            running += batch
        self._sync()
        t1 = time.perf_counter()
        steps = 0
        loop_rows = None
        use_loop = self.device_loop and forced is None
        while True:
            live = [s for s in running if not s.done]
            if not live:
                break
            appends = {}
            for j, s in enumerate(live):
                bid = self.pool.may_append(s)
                if bid is not None:
                    appends[j] = bid
            if use_loop:
                rows = [id(s) for s in live]
                if rows != loop_rows:  # (re)load the persistent device state; it already holds the new blocks
                    self.exe.begin_decode(live)
                    loop_rows, appends = rows, None
                toks = self.exe.step_decode(appends).tolist()
                emit(live, self.exe.graph_vars["outputs"][:len(live)], steps, toks)
            else:
                hidden = self.exe.run(live, False, use_graph=not self.enforce_eager)
                emit(live, hidden, steps)
            steps += 1
        self._sync()
        t2 = time.perf_counter()
        n_decode = sum(len(s.completion) for s in seqs) - len(seqs)
        self.stats = {"prefill_s": t1 - t0, "decode_s": t2 - t1, "prefill_tokens": n_prefill_tokens,
                      "decode_tokens": n_decode, "decode_steps": steps,
                      "prefix_hits": self.pool.hits, "prefix_misses": self.pool.misses}
        outs = [s.completion for s in seqs]
        for s in seqs:
            self.pool.deallocate(s)
        if return_logits:
            return outs, picked, logits_log
        return outs, picked
