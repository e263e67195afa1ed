"""Attention-metadata preparation, paged-cache allocation and the CUDA-Graph decode runner.

Mirrors the parts of the reference's ``InferenceExecutor`` that feed the hot path
(``minivllm/engine/inference_executor.py``): ``_allocate_kv_cache`` (:198-243),
``_assign_kv_cache_to_layers`` (:320-339), ``_prepare_prefill_input`` (:468-549),
``_prepare_decode_input`` (:551-617), ``_prepare_block_tables(_decode)`` (:619-688),
``capture_device_graphs`` / ``_capture_graph_for_batch_size`` (:823-920) and
``_execute_with_cuda_graph`` (:728-774).  Same tensors, names, dtypes and
block/slot layout; what changes:

* the cache holds this rank's KV heads (``num_kv_heads // tp``), not the global count;
* graph outputs are allocated in the model dtype and the whole ``[bs, hidden]``
  hidden state is copied (the reference copies row 0 of it);
* a fixed KV split count and its workspace are part of the captured state;
* host-side prep is vectorised (numpy) instead of per-token Python loops;
* optionally (``capture_graphs(token_fn=...)``) the decode graph carries its own bookkeeping: a
  ``decode_advance`` node updates ids/positions/slots/context_lens/block_tables on the device and a
  greedy-argmax node leaves the next tokens on the device, so a steady-state step costs the host one
  graph replay, one [B] token read-back and at most one [B] int32 upload.

Sequences are duck-typed: ``len(seq)``, ``seq.token_ids``, ``seq.num_cached_tokens``,
``seq.block_table`` (the reference's ``Sequence`` qualifies, engine/sequence.py:31).
"""

from __future__ import annotations

from typing import Any, Callable, Sequence as Seq

import numpy as np
import torch

from . import context as _context
from . import ops

GRAPH_MAX_BS = 512


def graph_batch_sizes(max_num_seqs: int) -> list[int]:
    """The reference's capture ladder (inference_executor.py:839-841), clipped to ``max_num_seqs``."""
    sizes = [1, 2, 4, 8, 12] + list(range(16, min(max_num_seqs, GRAPH_MAX_BS) + 1, 16))
    sizes = [s for s in sizes if s <= max_num_seqs]
    if max_num_seqs <= GRAPH_MAX_BS and max_num_seqs not in sizes:
        sizes.append(max_num_seqs)
    return sorted(set(sizes))


class HotPathExecutor:
    """Owns the paged KV cache and everything the attention kernels consume per step."""

    def __init__(
        self,
        model: torch.nn.Module | None,
        *,
        num_layers: int,
        num_kv_heads: int,
        head_dim: int,
        hidden_size: int,
        dtype: torch.dtype,
        device: torch.device | str,
        max_num_seqs: int,
        max_model_len: int,
        block_size: int = 64,
        num_heads: int | None = None,
        set_context: Callable[..., None] = _context.set_context,
        reset_context: Callable[[], None] = _context.reset_context,
        get_context: Callable[[], Any] = _context.get_context,
    ) -> None:
        if block_size % 64:
            raise ValueError(f"kvcache_block_size must be a multiple of 64, got {block_size}")  # config.py:178-185
        self.model = model
        self.num_layers = num_layers
        self.num_kv_heads = num_kv_heads  # per rank
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.hidden_size = hidden_size
        self.dtype = dtype
        self.device = torch.device(device)
        self.block_size = block_size
        self.max_num_seqs = max_num_seqs
        self.max_model_len = max_model_len
        self.max_blocks_per_seq = (max_model_len + block_size - 1) // block_size
        self._set_context = set_context
        self._reset_context = reset_context
        self._get_context = get_context
        self._pin = self.device.type == "cuda"
        self.kv_cache: torch.Tensor | None = None
        self.num_blocks = 0
        # persistent staging for decode (inference_executor.py:149-180)
        n, mb = max_num_seqs, self.max_blocks_per_seq
        self._h_ids = self._host(n, torch.int64)
        self._h_pos = self._host(n, torch.int64)
        self._h_slots = self._host(n, torch.int32)
        self._h_ctx = self._host(n, torch.int32)
        self._h_bt = self._host((n, mb), torch.int32)
        self._d_ids = torch.zeros(n, dtype=torch.int64, device=self.device)
        self._d_pos = torch.zeros(n, dtype=torch.int64, device=self.device)
        self._d_slots = torch.zeros(n, dtype=torch.int32, device=self.device)
        self._d_ctx = torch.zeros(n, dtype=torch.int32, device=self.device)
        self._d_bt = torch.full((n, mb), -1, dtype=torch.int32, device=self.device)
        # CUDA graphs
        self.graphs: dict[int, Any] = {}
        self.graph_bs: list[int] = []
        self.graph_vars: dict[str, torch.Tensor] = {}
        self.graph_splits: dict[int, int] = {}
        self.graph_pool = None
        # device-resident decode loop (capture_graphs(token_fn=...), begin_decode, step_decode)
        self._token_fn: Callable[[torch.Tensor, torch.Tensor], None] | None = None
        self._loop_b = self._loop_bs = 0
        self._h_app: torch.Tensor | None = None

    # ------------------------------------------------------------------ helpers
    def _host(self, shape, dtype) -> torch.Tensor:
        t = torch.zeros(shape, dtype=dtype)
        return t.pin_memory() if self._pin else t

    def _to_device(self, arr: np.ndarray, dtype: torch.dtype) -> torch.Tensor:
        t = torch.from_numpy(np.ascontiguousarray(arr)).to(dtype)
        if self._pin:
            t = t.pin_memory()
        return t.to(self.device, non_blocking=True)

    # ------------------------------------------------------------------ T4: cache
    def kv_block_bytes(self) -> int:
        return 2 * self.num_layers * self.block_size * self.num_kv_heads * self.head_dim * self.dtype.itemsize

    def allocate_kv_cache(self, num_blocks: int) -> torch.Tensor:
        """``[2, L, NB, BS, KVH_rank, D]`` (inference_executor.py:209-221), zero-initialised."""
        if num_blocks <= 0:
            raise ValueError(f"Calculated KV cache blocks ({num_blocks}) must be > 0.")
        self.num_blocks = num_blocks
        self.kv_cache = torch.zeros(
            2, self.num_layers, num_blocks, self.block_size, self.num_kv_heads, self.head_dim,
            dtype=self.dtype, device=self.device,
        )
        self.assign_kv_cache_to_layers()
        return self.kv_cache

    def num_blocks_for_budget(self, free_bytes: int) -> int:
        """``min(budget // block_bytes, max_num_seqs * ceil(max_model_len / BS))`` (:291-302)."""
        cap = self.max_num_seqs * self.max_blocks_per_seq
        return max(0, min(int(free_bytes // self.kv_block_bytes()), cap))

    def assign_kv_cache_to_layers(self) -> int:
        """Attribute discovery in ``model.modules()`` order (inference_executor.py:325-331)."""
        if self.kv_cache is None or self.model is None:
            return 0
        layer = 0
        for module in self.model.modules():
            if hasattr(module, "k_cache") and hasattr(module, "v_cache"):
                if layer >= self.num_layers:
                    raise RuntimeError("model has more attention layers than the cache has planes")
                module.k_cache = self.kv_cache[0, layer]
                module.v_cache = self.kv_cache[1, layer]
                layer += 1
        return layer

    # ------------------------------------------------------------------ T2: prefill metadata
    def prepare_prefill(self, seqs: Seq[Any]) -> tuple[torch.Tensor, torch.Tensor]:
        """Build ids/positions/cu_seqlens/slot_mapping/block_tables and ``set_context`` (:468-549)."""
        bs = self.block_size
        ids, pos, slots = [], [], []
        cu_q, cu_k = [0], [0]
        max_q = max_k = 0
        any_table = False
        for seq in seqs:
            n = len(seq)
            cached = seq.num_cached_tokens
            ids.append(np.asarray(seq.token_ids[cached:], dtype=np.int64))
            pos.append(np.arange(cached, n, dtype=np.int64))
            lq = n - cached
            cu_q.append(cu_q[-1] + lq)
            cu_k.append(cu_k[-1] + n)
            max_q, max_k = max(max_q, lq), max(max_k, n)
            if seq.block_table:
                any_table = True
                # slots of blocks num_cached_blocks .. num_blocks-1, the last one partial (:503-510)
                p = np.arange((cached // bs) * bs, n, dtype=np.int64)
                table = np.asarray(seq.block_table, dtype=np.int64)
                slots.append(table[p // bs] * bs + p % bs)
        cat = lambda xs, dt: np.concatenate(xs) if xs else np.zeros(0, dt)  # noqa: E731
        input_ids = self._to_device(cat(ids, np.int64), torch.int64)
        positions = self._to_device(cat(pos, np.int64), torch.int64)
        slot_mapping = self._to_device(cat(slots, np.int64), torch.int32) if slots else None
        block_tables = self.build_block_tables(seqs) if any_table else None
        self._set_context(
            is_prefill=True, max_seqlen_q=max_q, max_seqlen_k=max_k,
            cum_seqlens_q=self._to_device(np.asarray(cu_q), torch.int32),
            cum_seqlens_k=self._to_device(np.asarray(cu_k), torch.int32),
            slot_mapping=slot_mapping, context_lens=None, block_tables=block_tables,
        )
        return input_ids, positions

    def build_block_tables(self, seqs: Seq[Any]) -> torch.Tensor:
        """``[S, max_blocks]`` int32 padded with -1 (inference_executor.py:619-649)."""
        width = max((len(s.block_table) for s in seqs), default=0)
        if not seqs or width == 0:
            return torch.empty((len(seqs), 0), dtype=torch.int32, device=self.device)
        table = np.full((len(seqs), width), -1, dtype=np.int32)
        for i, s in enumerate(seqs):
            table[i, : len(s.block_table)] = s.block_table
        return self._to_device(table, torch.int32)

    # ------------------------------------------------------------------ T3: decode metadata
    def prepare_decode(self, seqs: Seq[Any]) -> tuple[torch.Tensor, torch.Tensor]:
        """Stage last token / position / slot / context_len / block table, one H2D copy each (:551-688)."""
        b, bs = len(seqs), self.block_size
        if b > self.max_num_seqs:
            raise ValueError(f"batch {b} exceeds max_num_seqs {self.max_num_seqs}")
        lens = np.fromiter((len(s) for s in seqs), dtype=np.int64, count=b)
        ids = np.fromiter((s.token_ids[-1] for s in seqs), dtype=np.int64, count=b)
        last_block = np.fromiter((s.block_table[-1] if s.block_table else 0 for s in seqs), dtype=np.int64, count=b)
        has_table = np.fromiter((1 if s.block_table else 0 for s in seqs), dtype=np.int64, count=b)
        # slot = block_table[-1]*BS + last_block_num_tokens - 1 (:568-572); 0 when there is no table (:574)
        slots = (last_block * bs + (lens - 1) % bs) * has_table
        width = max((len(s.block_table) for s in seqs), default=0)
        self._h_ids[:b] = torch.from_numpy(ids)
        self._h_pos[:b] = torch.from_numpy(lens - 1)
        self._h_slots[:b] = torch.from_numpy(slots.astype(np.int32))
        self._h_ctx[:b] = torch.from_numpy(lens.astype(np.int32))
        for dst, src in ((self._d_ids, self._h_ids), (self._d_pos, self._h_pos),
                         (self._d_slots, self._h_slots), (self._d_ctx, self._h_ctx)):
            dst[:b].copy_(src[:b], non_blocking=True)
        if width == 0:
            block_tables = torch.empty((b, 0), dtype=torch.int32, device=self.device)
        elif width <= self.max_blocks_per_seq:
            host = self._h_bt[:b, :width]
            host.fill_(-1)
            view = host.numpy()
            for i, s in enumerate(seqs):
                view[i, : len(s.block_table)] = s.block_table
            self._d_bt[:b, :width].copy_(host, non_blocking=True)
            block_tables = self._d_bt[:b, :width]
        else:
            block_tables = self.build_block_tables(seqs)
        self._set_context(
            is_prefill=False, slot_mapping=self._d_slots[:b], context_lens=self._d_ctx[:b],
            block_tables=block_tables,
        )
        return self._d_ids[:b], self._d_pos[:b]

    # ------------------------------------------------------------------ T5: CUDA graphs
    def capture_graphs(self, batch_sizes: list[int] | None = None, *,
                       token_fn: Callable[[torch.Tensor, torch.Tensor], None] | None = None,
                       capture: bool = True) -> list[int]:
        """Capture ``model(input_ids, positions)`` for the decode ladder, largest first (:823-920).

        Must run on every rank (the reference captures on rank 0 only, model_runner.py:113-115).
        ``token_fn(hidden[bs, hidden], next_tokens[bs] int64)`` — when given, each graph becomes a
        self-contained decode step: ``decode_advance`` -> model -> ``token_fn`` (see ``begin_decode``).
        ``capture=False`` only allocates the persistent buffers; ``step_decode`` then runs the same three
        stages eagerly (``enforce_eager``).
        """
        if self.model is None or self.device.type != "cuda":
            raise RuntimeError("CUDA graphs need a CUDA device and a model")
        sizes = batch_sizes or graph_batch_sizes(self.max_num_seqs)
        max_bs, mb = max(sizes), self.max_blocks_per_seq
        dev = self.device
        heads = self.num_heads or self.num_kv_heads
        self.graph_splits = {bs: ops.decode_num_splits(bs, self.num_kv_heads, mb * self.block_size, dev) for bs in sizes}
        ws = max(ops.decode_workspace_numel(bs, heads, self.head_dim, s) for bs, s in self.graph_splits.items())
        self.graph_vars = {
            "input_ids": torch.zeros(max_bs, dtype=torch.int64, device=dev),
            "positions": torch.zeros(max_bs, dtype=torch.int64, device=dev),
            "slot_mapping": torch.full((max_bs,), -1, dtype=torch.int32, device=dev),
            "context_lens": torch.zeros(max_bs, dtype=torch.int32, device=dev),
            "block_tables": torch.zeros(max_bs, mb, dtype=torch.int32, device=dev),
            "outputs": torch.zeros(max_bs, self.hidden_size, dtype=self.dtype, device=dev),
            "workspace": ops.decode_workspace(ws, dev),  # tickets / work counter zeroed once, self re-arming
            "next_tokens": torch.zeros(max_bs, dtype=torch.int64, device=dev),
            "appends": torch.full((max_bs,), -1, dtype=torch.int32, device=dev),
        }
        self.graph_bs = sorted(sizes)
        self._token_fn = token_fn
        gv = self.graph_vars
        if token_fn is not None:
            self.graph_vars["block_tables"].fill_(-1)

        def body(bs: int) -> None:
            if token_fn is not None:
                self._advance(bs)
            gv["outputs"][:bs].copy_(self.model(gv["input_ids"][:bs], gv["positions"][:bs]))
            if token_fn is not None:
                token_fn(gv["outputs"][:bs], gv["next_tokens"][:bs])

        for bs in reversed(self.graph_bs if capture else []):
            self._set_context(
                is_prefill=False, slot_mapping=gv["slot_mapping"][:bs], context_lens=gv["context_lens"][:bs],
                block_tables=gv["block_tables"][:bs], decode_num_splits=self.graph_splits[bs],
                decode_workspace=gv["workspace"],
            )
            try:
                side = torch.cuda.Stream(device=dev)
                side.wait_stream(torch.cuda.current_stream(dev))
                with torch.cuda.stream(side), torch.inference_mode():
                    body(bs)  # warm-up
                torch.cuda.current_stream(dev).wait_stream(side)
                torch.cuda.synchronize(dev)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, pool=self.graph_pool), torch.inference_mode():
                    body(bs)
                if self.graph_pool is None:
                    self.graph_pool = graph.pool()
                self.graphs[bs] = graph
            finally:
                self._reset_context()
        torch.cuda.synchronize(dev)
        if token_fn is not None:  # the warm-up passes advanced nothing (context_lens == 0) but leave a clean slate
            gv["context_lens"].zero_()
            gv["slot_mapping"].fill_(-1)
        return self.graph_bs

    # ------------------------------------------------------------------ device-resident decode loop
    def _advance(self, bs: int) -> None:
        gv = self.graph_vars
        ops.decode_advance(gv["next_tokens"][:bs], gv["appends"][:bs], gv["input_ids"][:bs], gv["positions"][:bs],
                           gv["slot_mapping"][:bs], gv["context_lens"][:bs], gv["block_tables"][:bs],
                           self.num_blocks, self.block_size)

    def begin_decode(self, seqs: Seq[Any]) -> int:
        """Load the persistent decode state of ``seqs`` (each already holding its newest token) into the
        graph buffers: block tables as of now, ``context_lens = len-1`` and ``next_tokens = last token``
        — the state ``decode_advance`` turns into exactly what ``_prepare_decode_input`` (:551-617)
        would build for this step.  Rows beyond ``len(seqs)`` of the bucket become padding.  Returns the
        graph bucket.  Call again whenever the set (or order) of running sequences changes."""
        if not self.graph_vars or self._token_fn is None:
            raise RuntimeError("begin_decode needs capture_graphs(token_fn=...)")
        b = len(seqs)
        bs = next((x for x in self.graph_bs if x >= b), None)
        if bs is None:
            raise ValueError(f"batch {b} exceeds the largest captured graph ({self.graph_bs[-1] if self.graph_bs else 0})")
        mb = self.max_blocks_per_seq
        gv = self.graph_vars
        lens = np.fromiter((len(s) for s in seqs), dtype=np.int64, count=b)
        last = np.fromiter((s.token_ids[-1] for s in seqs), dtype=np.int64, count=b)
        table = np.full((bs, mb), -1, dtype=np.int32)
        for i, s in enumerate(seqs):
            if len(s.block_table) > mb:
                raise ValueError(f"sequence needs {len(s.block_table)} blocks, max_model_len allows {mb}")
            table[i, : len(s.block_table)] = s.block_table
        ctx = np.zeros(bs, dtype=np.int32)
        ctx[:b] = lens - 1
        tok = np.zeros(bs, dtype=np.int64)
        tok[:b] = last
        gv["block_tables"][:bs].copy_(self._to_device(table, torch.int32))
        gv["context_lens"][:bs].copy_(self._to_device(ctx, torch.int32))
        gv["next_tokens"][:bs].copy_(self._to_device(tok, torch.int64))
        gv["appends"][:bs].fill_(-1)
        self._loop_b, self._loop_bs = b, bs
        # two pinned staging buffers + events: a caller that runs ahead of the device (pipelined sampling) must not
        # rewrite a buffer whose H2D copy has not executed yet
        self._h_app = [self._host(bs, torch.int32), self._host(bs, torch.int32)]
        self._h_app_done = [None, None]
        self._h_app_turn = 0
        return bs

    @torch.inference_mode()
    def step_decode(self, appends: dict[int, int] | None = None) -> torch.Tensor:
        """One decode step of the loop opened by ``begin_decode``.  ``appends`` maps row -> physical block
        id for the sequences whose newest token opened a block (BlockManager.may_append, len % BS == 1).
        Returns the device vector of the ``b`` tokens picked by ``token_fn``; hidden states of the step
        are in ``graph_vars['outputs'][:b]``."""
        b, bs = self._loop_b, self._loop_bs
        if bs == 0:
            raise RuntimeError("step_decode called before begin_decode")
        gv = self.graph_vars
        if appends:
            i = self._h_app_turn
            self._h_app_turn ^= 1
            if self._h_app_done[i] is not None:
                self._h_app_done[i].synchronize()  # the copy issued two uploads ago has read this buffer
            buf = self._h_app[i]
            buf.fill_(-1)
            view = buf.numpy()
            for row, block in appends.items():
                view[row] = block
            gv["appends"][:bs].copy_(buf, non_blocking=True)
            if buf.is_pinned():
                ev = self._h_app_done[i] or torch.cuda.Event()
                ev.record()
                self._h_app_done[i] = ev
        if bs in self.graphs:
            self.graphs[bs].replay()
        else:  # eager twin (enforce_eager)
            self._set_context(is_prefill=False, slot_mapping=gv["slot_mapping"][:bs], context_lens=gv["context_lens"][:bs],
                              block_tables=gv["block_tables"][:bs])
            try:
                self._advance(bs)
                gv["outputs"][:bs].copy_(self.model(gv["input_ids"][:bs], gv["positions"][:bs]))
                self._token_fn(gv["outputs"][:bs], gv["next_tokens"][:bs])
            finally:
                self._reset_context()
        return gv["next_tokens"][:b]

    def replay_decode(self, input_ids: torch.Tensor, positions: torch.Tensor) -> torch.Tensor | None:
        """Replay protocol of ``_execute_with_cuda_graph`` (:728-774); ``None`` if no graph fits."""
        b = input_ids.size(0)
        bs = next((x for x in self.graph_bs if x >= b), None)
        if bs is None or bs not in self.graphs:
            return None
        gv = self.graph_vars
        ctx = self._get_context()
        width = ctx.block_tables.size(1)
        gv["block_tables"][:b, :width] = ctx.block_tables  # stale columns persist beyond `width` (:766-768)
        gv["context_lens"].zero_()
        if self._token_fn is not None:
            # self-advancing graph: hand it the state *before* this step; its first node rebuilds
            # ids/positions/slots/context_lens from it (identical to what prepare_decode staged)
            gv["next_tokens"][:b] = input_ids
            gv["context_lens"][:b] = ctx.context_lens - 1
            gv["appends"].fill_(-1)
        else:
            gv["input_ids"][:b] = input_ids
            gv["positions"][:b] = positions
            gv["slot_mapping"].fill_(-1)
            gv["slot_mapping"][:b] = ctx.slot_mapping
            gv["context_lens"][:b] = ctx.context_lens
        self.graphs[bs].replay()
        return gv["outputs"][:b]

    # ------------------------------------------------------------------ one step
    @torch.inference_mode()
    def run(self, seqs: Seq[Any], is_prefill: bool, *, use_graph: bool = True) -> torch.Tensor:
        """Hidden states ``[N, hidden]`` for a batch (``execute_batch`` minus logits/sampling, :412-449)."""
        try:
            if is_prefill:
                ids, pos = self.prepare_prefill(seqs)
                return self.model(ids, pos)
            ids, pos = self.prepare_decode(seqs)
            if use_graph and self.graphs:
                out = self.replay_decode(ids, pos)
                if out is not None:
                    return out
            return self.model(ids, pos)
        finally:
            self._reset_context()
