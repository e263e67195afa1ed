"""Rebind the three executor subsystems north_star says change, on the reference's own classes.

``install_into_reference(executor=True)`` ends here.  Nothing in the reference is edited on disk: the
methods below are assigned onto ``minivllm.engine.inference_executor.InferenceExecutor`` at run time and
delegate to a ``HotPathExecutor`` (``executor.py``) that shares the reference executor's model, KV cache
and ``Context`` accessors.

| reference method (inference_executor.py)            | replaced by                                   | why |
|---|---|---|
| ``__init__`` (:59-95)                                | same + ``num_kv_heads //= tp``                | cache sized with the global KV-head count (defect 2) |
| ``_prepare_prefill_input`` (:468-549)                | ``HotPathExecutor.prepare_prefill``           | O(N) Python ``extend(range())`` per step -> numpy |
| ``_prepare_decode_input`` + ``_prepare_block_tables_decode`` (:551-688) | ``HotPathExecutor.prepare_decode`` | per-sequence ``torch.tensor`` per step -> one vectorised fill, pinned staging |
| ``capture_device_graphs`` (:823-920)                 | ``HotPathExecutor.capture_graphs``            | ``outputs.copy_(model(...)[0])`` row-0 broadcast, fp32 outputs buffer, frozen split count + workspace |
| ``_execute_with_cuda_graph`` (:728-774)              | ``HotPathExecutor.replay_decode`` + ``compute_logits`` | same replay protocol on the buffers above |
| ``_sample_tokens`` (:776-806), all-greedy batches only | ``ops.greedy_argmax``                         | the full sampling pipeline + multinomial on a one-hot distribution |
| ``initialize`` (:184-196), ``execute_batch`` (:412-449), TP > 1 only | same + capture on every rank; workers skip logits/sampling | graphs captured on rank 0 only (model_runner.py:113-115); ``ParallelLMHead`` returns ``None`` off rank 0 |

The scheduler, block manager, sequences, models, sampler and ``LLM.generate`` are untouched.
"""

from __future__ import annotations

import importlib
from typing import Any

import torch

from .executor import HotPathExecutor

_PATCHED = "_mvb200_patched"


def _hot(self) -> HotPathExecutor:
    """The ``HotPathExecutor`` twin of a reference ``InferenceExecutor`` (created on first use)."""
    hot = getattr(self, "_mvb200_hot", None)
    if hot is not None and hot.kv_cache is self.kv_cache:
        return hot
    ref_ctx = importlib.import_module("minivllm.utils.context")

    def set_context(**kw: Any) -> None:
        extra = {k: kw.pop(k) for k in ("decode_num_splits", "decode_workspace") if k in kw}
        ref_ctx.set_context(**kw)
        if extra:  # the reference's Context is a plain dataclass: optional fields ride on the instance
            ctx = ref_ctx.get_context()
            for k, v in extra.items():
                setattr(ctx, k, v)

    hf = self.config.hf_config
    tp = torch.distributed.get_world_size() if torch.distributed.is_initialized() else 1
    # the first call comes from the reference's warm-up, which runs under torch.inference_mode()
    # (inference_executor.py:396-404): persistent staging buffers must be ordinary tensors
    with torch.inference_mode(False):
        hot = _make_hot(self, hf, tp, set_context, ref_ctx)
    if self.kv_cache is not None:
        hot.kv_cache = self.kv_cache
        hot.num_blocks = self.kv_cache.size(2)
    self._mvb200_hot = hot
    return hot


def _make_hot(self, hf, tp, set_context, ref_ctx) -> HotPathExecutor:
    return HotPathExecutor(
        self.model,
        num_layers=self.num_layers,
        num_kv_heads=self.num_kv_heads,
        head_dim=self.head_dim,
        hidden_size=hf.hidden_size,
        dtype=self.dtype,
        device=self.device,
        max_num_seqs=self.config.max_num_seqs,
        max_model_len=self.config.max_model_len,
        block_size=self.block_size,
        num_heads=max(hf.num_attention_heads // tp, 1),
        set_context=set_context,
        reset_context=ref_ctx.reset_context,
        get_context=ref_ctx.get_context,
    )


def install_executor() -> None:
    mod = importlib.import_module("minivllm.engine.inference_executor")
    cls = mod.InferenceExecutor
    if getattr(cls, _PATCHED, False):
        return
    orig_init = cls.__init__

    def __init__(self, config, model) -> None:
        orig_init(self, config, model)
        tp = torch.distributed.get_world_size() if torch.distributed.is_initialized() else 1
        if tp > 1:
            if self.num_kv_heads % tp:
                raise ValueError(f"num_kv_heads ({self.num_kv_heads}) must be divisible by tp_size ({tp})")
            self.num_kv_heads //= tp  # this rank's share: what the sharded layers hand to Attention

    def _prepare_prefill_input(self, sequences):
        return _hot(self).prepare_prefill(sequences)

    def _prepare_decode_input(self, sequences):
        return _hot(self).prepare_decode(sequences)

    orig_initialize = cls.initialize
    orig_execute_batch = cls.execute_batch

    def initialize(self, max_num_batched_tokens: int, max_num_seqs: int) -> None:
        orig_initialize(self, max_num_batched_tokens, max_num_seqs)
        # graphs on EVERY rank: the reference captures on rank 0 only (model_runner.py:113-115), which cannot
        # work once the captured forward holds collectives
        if torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1:
            self.capture_device_graphs(max_num_seqs)

    def execute_batch(self, sequences, prefill: bool = True):
        tp_worker = torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1 \
            and torch.distributed.get_rank() != 0
        if not tp_worker or not sequences:
            return orig_execute_batch(self, sequences, prefill)
        # ParallelLMHead gathers the logits on rank 0 and returns None elsewhere (embed_head.py:217-229): workers
        # run the same forward (their half of every collective) and sample nothing
        ref_ctx = importlib.import_module("minivllm.utils.context")
        try:
            input_ids, positions = self._prepare_batch_input(sequences, prefill)
            self._execute_model(input_ids, positions, prefill)
            self.inference_count += 1
            return None, []
        finally:
            ref_ctx.reset_context()

    def capture_device_graphs(self, max_batch_size: int) -> None:
        if self.enforce_eager or self.device.type != "cuda" or self.graphs:
            return
        hot = _hot(self)
        with torch.inference_mode(False):
            hot.capture_graphs(hot_graph_sizes(max_batch_size))
        self.graphs, self.graph_bs, self.graph_vars, self.graph_pool = hot.graphs, hot.graph_bs, hot.graph_vars, hot.graph_pool

    def _execute_with_cuda_graph(self, input_ids, positions):
        out = _hot(self).replay_decode(input_ids, positions)
        if out is None:
            return self._execute_eager(input_ids, positions)
        with torch.inference_mode():
            return self.model.compute_logits(out)  # None on TP workers

    orig_sample = cls._sample_tokens

    def _sample_tokens(self, logits, sequences):
        """Greedy fast path: when every sequence asks for ``temperature == 0`` the
        reference still builds four per-step tensors and runs its whole temperature / top-k / top-p / min-p
        pipeline before ``torch.multinomial`` on a one-hot distribution (inference_executor.py:776-806,
        sampling/functional.py:434-436) — i.e. the argmax its sampler documents for this case
        (sampler.py:104-108).  One ``greedy_argmax`` launch instead; ties go to the lowest index."""
        if logits is None:  # TP worker: the logits live on rank 0 (embed_head.py:217-229)
            return []
        if logits.is_cuda and logits.dim() == 2 and all(seq.temperature == 0 for seq in sequences):
            from . import ops

            return ops.greedy_argmax(logits).tolist()
        return orig_sample(self, logits, sequences)

    cls._sample_tokens = _sample_tokens
    cls.__init__ = __init__
    cls.initialize = initialize
    cls.execute_batch = execute_batch
    cls._prepare_prefill_input = _prepare_prefill_input
    cls._prepare_decode_input = _prepare_decode_input
    cls.capture_device_graphs = capture_device_graphs
    cls._execute_with_cuda_graph = _execute_with_cuda_graph
    setattr(cls, _PATCHED, True)


def hot_graph_sizes(max_batch_size: int) -> list[int]:
    from .executor import graph_batch_sizes

    return graph_batch_sizes(max_batch_size)
