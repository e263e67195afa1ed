"""``torch.ops.mvb200.*`` — the hot-path operators registered as custom torch ops.

Same C-ABI launchers as ``ops.py`` (the registered functions call those wrappers), with the mutation of
``k_cache`` / ``v_cache`` / the decode metadata annotated and shape-only fake kernels, so that tracing
front-ends (``torch.compile``, export) see opaque, correctly aliased ops instead of ctypes calls.  The
``Attention`` layer itself calls ``ops.py`` directly: the dispatcher adds several microseconds per call that
a 10-microsecond decode launch does not need, and CUDA-Graph capture works either way.

Importing this module registers the ops; ``import minivllm_b200`` does so automatically.
"""

from __future__ import annotations

from typing import Optional

import torch
from torch import Tensor
from torch.library import custom_op

from . import ops

NAMESPACE = "mvb200"


@custom_op(f"{NAMESPACE}::store_kvcache", mutates_args=("k_cache", "v_cache"))
def store_kvcache(key: Tensor, value: Tensor, k_cache: Tensor, v_cache: Tensor, slot_mapping: Tensor) -> None:
    ops.store_kvcache(key, value, k_cache, v_cache, slot_mapping)


@store_kvcache.register_fake
def _(key, value, k_cache, v_cache, slot_mapping) -> None:
    return None


@custom_op(f"{NAMESPACE}::paged_decode", mutates_args=())
def paged_decode(q: Tensor, k_cache: Tensor, v_cache: Tensor, block_tables: Tensor, context_lens: Tensor,
                 scale: float, num_splits: int = 0) -> Tensor:
    return ops.paged_decode_attention(q, k_cache, v_cache, block_tables, context_lens, scale,
                                      num_splits=num_splits if num_splits > 0 else None)


@paged_decode.register_fake
def _(q, k_cache, v_cache, block_tables, context_lens, scale, num_splits=0):
    return q.new_empty(q.shape)


@custom_op(f"{NAMESPACE}::prefill_varlen", mutates_args=())
def prefill_varlen(q: Tensor, k: Optional[Tensor], v: Optional[Tensor], cu_seqlens_q: Tensor, cu_seqlens_k: Tensor,
                   max_seqlen_q: int, max_seqlen_k: int, scale: float, k_cache: Optional[Tensor] = None,
                   v_cache: Optional[Tensor] = None, block_tables: Optional[Tensor] = None) -> Tensor:
    return ops.prefill_varlen_attention(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, scale,
                                        k_cache=k_cache, v_cache=v_cache, block_tables=block_tables)


@prefill_varlen.register_fake
def _(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, scale, k_cache=None, v_cache=None,
      block_tables=None):
    return q.new_empty(q.shape)


@custom_op(f"{NAMESPACE}::decode_advance",
           mutates_args=("appends", "input_ids", "positions", "slot_mapping", "context_lens", "block_tables"))
def decode_advance(next_tokens: Tensor, appends: Tensor, input_ids: Tensor, positions: Tensor, slot_mapping: Tensor,
                   context_lens: Tensor, block_tables: Tensor, num_blocks: int, block_size: int = 64) -> None:
    ops.decode_advance(next_tokens, appends, input_ids, positions, slot_mapping, context_lens, block_tables,
                       num_blocks, block_size)


@decode_advance.register_fake
def _(next_tokens, appends, input_ids, positions, slot_mapping, context_lens, block_tables, num_blocks,
      block_size=64) -> None:
    return None


@custom_op(f"{NAMESPACE}::greedy_argmax", mutates_args=())
def greedy_argmax(logits: Tensor, index_offset: int = 0) -> Tensor:
    return ops.greedy_argmax(logits, index_offset=index_offset)


@greedy_argmax.register_fake
def _(logits, index_offset=0):
    return logits.new_empty((logits.shape[0],), dtype=torch.int64)


REGISTERED = ("store_kvcache", "paged_decode", "prefill_varlen", "decode_advance", "greedy_argmax")
