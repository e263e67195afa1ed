"""Per-step attention metadata, same object as the reference's.

Mirrors ``minivllm/utils/context.py:30-143``: a ``Context`` dataclass carried in a
``ContextVar`` with ``set_context / get_context / reset_context``.  Field names,
dtypes and meaning are unchanged (north-star requirement); two optional fields
are added for the CUDA-Graph decode path, both defaulting to "let the op
decide" so that code which builds ``Context(...)`` directly keeps working.
"""

from __future__ import annotations

from contextvars import ContextVar
from dataclasses import dataclass

import torch


@dataclass
class Context:
    is_prefill: bool = False
    max_seqlen_q: int = 0
    max_seqlen_k: int = 0
    cum_seqlens_q: torch.Tensor | None = None  # int32 [S+1]
    cum_seqlens_k: torch.Tensor | None = None  # int32 [S+1]
    slot_mapping: torch.Tensor | None = None   # int32 [N]
    context_lens: torch.Tensor | None = None   # int32 [B]
    block_tables: torch.Tensor | None = None   # int32 [B, MB]
    # additions (not in the reference): frozen split count + preallocated split workspace
    decode_num_splits: int | None = None
    decode_workspace: torch.Tensor | None = None


_CONTEXT_VAR: ContextVar[Context | None] = ContextVar("mvb200_inference_context", default=None)


def get_context() -> Context:
    ctx = _CONTEXT_VAR.get()
    return ctx if ctx is not None else Context()


def set_context(
    is_prefill: bool,
    max_seqlen_q: int = 0,
    max_seqlen_k: int = 0,
    cum_seqlens_q: torch.Tensor | None = None,
    cum_seqlens_k: torch.Tensor | None = None,
    slot_mapping: torch.Tensor | None = None,
    context_lens: torch.Tensor | None = None,
    block_tables: torch.Tensor | None = None,
    decode_num_splits: int | None = None,
    decode_workspace: torch.Tensor | None = None,
) -> None:
    _CONTEXT_VAR.set(
        Context(
            is_prefill=is_prefill,
            max_seqlen_q=max_seqlen_q,
            max_seqlen_k=max_seqlen_k,
            cum_seqlens_q=cum_seqlens_q,
            cum_seqlens_k=cum_seqlens_k,
            slot_mapping=slot_mapping,
            context_lens=context_lens,
            block_tables=block_tables,
            decode_num_splits=decode_num_splits,
            decode_workspace=decode_workspace,
        )
    )


def reset_context() -> None:
    _CONTEXT_VAR.set(Context())
