"""B200-native (sm_100a) attention hot path for mini-vLLM.

Scope: ``store_kvcache`` scatter, varlen / prefix-aware prefill
attention, paged split-KV GQA decode attention, behind the reference's
``Attention.forward(q, k, v)`` layer, its per-step ``Context`` and the
BlockManager's block/slot layout.  Kernels live in ``csrc/`` and are reached
through the C ABI in ``include/mvb200.h`` (``libmvb200.so``, built in-tree).
"""

from . import _native, ops, torch_ops  # torch_ops registers torch.ops.mvb200.*
from .attention import Attention, PagedDecodeAttention, install_into_reference
from .context import Context, get_context, reset_context, set_context
from .ops import paged_decode_attention, prefill_varlen_attention, store_kvcache

__all__ = [
    "Attention",
    "PagedDecodeAttention",
    "Context",
    "get_context",
    "set_context",
    "reset_context",
    "store_kvcache",
    "paged_decode_attention",
    "prefill_varlen_attention",
    "install_into_reference",
    "ops",
    "torch_ops",
]
__version__ = "0.1.0"
