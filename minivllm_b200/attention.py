"""Drop-in ``Attention`` layer backed by the sm_100a kernels.

Same class surface as the reference's ``minivllm.models.layers.Attention``
(attention.py:102-198): constructor ``(num_heads, head_dim, scale, num_kv_heads,
use_buffered_page_attention=False)``, ``scale == 0 -> 1/sqrt(head_dim)`` (:154),
attributes ``k_cache`` / ``v_cache`` (empty tensors until the executor assigns
views of the paged cache, inference_executor.py:325-331) and
``_cache_initialized``, ``forward(q, k, v) -> [N, H, D]``.  No parameters or
buffers are registered, so ``load_weights`` / ``state_dict`` are unaffected.

``forward`` keeps the reference's ordering contract (attention.py:225-455):
cast to the cache dtype, store the new K/V through ``context.slot_mapping``,
then attend — decode over ``block_tables``/``context_lens``, prefill over
``cu_seqlens`` reading keys back through ``block_tables`` when they are present.
"""

from __future__ import annotations

import math
from typing import Callable

import torch
from torch import nn

from . import context as _context
from . import ops

# Where forward() reads the per-step metadata from.  ``install_into_reference``
# re-points this at the reference's own ContextVar accessor.
_get_context: Callable[[], object] = _context.get_context


def use_context_getter(getter: Callable[[], object]) -> None:
    global _get_context
    _get_context = getter


class PagedDecodeAttention:
    """Callable twin of the reference's ``PageAttention`` (page_attention.py:25-133)."""

    def __call__(self, q, k_cache, v_cache, block_tables, context_lens, scale, **kw):
        if k_cache.numel() == 0 or v_cache.numel() == 0:
            raise ValueError("KV cache is empty — cannot compute page attention")
        if block_tables.size(1) == 0:  # engine warm-up (page_attention.py:62-65)
            return torch.zeros(q.size(0), q.size(1), q.size(2), device=q.device, dtype=q.dtype)
        return ops.paged_decode_attention(q, k_cache, v_cache, block_tables, context_lens, scale, **kw)


class B200AttentionBackend:
    """Stands where the reference keeps ``self.backend`` (attention.py:157-173)."""

    def store_kv_cache(self, key, value, k_cache, v_cache, slot_mapping) -> None:
        ops.store_kvcache(key, value, k_cache, v_cache, slot_mapping)


class Attention(nn.Module):
    def __init__(
        self,
        num_heads: int,
        head_dim: int,
        scale: float,
        num_kv_heads: int,
        use_buffered_page_attention: bool = False,  # accepted for signature parity; one decode kernel
    ) -> None:
        super().__init__()
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.scale = float(scale) if scale != 0.0 else 1.0 / math.sqrt(head_dim)
        self.num_kv_heads = num_kv_heads
        self.backend = B200AttentionBackend()
        self._page_attn = PagedDecodeAttention()
        self.k_cache: torch.Tensor = torch.tensor([])
        self.v_cache: torch.Tensor = torch.tensor([])
        self._cache_initialized = False

    def extra_repr(self) -> str:
        return (
            f"num_heads={self.num_heads}, head_dim={self.head_dim}, "
            f"num_kv_heads={self.num_kv_heads}, scale={self.scale}, backend=mvb200(sm_100a)"
        )

    def forward_fused(
        self,
        positions: torch.Tensor,
        q: torch.Tensor,
        k: torch.Tensor,
        v: torch.Tensor,
        cos_cache: torch.Tensor,
        sin_cache: torch.Tensor,
        q_norm_weight: torch.Tensor | None = None,
        k_norm_weight: torch.Tensor | None = None,
        eps: float = 1e-6,
    ) -> torch.Tensor:
        """``q_norm/k_norm -> rotary_emb -> self.attn(q, k, v)`` of qwen_base.py:181-186 with the
        norm, the rotation and the cache store done by ONE kernel.  ``q, k, v`` are
        the raw views of the fused qkv GEMM output; q and k are rotated in place."""
        context = _get_context()
        k_cache, v_cache = self.k_cache, self.v_cache
        has_cache = k_cache.numel() > 0 and v_cache.numel() > 0
        fusable = not has_cache or q.dtype == k_cache.dtype
        slots = context.slot_mapping if has_cache and fusable else None
        q, k = ops.qk_norm_rope_store(q, k, v, positions, cos_cache, sin_cache, q_norm_weight, k_norm_weight,
                                      eps, slots, k_cache if has_cache else None, v_cache if has_cache else None)
        return self._attend(q, k, v, stored=fusable)

    def forward(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
        # exactly the reference's signature (attention.py:196-198)
        return self._attend(q, k, v, stored=False)

    def _attend(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, *, stored: bool) -> torch.Tensor:
        context = _get_context()
        k_cache, v_cache = self.k_cache, self.v_cache
        has_cache = k_cache.numel() > 0 and v_cache.numel() > 0

        if has_cache and q.dtype != k_cache.dtype:  # attention.py:229-234
            q, k, v = q.to(k_cache.dtype), k.to(k_cache.dtype), v.to(k_cache.dtype)

        fuse_store = None
        # True only while the launch that precedes the decode kernel on the stream is known to write nothing the
        # run-ahead producer reads before its dependency wait (context_lens, block_tables, every K/V row but the
        # newest): the fused q/k path's own store (stored=True), our decode-shaped store_kvcache below, or — with the
        # store fused into the decode kernel — the caller's projection / rotary kernels, which produce q, k, v only.
        predecessor_known = stored
        if has_cache:
            self._cache_initialized = True
            slots = context.slot_mapping
            if not stored and slots is not None and slots.numel() > 0:  # attention.py:237-243
                if (not context.is_prefill and context.block_tables is not None and context.context_lens is not None
                        and context.block_tables.dim() == 2 and context.block_tables.size(1) > 0
                        and slots.numel() == q.size(0) and k.dim() == 3):
                    fuse_store = (k, v, slots)  # decode: the attention kernel stores the new token itself
                    predecessor_known = True
                else:
                    ops.store_kvcache(k, v, k_cache, v_cache, slots)
                    # one slot per sequence = the step's newest positions (inference_executor.py:568-572); any
                    # other slot shape may rewrite older rows, which a run-ahead producer could read stale
                    predecessor_known = (not context.is_prefill and context.context_lens is not None
                                         and slots.numel() == context.context_lens.numel())
        elif not context.is_prefill and not self._cache_initialized:
            raise RuntimeError(
                "KV cache has not been initialized. Ensure ModelRunner.allocate_kv_cache() "
                "is called before inference."
            )

        if context.is_prefill:
            if context.cum_seqlens_q is None or context.cum_seqlens_k is None:
                raise RuntimeError(
                    "Cumulative sequence lengths not set in context for prefill phase. "
                    "This is a bug in the inference pipeline."
                )
            paged = has_cache and context.block_tables is not None and context.block_tables.numel() > 0
            return ops.prefill_varlen_attention(
                q, k, v,
                context.cum_seqlens_q, context.cum_seqlens_k,
                context.max_seqlen_q, context.max_seqlen_k, self.scale,
                k_cache=k_cache if paged else None,
                v_cache=v_cache if paged else None,
                block_tables=context.block_tables if paged else None,
            )

        if not self._cache_initialized:
            raise RuntimeError(
                "KV cache must be initialized before decode phase. "
                "This indicates a problem with the inference pipeline."
            )
        if context.context_lens is None or context.block_tables is None:
            raise RuntimeError(
                "Context lengths or block tables not set for decode phase. "
                "This is a bug in the inference pipeline."
            )
        return self._page_attn(
            q, k_cache, v_cache, context.block_tables, context.context_lens, self.scale,
            num_splits=getattr(context, "decode_num_splits", None),
            workspace=getattr(context, "decode_workspace", None),
            # run-ahead only behind a predecessor we know (see above); the executor uploads the step's metadata
            # before the model runs, so no launch between it and this one writes context_lens / block_tables
            kv_prefix_stable=predecessor_known,
            **({"new_key": fuse_store[0], "new_value": fuse_store[1], "slot_mapping": fuse_store[2]} if fuse_store else {}),
        )


def install_into_reference(executor: bool = False, fuse_qk: bool = False, peer_allreduce: bool = False) -> None:
    """Swap this layer into an importable ``minivllm`` (the reference) before models are built.

    Every reference model imports the class by name from ``minivllm.models.layers``
    (layers/__init__.py:22; qwen_base.py, opt.py, gpt2.py), so the name is rebound
    in each of those modules, and ``forward`` is pointed at the reference's own
    ``get_context`` so that its executor's ``set_context`` calls are seen.

    ``executor=True`` additionally rebinds the attention-metadata preparation, the per-rank cache
    sizing and the CUDA-Graph capture/replay of the reference's ``InferenceExecutor``
    (``reference_patch.py``); without it only the layer changes and the engine must run with
    ``enforce_eager=True`` (the reference's own capture is broken).

    ``fuse_qk=True`` also rebinds ``QwenAttention.forward`` (models/qwen_base.py:146-190) to the one-kernel
    q/k RMSNorm + RoPE + cache store (``forward_fused``).
    ``peer_allreduce=True`` rebinds ``RowParallelLinear.forward`` (linear.py:314-319): decode-size all-reduces go
    through the one-shot NVLink peer kernel, everything else stays on NCCL.
    """
    import importlib

    layers = importlib.import_module("minivllm.models.layers")
    ref_ctx = importlib.import_module("minivllm.utils.context")
    use_context_getter(ref_ctx.get_context)
    layers.Attention = Attention
    for name in (
        "minivllm.models.layers.attention",
        "minivllm.models.qwen_base",
        "minivllm.models.opt",
        "minivllm.models.gpt2",
    ):
        try:
            mod = importlib.import_module(name)
        except Exception:  # optional model files
            continue
        if hasattr(mod, "Attention"):
            mod.Attention = Attention
    if executor:
        from . import reference_patch

        reference_patch.install_executor()
    if fuse_qk:
        from . import reference_patch

        reference_patch.install_fused_qk()
    if peer_allreduce:
        from . import reference_patch

        reference_patch.install_peer_allreduce()
