"""CPU oracle (numpy) for the caller-side sequence in front of the attention hot path.

TEST INFRASTRUCTURE ONLY: imported by ``tests/``, ``__graft_entry__.smoke()`` and the CPU-baseline
legs of ``bench.py``; never by the product path (``minivllm_b200/``).

Restates what ``QwenAttention.forward`` does between the fused qkv GEMM and ``self.attn(q, k, v)``
(reference ``minivllm/models/qwen_base.py:175-186``) plus the store ``Attention.forward`` starts with
(``minivllm/models/layers/attention.py:228-236``) — SURVEY.md §8(f) rank 2:

1. ``RMSNorm`` over ``head_dim`` on q and k, Qwen3 only (``layers/layernorm.py:81-93``):
   ``x.float()`` -> ``x * rsqrt(mean(x^2) + eps) * weight`` in fp32 -> cast back to the model dtype;
2. ``RotaryEmbedding.forward`` (``layers/rotary_embedding.py:166-193``) with ``apply_rotary_emb``
   (``:31-84``): ``x.float()``, halves ``x1, x2 = chunk(x, 2)``, ``y1 = x1*cos - x2*sin``,
   ``y2 = x2*cos + x1*sin`` with ``cos/sin = cache[positions]`` (fp32, ``[max_pos, D/2]``), cast back;
3. ``store_kv_cache`` of the rotated K and the untouched V at ``slot_mapping``
   (``layers/attention_backend.py:196-266``).

Pinned against ``tests/golden/qkrope.npz`` (recorded from the unmodified reference by
``tests/golden/gen_golden.py qkrope``).  Arithmetic is float32 with the reference's rounding points (one
rounding to the model dtype after the norm, one after the rotation); the only freedom left is the order
of the 64/128-term sum inside ``mean``, which can move a result by one unit in the last place of the
model dtype on a small fraction of elements.
"""

from __future__ import annotations

import numpy as np

from .paged_attention_oracle import bf16_bits_to_f32, f32_to_bf16_bits, store_kvcache


def to_f32(x: np.ndarray, dtype: str) -> np.ndarray:
    return bf16_bits_to_f32(x) if dtype == "bf16" else x.astype(np.float32)


def from_f32(x: np.ndarray, dtype: str) -> np.ndarray:
    if dtype == "bf16":
        return f32_to_bf16_bits(x)
    return x.astype(np.float16 if dtype == "f16" else np.float32)


def rope_cache(head_dim: int, max_pos: int, base: float) -> tuple[np.ndarray, np.ndarray]:
    """cos/sin tables ``[max_pos, D/2]`` fp32 (rotary_embedding.py:128-165, no scaling)."""
    inv_freq = (1.0 / (np.float32(base) ** (np.arange(0, head_dim, 2, dtype=np.float32) / np.float32(head_dim)))).astype(np.float32)
    freqs = np.arange(max_pos, dtype=np.float32)[:, None] * inv_freq[None, :]
    return np.cos(freqs).astype(np.float32), np.sin(freqs).astype(np.float32)


def rms_norm_heads(x: np.ndarray, weight: np.ndarray, eps: float, dtype: str) -> np.ndarray:
    """``RMSNorm.rms_forward`` over the last axis (layernorm.py:81-93); x, weight in the model dtype."""
    xf = to_f32(x, dtype)
    var = np.mean(xf * xf, axis=-1, keepdims=True, dtype=np.float32)
    inv = (np.float32(1.0) / np.sqrt(var + np.float32(eps))).astype(np.float32)
    return from_f32((xf * inv) * to_f32(weight, dtype), dtype)


def apply_rope(x: np.ndarray, cos: np.ndarray, sin: np.ndarray, dtype: str) -> np.ndarray:
    """``apply_rotary_emb`` fallback branch (rotary_embedding.py:78-84); cos/sin ``[N, D/2]`` fp32."""
    xf = to_f32(x, dtype)
    half = xf.shape[-1] // 2
    x1, x2 = xf[..., :half], xf[..., half:]
    c, s = cos[:, None, :], sin[:, None, :]
    y1 = (x1 * c).astype(np.float32) - (x2 * s).astype(np.float32)
    y2 = (x2 * c).astype(np.float32) + (x1 * s).astype(np.float32)
    return from_f32(np.concatenate([y1, y2], axis=-1), dtype)


def qk_norm_rope_store(q, k, v, positions, cos_cache, sin_cache, q_weight, k_weight, eps, slot_mapping,
                       k_cache, v_cache, dtype: str):
    """Returns ``(q_out, k_out)`` and scatters ``k_out`` / ``v`` into the caches in place.

    ``q [N,H,D]``, ``k, v [N,KVH,D]`` in the model dtype (bf16 as uint16 bits); ``q_weight/k_weight`` ``[D]``
    or ``None`` (Qwen2: no q/k norm); ``slot_mapping`` ``None`` skips the store (warm-up, no block tables).
    """
    if q_weight is not None:
        q = rms_norm_heads(q, q_weight, eps, dtype)
        k = rms_norm_heads(k, k_weight, eps, dtype)
    pos = np.asarray(positions).astype(np.int64)
    cos, sin = cos_cache[pos], sin_cache[pos]
    q_out = apply_rope(q, cos, sin, dtype)
    k_out = apply_rope(k, cos, sin, dtype)
    if slot_mapping is not None:
        store_kvcache(k_out, np.ascontiguousarray(v), k_cache, v_cache, slot_mapping)
    return q_out, k_out
