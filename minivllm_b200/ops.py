"""The three hot-path operators, called through the C ABI of libmvb200.so.

Python keeps only what needs host knowledge: argument checks with the
reference's exception types, stride bookkeeping, output/workspace allocation
and the current CUDA stream.  Every tensor must live on a CUDA device; there is
no CPU path.

Reference call sites (relative to the reference repository):
  store_kvcache              attention_backend.py:196-266 (+ Triton kernel :52-78)
  paged_decode_attention     attention.py:439-447, page_attention.py:35-133
  prefill_varlen_attention   attention.py:396-433, :467-504
  qk_norm_rope_store         qwen_base.py:181-186 + attention.py:228-243 (caller-side fusion)
  decode_advance             inference_executor.py:551-617, :651-688, :751-768 (device-side metadata)
  greedy_argmax              sampling/sampler.py:104-106, :118-130 (temperature == 0)
"""

from __future__ import annotations

import ctypes
import os

import torch

from . import _native

_DTYPE_CODE = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}
_SM_COUNT: dict[int, int] = {}


def _code(dtype: torch.dtype) -> int:
    try:
        return _DTYPE_CODE[dtype]
    except KeyError:
        raise ValueError(f"mvb200: unsupported dtype {dtype}") from None


def _require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "mvb200 ops run on CUDA tensors only (got a "
                f"{t.device.type} tensor); there is no CPU fallback"
            )


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _sm_count(device: torch.device) -> int:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _SM_COUNT:
        _SM_COUNT[idx] = torch.cuda.get_device_properties(idx).multi_processor_count
    return _SM_COUNT[idx]


def _rows(t: torch.Tensor) -> torch.Tensor:
    """[N, heads, D] with dense inner dims (row stride free); copy only if needed."""
    if t.dim() != 3:
        raise ValueError(f"expected a [N, heads, head_dim] tensor, got shape {tuple(t.shape)}")
    if t.size(0) > 0 and (t.stride(2) != 1 or t.stride(1) != t.size(2)):
        t = t.contiguous()
    return t


def _row_stride(t: torch.Tensor) -> int:
    return t.stride(0) if t.size(0) > 1 else t.size(1) * t.size(2)


def _i32(t: torch.Tensor, device: torch.device) -> torch.Tensor:
    if t.device != device:
        t = t.to(device)
    if t.dtype != torch.int32:
        t = t.to(torch.int32)
    return t


# ---------------------------------------------------------------------------
# store_kvcache
# ---------------------------------------------------------------------------


def store_kvcache(
    key: torch.Tensor,
    value: torch.Tensor,
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
) -> None:
    """Scatter ``key[i]``/``value[i]`` to cache row ``slot_mapping[i]`` (``< 0`` skips).

    Same contract as ``StandardAttentionBackend.store_kv_cache``
    (attention_backend.py:196-266): ``ValueError`` when ``slot_mapping`` and the
    batch disagree (:209-212); values are cast to the cache dtype first (:254-257).
    """
    _require_cuda(key, value, k_cache, v_cache, slot_mapping)
    n, kvh, d = key.shape
    if slot_mapping.numel() != n:
        raise ValueError(f"slot_mapping size {slot_mapping.numel()} != batch_size {n}")
    if value.shape != key.shape:
        raise ValueError(f"key {tuple(key.shape)} and value {tuple(value.shape)} differ")
    if n == 0:
        return
    if key.dtype != k_cache.dtype:
        key = key.to(k_cache.dtype)
    if value.dtype != v_cache.dtype:
        value = value.to(v_cache.dtype)
    if not (k_cache.is_contiguous() and v_cache.is_contiguous()):
        raise ValueError("k_cache / v_cache must be contiguous")
    row = kvh * d
    if k_cache.numel() % row or v_cache.numel() != k_cache.numel():
        raise ValueError("cache size is not a multiple of num_kv_heads*head_dim")
    key, value = _rows(key), _rows(value)
    slots = _i32(slot_mapping.reshape(-1), key.device)
    if not slots.is_contiguous():
        slots = slots.contiguous()
    rc = _native.lib().mvb200_store_kvcache(
        key.data_ptr(), value.data_ptr(), _row_stride(key), _row_stride(value),
        k_cache.data_ptr(), v_cache.data_ptr(), slots.data_ptr(),
        n, row, k_cache.numel() // row, _code(k_cache.dtype), _stream(),
    )
    _native.check(rc, "mvb200_store_kvcache")


# ---------------------------------------------------------------------------
# q/k RMSNorm + RoPE + store (caller-side fusion)
# ---------------------------------------------------------------------------


def qk_norm_rope_store(
    q: torch.Tensor,
    k: torch.Tensor,
    v: torch.Tensor,
    positions: torch.Tensor,
    cos_cache: torch.Tensor,
    sin_cache: torch.Tensor,
    q_norm_weight: torch.Tensor | None = None,
    k_norm_weight: torch.Tensor | None = None,
    eps: float = 1e-6,
    slot_mapping: torch.Tensor | None = None,
    k_cache: torch.Tensor | None = None,
    v_cache: torch.Tensor | None = None,
    *,
    inplace: bool = True,
) -> tuple[torch.Tensor, torch.Tensor]:
    """One kernel for ``q_norm``/``k_norm`` -> ``rotary_emb`` -> ``store_kv_cache``.

    Replaces qwen_base.py:181-186 plus the store of attention.py:228-243: RMSNorm over
    ``head_dim`` (layernorm.py:81-93, skipped when the weights are ``None`` — Qwen2), rotation with
    ``cos_cache/sin_cache[positions]`` (fp32 ``[max_pos, D/2]``, rotary_embedding.py:150-193), and
    the scatter of the rotated ``k`` and of ``v`` to ``slot_mapping`` (``None`` or empty: no store).
    Returns ``(q, k)`` rotated — the inputs themselves when ``inplace`` (they are views of the
    fused qkv GEMM output that nothing else reads), fresh tensors otherwise.
    """
    _require_cuda(q, k, v, positions, cos_cache, sin_cache, q_norm_weight, k_norm_weight, slot_mapping, k_cache, v_cache)
    n, h, d = q.shape
    kvh = k.size(1)
    if k.shape != (n, kvh, d) or v.shape != (n, kvh, d):
        raise ValueError(f"q {tuple(q.shape)}, k {tuple(k.shape)}, v {tuple(v.shape)} are inconsistent")
    if positions.numel() != n:
        raise ValueError(f"positions size {positions.numel()} != num_tokens {n}")
    if k.dtype != q.dtype or v.dtype != q.dtype:
        raise ValueError("q, k, v must share a dtype")
    if cos_cache.dtype != torch.float32 or sin_cache.dtype != torch.float32:
        raise ValueError("cos/sin caches must be float32 (rotary_embedding.py:157-165)")
    if cos_cache.dim() != 2 or cos_cache.size(1) * 2 != d or sin_cache.shape != cos_cache.shape:
        raise ValueError(f"cos/sin caches must be [max_pos, {d // 2}], got {tuple(cos_cache.shape)}")
    if (q_norm_weight is None) != (k_norm_weight is None):
        raise ValueError("q_norm_weight and k_norm_weight must be given together")
    for w in (q_norm_weight, k_norm_weight):
        if w is not None and (w.numel() != d or w.dtype != q.dtype):
            raise ValueError("norm weights must be [head_dim] in the activation dtype")
    store = slot_mapping is not None and slot_mapping.numel() > 0 and k_cache is not None and k_cache.numel() > 0
    num_slots = 0
    if store:
        if slot_mapping.numel() != n:
            raise ValueError(f"slot_mapping size {slot_mapping.numel()} != batch_size {n}")
        if k_cache.dtype != q.dtype or v_cache.dtype != q.dtype:
            raise ValueError("cache dtype must equal the activation dtype for the fused store")
        if not (k_cache.is_contiguous() and v_cache.is_contiguous()):
            raise ValueError("k_cache / v_cache must be contiguous")
        num_slots = k_cache.numel() // (kvh * d)
    if n == 0:
        return q, k
    q, k, v = _rows(q), _rows(k), _rows(v)
    q_out, k_out = (q, k) if inplace else (torch.empty_like(q, memory_format=torch.contiguous_format),
                                           torch.empty_like(k, memory_format=torch.contiguous_format))
    pos = positions.reshape(-1)
    if pos.dtype != torch.int64 or not pos.is_contiguous():
        pos = pos.to(torch.int64).contiguous()
    slots = None
    if store:
        slots = _i32(slot_mapping.reshape(-1), q.device).contiguous()
    cos_cache, sin_cache = cos_cache.contiguous(), sin_cache.contiguous()
    rc = _native.lib().mvb200_qk_norm_rope_store(
        q.data_ptr(), _row_stride(q), k.data_ptr(), _row_stride(k), v.data_ptr(), _row_stride(v),
        q_out.data_ptr(), _row_stride(q_out), k_out.data_ptr(), _row_stride(k_out),
        pos.data_ptr(), cos_cache.data_ptr(), sin_cache.data_ptr(), cos_cache.size(0),
        q_norm_weight.data_ptr() if q_norm_weight is not None else None,
        k_norm_weight.data_ptr() if k_norm_weight is not None else None, float(eps),
        slots.data_ptr() if store else None,
        k_cache.data_ptr() if store else None, v_cache.data_ptr() if store else None, num_slots,
        n, h, kvh, d, _code(q.dtype), _stream(),
    )
    _native.check(rc, "mvb200_qk_norm_rope_store")
    return q_out, k_out


# ---------------------------------------------------------------------------
# paged decode attention
# ---------------------------------------------------------------------------


def decode_num_splits(batch: int, num_kv_heads: int, max_context: int, device: torch.device) -> int:
    """Split count for a batch; fixed per captured batch size under CUDA Graphs."""
    return int(_native.lib().mvb200_decode_suggest_splits(batch, num_kv_heads, max_context, _sm_count(device)))


def decode_workspace_numel(batch: int, num_heads: int, head_dim: int, num_splits: int) -> int:
    """float32 elements of decode workspace: split partials (when ``num_splits > 1``) plus the fixed tail
    that holds the arrival tickets and the work counter of the persistent kernel."""
    return (int(_native.lib().mvb200_decode_workspace_bytes(batch, num_heads, head_dim, num_splits)) + 3) // 4


def decode_workspace(numel: int, device: torch.device) -> torch.Tensor:
    """Allocate AND initialise a decode workspace of ``numel`` float32 (``mvb200_decode_workspace_init``:
    tickets and counters zeroed, buffer stamped).  The kernel re-arms it after every launch, so one buffer
    serves any number of later calls — on ONE stream at a time."""
    numel = (max(int(numel), decode_workspace_numel(1, 1, 64, 1)) + 3) // 4 * 4
    ws = torch.empty(numel, dtype=torch.float32, device=device)
    rc = _native.lib().mvb200_decode_workspace_init(ws.data_ptr(), ws.numel() * 4, _stream())
    _native.check(rc, "mvb200_decode_workspace_init")
    return ws


_DEFAULT_WS: dict[tuple[int, int], torch.Tensor] = {}


def _default_decode_workspace(need: int, device: torch.device) -> torch.Tensor:
    """Per (device, stream) workspace for eager calls that pass none (grown on demand)."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), _stream())
    ws = _DEFAULT_WS.get(key)
    if ws is None or ws.numel() < need:
        ws = decode_workspace(max(need, 1 << 20), device)
        _DEFAULT_WS[key] = ws
    return ws


_DEBUG_CHECKS = os.environ.get("MVB200_DEBUG_CHECKS", "") not in ("", "0")


def _check_decode_slots(slot_mapping: torch.Tensor, block_tables: torch.Tensor, context_lens: torch.Tensor,
                        block_size: int) -> None:
    """The fused store's contract (include/mvb200.h, ``mvb200_decode_store``): row ``b`` goes to the slot of position
    ``context_lens[b] - 1`` of sequence ``b``.  A slot elsewhere would be patched into the wrong smem row and the
    attention output would differ from the cache.  Padding rows (slot < 0 or context <= 0) are exempt."""
    pos = context_lens.long() - 1
    live = (slot_mapping >= 0) & (pos >= 0)
    col = (pos.clamp(min=0) // block_size).clamp(max=block_tables.size(1) - 1)
    want_block = block_tables.gather(1, col.unsqueeze(1)).squeeze(1).long()
    ok = (slot_mapping.long() // block_size == want_block) & (slot_mapping.long() % block_size == pos % block_size)
    bad = (live & ~ok).nonzero().flatten()
    if bad.numel():
        b = int(bad[0])
        raise ValueError(f"paged_decode: slot_mapping[{b}] = {int(slot_mapping[b])} is not the slot of position "
                         f"context_lens[{b}] - 1 = {int(pos[b])} (block {int(want_block[b])}, block_size {block_size})")


def paged_decode_attention(
    q: torch.Tensor,
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    block_tables: torch.Tensor,
    context_lens: torch.Tensor,
    scale: float,
    *,
    num_splits: int | None = None,
    workspace: torch.Tensor | None = None,
    out: torch.Tensor | None = None,
    kv_prefix_stable: bool = False,
    new_key: torch.Tensor | None = None,
    new_value: torch.Tensor | None = None,
    slot_mapping: torch.Tensor | None = None,
) -> torch.Tensor:
    """``softmax(scale * q K^T) V`` of one query token per sequence over paged KV.

    ``kv_prefix_stable=True`` (``MVB200_DECODE_KV_PREFIX_STABLE``): the kernel launched just before this call
    writes nothing but the newest token of each sequence (and q) — in particular neither ``context_lens`` nor
    ``block_tables``, which must be in place before that predecessor starts.  The kernel may then read the
    metadata and fetch older K/V tiles before the predecessor has finished (programmatic dependent launch).
    Leave it off when a metadata update or a store of older rows directly precedes the call.

    ``MVB200_DEBUG_CHECKS=1`` in the environment makes a call with a fused store verify (host sync, not
    capturable) that ``slot_mapping[b] // block_size == block_tables[b, (context_lens[b]-1) // block_size]``.

    ``new_key`` / ``new_value`` ``[B, KVH, D]`` + ``slot_mapping [B]``: the step's new tokens, stored into the
    caches by the decode kernel itself (``store_kvcache`` fused in: bit-identical cache contents, one launch
    less per layer).  ``slot_mapping[b]`` must be the slot of position ``context_lens[b] - 1`` (or negative).

    ``q [B, H, D]``, caches ``[NB, BS, KVH, D]``, ``block_tables [B, MB]`` int32,
    ``context_lens [B]`` int32 (current token included).  Returns ``[B, H, D]``.
    Error contract follows ``PageAttention.__call__`` (page_attention.py:45-46,
    :62-65) except for the data-dependent "no valid entry" check (:88-91), which
    would need a host sync and is not capturable: such rows return zeros.
    """
    if k_cache.numel() == 0 or v_cache.numel() == 0:
        raise ValueError("KV cache is empty — cannot compute page attention")
    _require_cuda(q, k_cache, v_cache)
    if k_cache.dim() != 4 or k_cache.shape != v_cache.shape:
        raise ValueError("k_cache / v_cache must be [num_blocks, block_size, num_kv_heads, head_dim]")
    if not (k_cache.is_contiguous() and v_cache.is_contiguous()):
        raise ValueError("k_cache / v_cache must be contiguous")
    q = _rows(q)
    b, h, d = q.shape
    nb, bs, kvh, dc = k_cache.shape
    if dc != d:
        raise ValueError(f"head_dim mismatch: q has {d}, cache has {dc}")
    if h % kvh:
        raise ValueError(f"num_heads {h} is not a multiple of num_kv_heads {kvh}")
    if q.dtype != k_cache.dtype:
        q = q.to(k_cache.dtype)
    if out is None:
        out = torch.empty((b, h, d), dtype=q.dtype, device=q.device)
    elif out.shape != (b, h, d) or out.dtype != q.dtype or not out.is_contiguous():
        raise ValueError("out must be a contiguous [B, H, D] tensor of q's dtype")
    if b == 0:
        return out
    block_tables = _i32(block_tables, q.device)
    context_lens = _i32(context_lens.reshape(-1), q.device)
    if block_tables.dim() != 2 or block_tables.size(0) != b or context_lens.numel() != b:
        raise ValueError("block_tables must be [B, MB] and context_lens [B]")
    mb = block_tables.size(1)
    if mb > 0 and block_tables.stride(1) != 1:
        block_tables = block_tables.contiguous()
    if not context_lens.is_contiguous():
        context_lens = context_lens.contiguous()
    if num_splits is None:
        num_splits = decode_num_splits(b, kvh, mb * bs, q.device) if mb > 0 else 1
    ws_ptr, ws_bytes = 0, 0
    if mb > 0:
        need = decode_workspace_numel(b, h, d, num_splits)
        if workspace is None:
            if torch.cuda.is_current_stream_capturing():
                # a captured launch cannot share the eager default workspace (its lifetime is not the
                # graph's): without one, capture the split-free, statically scheduled variant
                num_splits = 1
            else:
                workspace = _default_decode_workspace(need, q.device)
        elif workspace.numel() * workspace.element_size() < need * 4 or not workspace.is_contiguous():
            raise ValueError(f"decode workspace too small: need {need} float32 (ops.decode_workspace)")
        if workspace is not None:
            ws_ptr, ws_bytes = workspace.data_ptr(), workspace.numel() * workspace.element_size()
    store = None
    if new_key is not None:
        if new_value is None or slot_mapping is None:
            raise ValueError("new_key needs new_value and slot_mapping")
        _require_cuda(new_key, new_value, slot_mapping)
        new_key, new_value = _rows(new_key), _rows(new_value)
        if new_key.shape != (b, kvh, d) or new_value.shape != (b, kvh, d):
            raise ValueError(f"new_key / new_value must be [B, KVH, D] = {(b, kvh, d)}")
        if new_key.dtype != k_cache.dtype or new_value.dtype != k_cache.dtype:
            raise ValueError("new_key / new_value must have the cache dtype")
        slot_mapping = _i32(slot_mapping.reshape(-1), q.device)
        if slot_mapping.numel() != b:  # attention_backend.py:248-250
            raise ValueError(f"slot_mapping size ({slot_mapping.numel()}) must match number of tokens ({b})")
        if not slot_mapping.is_contiguous():
            slot_mapping = slot_mapping.contiguous()
        if _DEBUG_CHECKS and mb > 0 and not torch.cuda.is_current_stream_capturing():
            _check_decode_slots(slot_mapping, block_tables, context_lens, bs)
        store = _native.DecodeStore(new_key.data_ptr(), new_value.data_ptr(), _row_stride(new_key), _row_stride(new_value),
                                    slot_mapping.data_ptr())
    rc = _native.lib().mvb200_paged_decode_ex(
        q.data_ptr(), _row_stride(q), k_cache.data_ptr(), v_cache.data_ptr(),
        block_tables.data_ptr() if mb > 0 else 0, block_tables.stride(0) if mb > 0 else 0,
        context_lens.data_ptr(), out.data_ptr(), h * d, ws_ptr, ws_bytes,
        b, h, kvh, d, bs, nb, mb, num_splits, float(scale), _code(q.dtype), 1 if kv_prefix_stable else 0,
        ctypes.byref(store) if store is not None else None, _stream(),
    )
    _native.check(rc, "mvb200_paged_decode_ex")
    return out


# ---------------------------------------------------------------------------
# varlen causal prefill attention
# ---------------------------------------------------------------------------


def prefill_varlen_attention(
    q: torch.Tensor,
    k: torch.Tensor | None,
    v: torch.Tensor | None,
    cu_seqlens_q: torch.Tensor,
    cu_seqlens_k: torch.Tensor,
    max_seqlen_q: int,
    max_seqlen_k: int,
    scale: float,
    *,
    k_cache: torch.Tensor | None = None,
    v_cache: torch.Tensor | None = None,
    block_tables: torch.Tensor | None = None,
    out: torch.Tensor | None = None,
) -> torch.Tensor:
    """Varlen causal attention over packed sequences (bottom-right aligned mask).

    With ``block_tables`` the keys/values are the paged caches (prefix-cache
    path, attention.py:399-400); without, the dense ``k``/``v`` rows addressed by
    ``cu_seqlens_k`` (attention.py:481-488).  Returns ``[N, H, D]``.
    """
    _require_cuda(q, k, v, k_cache, v_cache, block_tables)
    if cu_seqlens_q is None or cu_seqlens_k is None:
        raise RuntimeError(
            "Cumulative sequence lengths not set in context for prefill phase. "
            "This is a bug in the inference pipeline."
        )
    q = _rows(q)
    n, h, d = q.shape
    if out is None:
        out = torch.empty((n, h, d), dtype=q.dtype, device=q.device)
    elif out.shape != (n, h, d) or out.dtype != q.dtype or not out.is_contiguous():
        raise ValueError("out must be a contiguous [N, H, D] tensor of q's dtype")
    cu_q = _i32(cu_seqlens_q.reshape(-1), q.device).contiguous()
    cu_k = _i32(cu_seqlens_k.reshape(-1), q.device).contiguous()
    s = cu_q.numel() - 1
    if cu_k.numel() != s + 1:
        raise ValueError("cu_seqlens_q and cu_seqlens_k must have the same length")
    if n == 0 or s <= 0:
        return out
    paged = block_tables is not None and block_tables.numel() > 0
    if paged:
        if k_cache is None or v_cache is None or k_cache.numel() == 0:
            raise ValueError("block_tables given but the KV cache is empty")
        if not (k_cache.is_contiguous() and v_cache.is_contiguous()):
            raise ValueError("k_cache / v_cache must be contiguous")
        nb, bs, kvh, dc = k_cache.shape
        if q.dtype != k_cache.dtype:
            q = q.to(k_cache.dtype)
            out_cast = out
            out = torch.empty_like(q)
        else:
            out_cast = None
        bt = _i32(block_tables, q.device)
        if bt.stride(1) != 1:
            bt = bt.contiguous()
        if bt.size(0) != s:
            raise ValueError("block_tables must have one row per sequence")
        args_kv = (0, 0, 0, 0, 0, k_cache.data_ptr(), v_cache.data_ptr(), bt.data_ptr(), bt.stride(0))
        mb = bt.size(1)
    else:
        if k is None or v is None:
            raise ValueError("dense prefill needs k and v")
        k, v = _rows(k), _rows(v)
        if k.dtype != q.dtype or v.dtype != q.dtype:
            k, v = k.to(q.dtype), v.to(q.dtype)
        kvh, dc = k.size(1), k.size(2)
        nb, bs, mb = 0, 0, 0
        out_cast = None
        args_kv = (k.data_ptr(), v.data_ptr(), _row_stride(k), _row_stride(v), k.size(0), 0, 0, 0, 0)
    if dc != d:
        raise ValueError(f"head_dim mismatch: q has {d}, k/v have {dc}")
    if h % kvh:
        raise ValueError(f"num_heads {h} is not a multiple of num_kv_heads {kvh}")
    rc = _native.lib().mvb200_prefill_varlen(
        q.data_ptr(), _row_stride(q), *args_kv, cu_q.data_ptr(), cu_k.data_ptr(),
        out.data_ptr(), h * d, s, n, int(max_seqlen_q), int(max_seqlen_k),
        h, kvh, d, bs, nb, mb, float(scale), _code(q.dtype), _stream(),
    )
    _native.check(rc, "mvb200_prefill_varlen")
    if out_cast is not None:
        out_cast.copy_(out)
        return out_cast
    return out


# ---------------------------------------------------------------------------
# device-side decode bookkeeping
# ---------------------------------------------------------------------------


def decode_advance(
    next_tokens: torch.Tensor,
    appends: torch.Tensor,
    input_ids: torch.Tensor,
    positions: torch.Tensor,
    slot_mapping: torch.Tensor,
    context_lens: torch.Tensor,
    block_tables: torch.Tensor,
    num_blocks: int,
    block_size: int = 64,
) -> None:
    """Advance the persistent decode metadata by one token, in place, on the device.

    What ``_prepare_decode_input`` + ``_prepare_block_tables_decode`` + the graph-buffer refill
    (inference_executor.py:551-617, :651-688, :751-768) recompute on the host every step:
    ``input_ids = next_tokens``, ``positions = len-1``, ``context_lens = len``,
    ``slot = block_tables[(len-1)//BS]*BS + (len-1)%BS`` after writing ``appends[b]`` (a block id or -1)
    into column ``(len-1)//BS``.  Rows with ``context_lens <= 0`` are padding.  ``appends`` is reset to -1.
    """
    _require_cuda(next_tokens, appends, input_ids, positions, slot_mapping, context_lens, block_tables)
    b = context_lens.numel()
    for name, t, dt in (("next_tokens", next_tokens, torch.int64), ("appends", appends, torch.int32),
                        ("input_ids", input_ids, torch.int64), ("positions", positions, torch.int64),
                        ("slot_mapping", slot_mapping, torch.int32), ("context_lens", context_lens, torch.int32)):
        if t.dtype != dt or t.dim() != 1 or t.numel() != b or not t.is_contiguous():
            raise ValueError(f"decode_advance: {name} must be a contiguous {dt} vector of {b} rows, "
                             f"got {t.dtype} {tuple(t.shape)}")
    if block_tables.dtype != torch.int32 or block_tables.dim() != 2 or block_tables.size(0) != b:
        raise ValueError(f"decode_advance: block_tables must be int32 [{b}, max_blocks], got "
                         f"{block_tables.dtype} {tuple(block_tables.shape)}")
    if block_tables.size(1) > 0 and block_tables.stride(1) != 1:
        raise ValueError("decode_advance: block_tables rows must be dense")
    if b == 0:
        return
    mb = block_tables.size(1)
    stride = block_tables.stride(0) if b > 1 else max(mb, block_tables.stride(0))
    rc = _native.lib().mvb200_decode_advance(
        next_tokens.data_ptr(), appends.data_ptr(), input_ids.data_ptr(), positions.data_ptr(),
        slot_mapping.data_ptr(), context_lens.data_ptr(), block_tables.data_ptr(), stride,
        b, mb, block_size, num_blocks, _stream(),
    )
    _native.check(rc, "mvb200_decode_advance")


def greedy_argmax_workspace_numel(rows: int, vocab: int) -> int:
    """int64 elements of scratch ``greedy_argmax`` needs for ``rows x vocab`` logits."""
    return (int(_native.lib().mvb200_greedy_argmax_workspace_bytes(rows, vocab)) + 7) // 8


def greedy_argmax(
    logits: torch.Tensor,
    out: torch.Tensor | None = None,
    *,
    out_value: torch.Tensor | None = None,
    index_offset: int = 0,
    workspace: torch.Tensor | None = None,
) -> torch.Tensor:
    """``index_offset + argmax(logits, dim=-1)`` as int64 — the sampler's temperature == 0 branch
    (sampling/sampler.py:104-106).  Ties go to the lowest index, NaN is the maximum (torch.argmax).
    ``out_value`` (fp32 ``[rows]``) receives the winning logit; ``workspace`` (int64) must be passed
    under CUDA-graph capture when the caller wants no allocation inside the graph."""
    _require_cuda(logits, out, out_value, workspace)
    if logits.dim() != 2:
        raise ValueError(f"greedy_argmax: expected [rows, vocab] logits, got {tuple(logits.shape)}")
    rows, vocab = logits.shape
    if vocab == 0:
        raise ValueError("greedy_argmax: empty vocabulary")
    if logits.stride(1) != 1:
        logits = logits.contiguous()
    if out is None:
        out = torch.empty(rows, dtype=torch.int64, device=logits.device)
    elif out.dtype != torch.int64 or out.numel() != rows or not out.is_contiguous():
        raise ValueError("greedy_argmax: out must be a contiguous int64 vector with one entry per row")
    if out_value is not None and (out_value.dtype != torch.float32 or out_value.numel() != rows
                                  or not out_value.is_contiguous()):
        raise ValueError("greedy_argmax: out_value must be a contiguous float32 vector with one entry per row")
    if rows == 0:
        return out
    need = greedy_argmax_workspace_numel(rows, vocab)
    if workspace is None:
        workspace = torch.empty(max(need, 1), dtype=torch.int64, device=logits.device)
    elif workspace.dtype != torch.int64 or workspace.numel() < need:
        raise ValueError(f"greedy_argmax: workspace must be int64 with at least {need} elements")
    rc = _native.lib().mvb200_greedy_argmax(
        logits.data_ptr(), logits.stride(0) if rows > 1 else max(vocab, logits.stride(0)), rows, vocab,
        _code(logits.dtype), workspace.data_ptr(), workspace.numel() * 8, out.data_ptr(),
        out_value.data_ptr() if out_value is not None else None, index_offset, _stream(),
    )
    _native.check(rc, "mvb200_greedy_argmax")
    return out


def last_kernel() -> str:
    """Kernel family the last call on this thread dispatched to (for tests / bench)."""
    return _native.last_kernel()


def kernels_seen() -> list[str]:
    """Distinct kernel families this process has launched so far (``mvb200_kernels_seen``)."""
    return _native.kernels_seen()
