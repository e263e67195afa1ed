"""Torch-CPU port of the reference's eager hot path.  TEST / BASELINE INFRASTRUCTURE ONLY.

``paged_attention_oracle.py`` is the float64 numpy oracle used for parity.  This
module restates the same three functions with the *operator sequence the
reference really executes on a CPU* (torch advanced-index gather into zero-filled
dense buffers, ``repeat_interleave`` for GQA, ``F.scaled_dot_product_attention``
with a boolean padding mask; boolean-mask + ``index_put`` for the store; a Python
loop of matmul/softmax/matmul per sequence for prefill), so that timing it on
the GPU box's host cores is a fair stand-in for "the reference's CPU path"
(``cpu_baseline.kind == "port"``; ``/root/reference`` does not exist there).

Only ``bench.py`` (``cpu_baseline`` leg and ``--impl reference``) and ``tests/``
import it.  Validated against the golden vectors recorded from the reference in
``tests/test_oracle.py::test_eager_port_*``.

Follows, relative to the reference repository:
  store   minivllm/models/layers/attention_backend.py:237-266
  decode  minivllm/models/layers/page_attention.py:35-133
  prefill minivllm/models/layers/attention.py:467-504, :530-597
"""

from __future__ import annotations

import torch
import torch.nn.functional as F


def store_kvcache(key, value, k_cache, v_cache, slot_mapping) -> None:
    n, kvh, d = key.shape
    width = kvh * d
    if slot_mapping.numel() != n:
        raise ValueError(f"slot_mapping size {slot_mapping.numel()} != batch_size {n}")
    keep = slot_mapping >= 0
    if not bool(keep.any()):
        return
    rows = slot_mapping[keep]
    k_rows = key[keep].reshape(-1, width).to(k_cache.dtype)
    v_rows = value[keep].reshape(-1, width).to(v_cache.dtype)
    k_cache.view(-1, width)[rows] = k_rows
    v_cache.view(-1, width)[rows] = v_rows


def paged_decode(q, k_cache, v_cache, block_tables, context_lens, scale):
    if k_cache.numel() == 0 or v_cache.numel() == 0:
        raise ValueError("KV cache is empty — cannot compute page attention")
    b, h, d = q.shape
    bs, kvh = k_cache.size(1), k_cache.size(2)
    if block_tables.size(1) == 0:
        return q.new_zeros(b, h, d)
    longest = int(context_lens.max())
    pos = torch.arange(longest, dtype=torch.int64).unsqueeze(0)
    in_ctx = pos < context_lens.unsqueeze(1)
    col = (pos // bs).clamp(0, block_tables.size(1) - 1).expand(b, -1)
    page = torch.gather(block_tables.long(), 1, col)
    take = in_ctx & (page >= 0)
    if not bool(take.any()):
        raise RuntimeError("No valid KV cache entries found. Check block_tables and context_lens.")
    dense_k = q.new_zeros(b, longest, kvh, d)
    dense_v = q.new_zeros(b, longest, kvh, d)
    off = (pos % bs).expand(b, -1)
    dense_k[take] = k_cache[page[take], off[take]]
    dense_v[take] = v_cache[page[take], off[take]]
    if kvh != h:
        dense_k = dense_k.repeat_interleave(h // kvh, dim=2)
        dense_v = dense_v.repeat_interleave(h // kvh, dim=2)
    o = F.scaled_dot_product_attention(
        q.unsqueeze(2), dense_k.permute(0, 2, 1, 3), dense_v.permute(0, 2, 1, 3),
        attn_mask=in_ctx[:, None, None, :], scale=scale,
    )
    return o.squeeze(2)


def prefill_dense(q, k, v, cu_seqlens_q, cu_seqlens_k, scale):
    h, kvh = q.size(1), k.size(1)
    pieces = []
    for s in range(cu_seqlens_q.numel() - 1):
        qa, qb = int(cu_seqlens_q[s]), int(cu_seqlens_q[s + 1])
        ka, kb = int(cu_seqlens_k[s]), int(cu_seqlens_k[s + 1])
        qs, ks, vs = q[qa:qb], k[ka:kb], v[ka:kb]
        if kvh != h:
            ks = ks.repeat_interleave(h // kvh, dim=1)
            vs = vs.repeat_interleave(h // kvh, dim=1)
        w = torch.matmul(qs.transpose(0, 1), ks.transpose(0, 1).transpose(-2, -1)) * scale
        lq, lk = qb - qa, kb - ka
        if lq > 1:
            hide = torch.triu(torch.ones(lq, lk, dtype=torch.bool), diagonal=lk - lq + 1)
            w = w.masked_fill(hide, float("-inf"))
        pieces.append(torch.matmul(torch.softmax(w, dim=-1), vs.transpose(0, 1)).transpose(0, 1))
    return torch.cat(pieces, dim=0)


def prefill_paged(q, k_cache, v_cache, block_tables, cu_seqlens_q, cu_seqlens_k, scale):
    """Prefix-aware prefill as the flash call site defines it (attention.py:396-433): gather each
    sequence's keys through its block table, then the same masked softmax as ``prefill_dense``."""
    bs = k_cache.size(1)
    ks, vs, cu = [], [], [0]
    for s in range(cu_seqlens_q.numel() - 1):
        lk = int(cu_seqlens_k[s + 1]) - int(cu_seqlens_k[s])
        pos = torch.arange(lk)
        page = block_tables[s, pos // bs].long()
        ks.append(k_cache[page, pos % bs])
        vs.append(v_cache[page, pos % bs])
        cu.append(cu[-1] + lk)
    return prefill_dense(q, torch.cat(ks), torch.cat(vs), cu_seqlens_q, torch.tensor(cu), scale)
