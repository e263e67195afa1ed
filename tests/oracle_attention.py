"""Test double: the Attention layer computed by the CPU oracle port (tests only).

Lets the harness (model + engine + executor) run on a CPU-only box so that it can be compared
with the reference engine; the product layer ``minivllm_b200.Attention`` has no CPU path.
With ``block_tables`` present, prefill keys are read back through them (flash call-site semantics,
attention.py:399-400) — which is also what makes prefix-cache hits correct.
"""

import math

import torch
from torch import nn

from minivllm_b200 import get_context
from oracle import eager_port as port


class OracleAttention(nn.Module):
    def __init__(self, num_heads, head_dim, scale, num_kv_heads, use_buffered_page_attention=False):
        super().__init__()
        self.num_heads, self.head_dim, self.num_kv_heads = num_heads, head_dim, num_kv_heads
        self.scale = float(scale) if scale != 0.0 else 1.0 / math.sqrt(head_dim)
        self.k_cache = torch.tensor([])
        self.v_cache = torch.tensor([])

    def forward(self, q, k, v):
        ctx = get_context()
        dev = q.device
        q, k, v = q.cpu(), k.cpu(), v.cpu()
        has_cache = self.k_cache.numel() > 0
        kc, vc = (self.k_cache.cpu(), self.v_cache.cpu()) if has_cache else (None, None)
        if has_cache and ctx.slot_mapping is not None and ctx.slot_mapping.numel() > 0:
            port.store_kvcache(k, v, kc, vc, ctx.slot_mapping.cpu().long())
            if self.k_cache.device.type != "cpu":
                self.k_cache.copy_(kc)
                self.v_cache.copy_(vc)
        if ctx.is_prefill:
            cq, ck = ctx.cum_seqlens_q.cpu(), ctx.cum_seqlens_k.cpu()
            if has_cache and ctx.block_tables is not None and ctx.block_tables.numel() > 0:
                out = port.prefill_paged(q, kc, vc, ctx.block_tables.cpu(), cq, ck, self.scale)
            else:
                out = port.prefill_dense(q, k, v, cq, ck, self.scale)
        elif ctx.block_tables.size(1) == 0:
            out = torch.zeros_like(q)
        else:
            out = port.paged_decode(q, kc, vc, ctx.block_tables.cpu(), ctx.context_lens.cpu(), self.scale)
        return out.to(dev)
