"""Plain-torch Qwen2/Qwen3 decoder hosting ``minivllm_b200.Attention`` (harness, not product).

Structure, parameter names and numerics follow the reference's ``QwenForCausalLM``
(minivllm/models/qwen_base.py:63-387): fused ``qkv_proj``/``gate_up_proj``, per-head q/k RMSNorm when
there is no qkv bias (:139-141), neox RoPE in fp32 (layers/rotary_embedding.py), fp32-internal RMSNorm with
fused residual add (layers/layernorm.py:81-93), head split by TP rank (:95-108), all-reduce after the
row-parallel ``o_proj``/``down_proj`` (layers/linear.py:316-319).
"""

from __future__ import annotations

import os
import zlib
from dataclasses import dataclass

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch import nn

from . import fused


@dataclass
class ModelConfig:
    name: str
    vocab_size: int
    hidden_size: int
    num_layers: int
    num_heads: int
    num_kv_heads: int
    head_dim: int
    intermediate_size: int
    qkv_bias: bool = False          # Qwen2: True (and no q/k norm); Qwen3: False (q/k norm)
    rms_eps: float = 1e-6
    rope_theta: float = 1e6
    max_position: int = 40960
    tie_embeddings: bool = True


PRESETS = {
    "qwen3-0.6b": ModelConfig("Qwen3-0.6B", 151936, 1024, 28, 16, 8, 128, 3072),
    "qwen3-4b": ModelConfig("Qwen3-4B", 151936, 2560, 36, 32, 8, 128, 9728),
    "qwen2-7b": ModelConfig("Qwen2-7B", 152064, 3584, 28, 28, 4, 128, 18944, qkv_bias=True, tie_embeddings=False),
    "qwen2-72b": ModelConfig("Qwen2-72B", 152064, 8192, 80, 64, 8, 128, 29568, qkv_bias=True, tie_embeddings=False),
    # what ONE rank of Qwen2-72B at TP=8 computes (heads, kv heads, MLP and vocabulary divided by 8): lets a single
    # GPU run cfg5's per-rank engine workload; only the hidden-size all-reduces are missing
    "qwen2-72b-rank8": ModelConfig("Qwen2-72B (one TP8 rank)", 19008, 8192, 80, 8, 1, 128, 3696, qkv_bias=True, tie_embeddings=False),
    # tiny fp32-able configs for parity tests (head_dim stays 128 so the fast kernels are exercised in bf16)
    "tiny-qwen3": ModelConfig("tiny-qwen3", 512, 256, 2, 4, 2, 128, 512, max_position=2048, tie_embeddings=False),
    "tiny-qwen2": ModelConfig("tiny-qwen2", 512, 256, 2, 4, 2, 64, 512, qkv_bias=True, max_position=2048, tie_embeddings=False),
}


def seeded_tensor(name: str, shape, seed: int = 1234) -> torch.Tensor:
    """Deterministic CPU init keyed by the parameter name (same values here and in the golden generator)."""
    g = torch.Generator().manual_seed(seed + zlib.crc32(name.encode()))
    if name.endswith("norm.weight"):
        return 1.0 + 0.05 * torch.randn(shape, generator=g)
    std = 0.02 if not name.endswith(".bias") else 0.02
    return std * torch.randn(shape, generator=g)


def rms_norm(x, weight, eps, residual=None):
    orig = x.dtype
    xf = x.float()
    if residual is not None:
        xf = xf + residual.float()
        residual = xf.to(orig)
    var = xf.pow(2).mean(dim=-1, keepdim=True)
    y = (xf * torch.rsqrt(var + eps) * weight).to(orig)
    return y, residual


class Rotary(nn.Module):
    def __init__(self, head_dim, max_position, base):
        super().__init__()
        inv = 1.0 / (base ** (torch.arange(0, head_dim, 2, dtype=torch.float) / head_dim))
        freqs = torch.einsum("i,j->ij", torch.arange(max_position, dtype=torch.float), inv)
        self.register_buffer("cos", freqs.cos(), persistent=False)
        self.register_buffer("sin", freqs.sin(), persistent=False)

    def forward(self, positions, q, k):
        cos = self.cos[positions].unsqueeze(1)
        sin = self.sin[positions].unsqueeze(1)

        def rot(x):
            x1, x2 = torch.chunk(x.float(), 2, dim=-1)
            return torch.cat((x1 * cos - x2 * sin, x2 * cos + x1 * sin), dim=-1).to(x.dtype)

        return rot(q), rot(k)


class DecoderLayer(nn.Module):
    def __init__(self, cfg: ModelConfig, tp: int, attention_cls, rotary):
        super().__init__()
        if cfg.num_heads % tp or cfg.num_kv_heads % tp:  # qwen_base.py:95-104
            raise ValueError(f"heads ({cfg.num_heads}/{cfg.num_kv_heads}) must be divisible by tp_size ({tp})")
        self.cfg, self.tp = cfg, tp
        self.h, self.kvh, self.d = cfg.num_heads // tp, cfg.num_kv_heads // tp, cfg.head_dim
        qkv_out = (self.h + 2 * self.kvh) * self.d
        self.qkv_w = nn.Parameter(torch.empty(qkv_out, cfg.hidden_size), requires_grad=False)
        self.qkv_b = nn.Parameter(torch.empty(qkv_out), requires_grad=False) if cfg.qkv_bias else None
        self.o_w = nn.Parameter(torch.empty(cfg.hidden_size, self.h * self.d), requires_grad=False)
        self.gate_up_w = nn.Parameter(torch.empty(2 * cfg.intermediate_size // tp, cfg.hidden_size), requires_grad=False)
        self.down_w = nn.Parameter(torch.empty(cfg.hidden_size, cfg.intermediate_size // tp), requires_grad=False)
        self.in_norm = nn.Parameter(torch.ones(cfg.hidden_size), requires_grad=False)
        self.post_norm = nn.Parameter(torch.ones(cfg.hidden_size), requires_grad=False)
        if not cfg.qkv_bias:
            self.q_norm = nn.Parameter(torch.ones(self.d), requires_grad=False)
            self.k_norm = nn.Parameter(torch.ones(self.d), requires_grad=False)
        else:
            self.q_norm = self.k_norm = None
        self.rotary = rotary  # shared cos/sin tables
        self.attn = attention_cls(self.h, self.d, self.d ** -0.5, self.kvh)
        self.comm = [None]  # shared one-element box, filled by QwenLM.enable_peer_allreduce
        self.fused_rope = os.environ.get("MVB200_HARNESS_UNFUSED", "0") in ("", "0")

    def _norm(self, x, weight, residual, reduce=False):
        """RMSNorm with residual; ``reduce``: ``x`` is this rank's partial sum of a row-parallel GEMM and is
        all-reduced first (linear.py:316-319) — for decode-size messages by the fused peer kernel."""
        if reduce and self.tp > 1:
            comm = self.comm[0]
            if comm is not None and x.is_cuda and residual is not None and comm.fits(x):
                # the residual stream is consumed by these calls only: leave it row-sharded (two-shot sizes)
                return comm.allreduce_add_rms_norm(x, weight, self.cfg.rms_eps, residual, gather_residual=False)
            dist.all_reduce(x)
        if self.fused_rope and x.is_cuda:  # harness/csrc/fused_elementwise.cu
            return fused.add_rms_norm(x, weight, self.cfg.rms_eps, residual)
        if residual is None:
            return rms_norm(x, weight, self.cfg.rms_eps)[0], x
        return rms_norm(x, weight, self.cfg.rms_eps, residual)

    def forward(self, positions, x, residual, pending_reduce=False):
        """``x`` arrives as the previous layer's un-reduced ``down_proj`` output when ``pending_reduce``;
        this layer's own ``down_proj`` output is returned un-reduced (the next norm reduces it)."""
        cfg = self.cfg
        x, residual = self._norm(x, self.in_norm, residual, reduce=pending_reduce)
        n = x.size(0)
        qkv = F.linear(x, self.qkv_w, self.qkv_b)
        q, k, v = qkv.split([self.h * self.d, self.kvh * self.d, self.kvh * self.d], dim=-1)
        q, k, v = q.view(n, self.h, self.d), k.view(n, self.kvh, self.d), v.view(n, self.kvh, self.d)
        if self.fused_rope and q.is_cuda and hasattr(self.attn, "forward_fused"):
            # one kernel for q/k norm + RoPE + cache store
            o = self.attn.forward_fused(positions, q, k, v, self.rotary.cos, self.rotary.sin,
                                        self.q_norm, self.k_norm, cfg.rms_eps).reshape(n, self.h * self.d)
        else:
            if self.q_norm is not None:
                q, _ = rms_norm(q, self.q_norm, cfg.rms_eps)
                k, _ = rms_norm(k, self.k_norm, cfg.rms_eps)
            q, k = self.rotary(positions, q, k)
            o = self.attn(q, k, v).reshape(n, self.h * self.d)
        o = F.linear(o, self.o_w)
        x, residual = self._norm(o, self.post_norm, residual, reduce=True)
        gu = F.linear(x, self.gate_up_w)
        if self.fused_rope and gu.is_cuda:
            act = fused.silu_mul(gu)
        else:
            g, u = gu.chunk(2, dim=-1)
            act = F.silu(g) * u
        return F.linear(act, self.down_w), residual


class QwenLM(nn.Module):
    def __init__(self, cfg: ModelConfig, attention_cls, tp: int = 1, rank: int = 0):
        super().__init__()
        self.cfg, self.tp, self.rank = cfg, tp, rank
        self.embed = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size), requires_grad=False)
        self.rotary = Rotary(cfg.head_dim, cfg.max_position, cfg.rope_theta)
        self.layers = nn.ModuleList([DecoderLayer(cfg, tp, attention_cls, self.rotary) for _ in range(cfg.num_layers)])
        self.final_norm = nn.Parameter(torch.ones(cfg.hidden_size), requires_grad=False)
        self.lm_head = None if cfg.tie_embeddings else nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size), requires_grad=False)

    def forward(self, input_ids, positions):
        x = F.embedding(input_ids, self.embed)
        residual = None
        for i, layer in enumerate(self.layers):
            x, residual = layer(positions, x, residual, pending_reduce=i > 0)
        return self.layers[-1]._norm(x, self.final_norm, residual, reduce=True)[0]

    def enable_peer_allreduce(self, max_rows: int, group=None):
        """TP > 1 on CUDA: route the decode-size all-reduce + add + RMSNorm through the one-shot NVLink
        peer kernel (minivllm_b200.comm).  Collective: call on every rank."""
        from minivllm_b200.comm import PeerAllReduce

        if self.tp == 1:
            return None
        # decode-size messages only (one-shot up to 128 KiB, two-shot above); prefill stays on NCCL
        nbytes = min(max_rows * self.cfg.hidden_size * self.embed.element_size(), 16 << 20)
        comm = PeerAllReduce(self.rank, self.tp, nbytes, group=group, device=self.embed.device)
        box = self.layers[0].comm
        box[0] = comm
        for layer in self.layers:
            layer.comm = box
        return comm

    def compute_logits(self, hidden):
        """Full-vocabulary logits (TP=1) — used by parity tests."""
        w = self.embed if self.lm_head is None else self.lm_head
        return F.linear(hidden, w)

    def greedy_tokens(self, hidden):
        """argmax over the vocabulary; vocab-sharded across TP ranks with a tiny all-gather."""
        w = self.embed if self.lm_head is None else self.lm_head
        if self.tp == 1:
            return F.linear(hidden, w).argmax(dim=-1)
        per = (self.cfg.vocab_size + self.tp - 1) // self.tp
        lo, hi = self.rank * per, min(self.cfg.vocab_size, (self.rank + 1) * per)
        val, idx = F.linear(hidden, w[lo:hi]).float().max(dim=-1)
        pack = torch.stack([val, (idx + lo).float()], dim=-1)
        allp = [torch.empty_like(pack) for _ in range(self.tp)]
        dist.all_gather(allp, pack)
        allp = torch.stack(allp)  # [tp, B, 2]
        win = allp[..., 0].argmax(dim=0)
        return allp[win, torch.arange(hidden.size(0), device=hidden.device), 1].long()

    def pick_tokens(self, hidden, out):
        """Greedy pick on the device into ``out`` (int64 [B]) with the mvb200 argmax kernel; capturable.
        Vocabulary-sharded across TP ranks: every rank reduces its slice, a tiny all-gather of (value,
        index) pairs picks the winner (lowest rank, hence lowest index, on ties)."""
        from minivllm_b200 import ops

        w = self.embed if self.lm_head is None else self.lm_head
        if self.tp == 1:
            ops.greedy_argmax(F.linear(hidden, w), out)
            return
        per = (self.cfg.vocab_size + self.tp - 1) // self.tp
        lo, hi = self.rank * per, min(self.cfg.vocab_size, (self.rank + 1) * per)
        b = hidden.size(0)
        val = torch.empty(b, dtype=torch.float32, device=hidden.device)
        idx = torch.empty(b, dtype=torch.int64, device=hidden.device)
        ops.greedy_argmax(F.linear(hidden, w[lo:hi]), idx, out_value=val, index_offset=lo)
        pack = torch.stack([val, idx.float()], dim=-1)
        allp = torch.empty(self.tp, b, 2, dtype=torch.float32, device=hidden.device)
        dist.all_gather_into_tensor(allp.view(self.tp * b, 2), pack)
        win = allp[..., 0].argmax(dim=0)
        out.copy_(allp[win, torch.arange(b, device=hidden.device), 1].long())

    # ---- weights -------------------------------------------------------------------------------
    def reference_names(self):
        """(harness parameter, reference parameter name, TP slicing rule) for every tensor."""
        out = [(self.embed, "model.embed_tokens.weight", None), (self.final_norm, "model.norm.weight", None)]
        if self.lm_head is not None:
            out.append((self.lm_head, "lm_head.weight", None))
        for i, l in enumerate(self.layers):
            p = f"model.layers.{i}."
            out += [(l.qkv_w, p + "self_attn.qkv_proj.weight", "qkv"), (l.o_w, p + "self_attn.o_proj.weight", "row"),
                    (l.gate_up_w, p + "mlp.gate_up_proj.weight", "gate_up"), (l.down_w, p + "mlp.down_proj.weight", "row"),
                    (l.in_norm, p + "input_layernorm.weight", None), (l.post_norm, p + "post_attention_layernorm.weight", None)]
            if l.qkv_b is not None:
                out.append((l.qkv_b, p + "self_attn.qkv_proj.bias", "qkv"))
            if l.q_norm is not None:
                out += [(l.q_norm, p + "self_attn.q_norm.weight", None), (l.k_norm, p + "self_attn.k_norm.weight", None)]
        return out

    def _slice(self, full, rule):
        cfg, tp, r = self.cfg, self.tp, self.rank
        if rule is None or tp == 1:
            return full
        d = cfg.head_dim
        if rule == "qkv":
            qs, ks = cfg.num_heads * d, cfg.num_kv_heads * d
            q, k, v = full[:qs], full[qs:qs + ks], full[qs + ks:]
            return torch.cat([t.chunk(tp, dim=0)[r] for t in (q, k, v)], dim=0)
        if rule == "gate_up":
            g, u = full.chunk(2, dim=0)
            return torch.cat([g.chunk(tp, dim=0)[r], u.chunk(tp, dim=0)[r]], dim=0)
        if rule == "row":
            return full.chunk(tp, dim=1)[r]
        raise ValueError(rule)

    def full_shape(self, param, rule):
        cfg, d = self.cfg, self.cfg.head_dim
        if rule == "qkv":
            n = (cfg.num_heads + 2 * cfg.num_kv_heads) * d
            return (n,) + tuple(param.shape[1:])
        if rule == "gate_up":
            return (2 * cfg.intermediate_size, cfg.hidden_size)
        if rule == "row":
            return (param.shape[0], param.shape[1] * self.tp)
        return tuple(param.shape)

    @torch.no_grad()
    def init_weights(self, mode: str = "seeded", seed: int = 1234):
        """``seeded``: CPU, keyed by reference parameter name (bit-identical to the golden generator);
        ``fast``: on-device normal init (benchmarks)."""
        pos = torch.arange(self.cfg.max_position, dtype=torch.float, device=self.rotary.cos.device)
        inv = 1.0 / (self.cfg.rope_theta ** (torch.arange(0, self.cfg.head_dim, 2, dtype=torch.float, device=pos.device) / self.cfg.head_dim))
        freqs = torch.einsum("i,j->ij", pos, inv)
        self.rotary.cos = freqs.cos()  # keep the tables fp32 whatever the model dtype
        self.rotary.sin = freqs.sin()
        for param, name, rule in self.reference_names():
            shape = self.full_shape(param, rule)
            if mode == "seeded":
                full = seeded_tensor(name, shape, seed)
                param.copy_(self._slice(full, rule).to(param.dtype))
            else:
                if name.endswith("norm.weight"):
                    param.fill_(1.0)
                else:
                    param.normal_(0.0, 0.02)
        return self
