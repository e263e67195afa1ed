#!/usr/bin/env python
"""Generate golden input/output vectors by running the UNMODIFIED reference on CPU.

Run in the build container only (``/root/reference`` does not exist on the GPU
box); the ``.npz`` files it writes next to itself are committed and are what the
tests read.  Usage::

    python tests/golden/gen_golden.py

The reference needs: a ``colorama`` stand-in (its only missing dependency),
``flash_attn`` hidden so that ``Attention.forward`` takes the eager branch
(attention.py:80-87,450-453), and ``MINIVLLM_DEVICE=cpu``.

Every case is seeded (``torch.Generator().manual_seed``).  bf16 tensors are
stored as uint16 bit patterns.  What is recorded:

* ``store_*``   : ``StandardAttentionBackend.store_kv_cache`` (attention_backend.py:196-266)
* ``decode_*``  : ``PageAttention.__call__`` (page_attention.py:35-133); the five
                  fp32 shapes are the reference's own tests/test_page_attention.py:95-276
* ``prefill_*`` : ``Attention._fallback_attention`` prefill branch (attention.py:467-504)
* ``layer_*``   : ``Attention.forward`` prefill then decode steps through the
                  real ``Context`` (attention.py:196-455), caches included
* ``qkrope``    : ``RMSNorm`` on q/k + ``RotaryEmbedding`` + store, the caller-side sequence of
                  ``QwenAttention.forward`` (qwen_base.py:175-186)
"""

from __future__ import annotations

import os
import sys
import tempfile
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE = os.environ.get("MINIVLLM_REFERENCE", "/root/reference")


def _import_reference():
    os.environ.setdefault("MINIVLLM_DEVICE", "cpu")
    os.environ["CUDA_VISIBLE_DEVICES"] = ""
    shim = types.ModuleType("colorama")

    class _Blank:
        def __getattr__(self, _name):
            return ""

    shim.Fore = shim.Style = shim.Back = _Blank()
    shim.init = lambda *a, **k: None
    sys.modules.setdefault("colorama", shim)
    sys.modules["flash_attn"] = None  # force the eager branch
    sys.path.insert(0, REFERENCE)
    import minivllm  # noqa: F401


_import_reference()

import numpy as np  # noqa: E402
import torch  # noqa: E402
from minivllm.models.layers.attention import Attention  # noqa: E402
from minivllm.models.layers.attention_backend import StandardAttentionBackend  # noqa: E402
from minivllm.models.layers.page_attention import PageAttention  # noqa: E402
from minivllm.utils.context import reset_context, set_context  # noqa: E402

torch.set_num_threads(4)


def to_np(t: torch.Tensor) -> np.ndarray:
    t = t.detach().contiguous()
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).numpy().view(np.uint16).copy()
    if t.dtype == torch.float16:
        return t.view(torch.int16).numpy().view(np.float16).copy()
    return t.numpy().copy()


def randn(gen, *shape, dtype=torch.float32):
    return torch.randn(*shape, generator=gen, dtype=torch.float32).to(dtype)


def dtype_name(dt):
    return {torch.float32: "f32", torch.bfloat16: "bf16", torch.float16: "f16"}[dt]


def save(name, **arrays):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **arrays)
    print(f"  wrote {name}.npz  ({os.path.getsize(path) / 1024:.0f} KiB)")


# ---------------------------------------------------------------------------
# store
# ---------------------------------------------------------------------------


def gen_store():
    backend = StandardAttentionBackend()
    out = {}
    cases = [
        # name, N, KVH, D, num_slots, dtype, slots-kind, strided
        ("ref_basic", 2, 4, 16, 100, torch.float32, [10, 20], False),
        ("ref_noncontig", 3, 4, 16, 50, torch.float32, [0, 25, 49], False),
        ("bf16_prefill", 150, 2, 128, 256, torch.bfloat16, "prefill", True),
        ("bf16_decode_pad", 16, 8, 128, 96, torch.bfloat16, "decode_pad", True),
        ("f16_opt", 37, 12, 64, 128, torch.float16, "perm", False),
        ("f32_odd", 9, 3, 20, 64, torch.float32, "perm", True),
    ]
    for idx, (name, n, kvh, d, slots_n, dt, kind, strided) in enumerate(cases):
        g = torch.Generator().manual_seed(100 + idx)
        if strided:
            # k and v as row-strided views of one fused projection output
            fused = randn(g, n, 3 * kvh * d + 8, dtype=dt)
            key = fused[:, kvh * d : 2 * kvh * d].view(n, kvh, d)
            value = fused[:, 2 * kvh * d : 3 * kvh * d].view(n, kvh, d)
        else:
            key = randn(g, n, kvh, d, dtype=dt)
            value = randn(g, n, kvh, d, dtype=dt)
        if isinstance(kind, list):
            slots = torch.tensor(kind, dtype=torch.int64)
        elif kind == "prefill":
            # blocks 3,1,2 of 64 slots, last partial: 64+64+22 = 150 tokens
            slots = torch.tensor(
                list(range(192, 256)) + list(range(64, 128)) + list(range(128, 150)),
                dtype=torch.int32,
            )
        elif kind == "decode_pad":
            perm = torch.randperm(slots_n, generator=g)[:n].to(torch.int32)
            perm[n - 5 :] = -1  # graph padding rows
            perm[2] = -1
            slots = perm
        else:
            slots = torch.randperm(slots_n, generator=g)[:n].to(torch.int32)
        k_cache = randn(g, slots_n, kvh * d, dtype=dt)
        v_cache = randn(g, slots_n, kvh * d, dtype=dt)
        out[f"{name}.key"] = to_np(key)
        out[f"{name}.value"] = to_np(value)
        out[f"{name}.slots"] = to_np(slots)
        out[f"{name}.k_cache_in"] = to_np(k_cache)
        out[f"{name}.v_cache_in"] = to_np(v_cache)
        backend.store_kv_cache(key, value, k_cache, v_cache, slots)
        out[f"{name}.k_cache_out"] = to_np(k_cache)
        out[f"{name}.v_cache_out"] = to_np(v_cache)
        out[f"{name}.meta"] = np.array([n, kvh, d, slots_n, int(strided)])
        out[f"{name}.dtype"] = np.array(dtype_name(dt))
    out["cases"] = np.array([c[0] for c in cases])
    save("store", **out)


# ---------------------------------------------------------------------------
# decode
# ---------------------------------------------------------------------------


def build_paged(g, seqlens, block_tables, bs, kvh, d, dt, num_blocks=None):
    """Scatter per-sequence contiguous K/V into a paged cache (garbage elsewhere)."""
    bt = torch.tensor(block_tables, dtype=torch.int32)
    nb = num_blocks or int(bt.max().item()) + 1
    # unwritten slots hold finite garbage so that masking mistakes show up
    k_cache = randn(g, nb, bs, kvh, d, dtype=dt) * 3
    v_cache = randn(g, nb, bs, kvh, d, dtype=dt) * 3
    for b, sl in enumerate(seqlens):
        kk = randn(g, sl, kvh, d, dtype=dt)
        vv = randn(g, sl, kvh, d, dtype=dt)
        for p in range(sl):
            blk = int(bt[b, p // bs])
            k_cache[blk, p % bs] = kk[p]
            v_cache[blk, p % bs] = vv[p]
    return k_cache, v_cache, bt


def gen_decode():
    pa = PageAttention()
    out = {}
    cases = [
        # name, H, KVH, D, BS, seqlens, block_tables, dtype
        ("ref_single_block", 8, 8, 64, 16, [12], [[0]], torch.float32),
        ("ref_multi_block", 8, 8, 64, 4, [10], [[5, 2, 7]], torch.float32),
        ("ref_multi_seq", 8, 8, 64, 8, [15, 8, 20], [[3, 1, 0], [2, -1, -1], [5, 4, 6]], torch.float32),
        ("ref_gqa", 8, 2, 64, 8, [16, 10], [[0, 1], [2, 3]], torch.float32),
        ("ref_varying", 4, 4, 32, 4, [4, 7, 3], [[0, 1], [2, 3], [4, -1]], torch.float32),
        # engine-shaped: BS=64, D=128 (Qwen) / 64 (OPT), bf16
        ("qwen3_g2", 4, 2, 128, 64, [1, 63, 64, 65, 300], None, torch.bfloat16),
        ("qwen3_g4", 8, 2, 128, 64, [200, 129, 5], None, torch.bfloat16),
        ("qwen2_g7", 7, 1, 128, 64, [257, 64, 33], None, torch.bfloat16),
        ("qwen2_g8", 8, 1, 128, 64, [513, 100], None, torch.bfloat16),
        ("opt_mha", 4, 4, 64, 64, [70, 3, 128], None, torch.float16),
        ("opt_mha_f32", 4, 4, 64, 64, [70, 3, 128], None, torch.float32),
    ]
    for idx, (name, h, kvh, d, bs, seqlens, bt, dt) in enumerate(cases):
        g = torch.Generator().manual_seed(200 + idx)
        if bt is None:
            # non-contiguous physical pages, -1 padded like _prepare_block_tables
            nblk = [(s + bs - 1) // bs for s in seqlens]
            total = sum(nblk) + 3
            perm = torch.randperm(total, generator=g).tolist()
            width = max(nblk)
            bt, c = [], 0
            for nb in nblk:
                bt.append(perm[c : c + nb] + [-1] * (width - nb))
                c += nb
            num_blocks = total
        else:
            num_blocks = None
        k_cache, v_cache, btt = build_paged(g, seqlens, bt, bs, kvh, d, dt, num_blocks)
        q = randn(g, len(seqlens), h, d, dtype=dt)
        ctx = torch.tensor(seqlens, dtype=torch.int32)
        scale = 1.0 / (d**0.5)
        ref = pa(q, k_cache, v_cache, btt, ctx, scale)
        out[f"{name}.q"] = to_np(q)
        out[f"{name}.k_cache"] = to_np(k_cache)
        out[f"{name}.v_cache"] = to_np(v_cache)
        out[f"{name}.block_tables"] = to_np(btt)
        out[f"{name}.context_lens"] = to_np(ctx)
        out[f"{name}.out"] = to_np(ref)
        out[f"{name}.scale"] = np.array(scale)
        out[f"{name}.dtype"] = np.array(dtype_name(dt))
    out["cases"] = np.array([c[0] for c in cases])
    save("decode", **out)


# ---------------------------------------------------------------------------
# dense prefill (no prefix hits): Attention._fallback_attention
# ---------------------------------------------------------------------------


def gen_prefill():
    out = {}
    cases = [
        ("f32_gqa4", 8, 2, 64, [5, 17, 1], torch.float32),
        ("f32_mha", 4, 4, 32, [33, 2], torch.float32),
        ("bf16_qwen3", 4, 2, 128, [130, 64, 257, 1], torch.bfloat16),
        ("bf16_g7", 7, 1, 128, [129, 200], torch.bfloat16),
        ("f16_opt", 4, 4, 64, [100, 31], torch.float16),
    ]
    for idx, (name, h, kvh, d, seqlens, dt) in enumerate(cases):
        g = torch.Generator().manual_seed(300 + idx)
        n = sum(seqlens)
        q = randn(g, n, h, d, dtype=dt)
        k = randn(g, n, kvh, d, dtype=dt)
        v = randn(g, n, kvh, d, dtype=dt)
        cu = torch.tensor([0] + list(np.cumsum(seqlens)), dtype=torch.int32)
        scale = 1.0 / (d**0.5)
        attn = Attention(h, d, scale, kvh)
        set_context(
            is_prefill=True,
            max_seqlen_q=max(seqlens),
            max_seqlen_k=max(seqlens),
            cum_seqlens_q=cu,
            cum_seqlens_k=cu,
        )
        try:
            with torch.no_grad():
                ref = attn(q, k, v)  # caches empty -> no store; eager prefill
        finally:
            reset_context()
        out[f"{name}.q"] = to_np(q)
        out[f"{name}.k"] = to_np(k)
        out[f"{name}.v"] = to_np(v)
        out[f"{name}.cu_seqlens"] = to_np(cu)
        out[f"{name}.out"] = to_np(ref)
        out[f"{name}.scale"] = np.array(scale)
        out[f"{name}.dtype"] = np.array(dtype_name(dt))
    out["cases"] = np.array([c[0] for c in cases])
    save("prefill", **out)


# ---------------------------------------------------------------------------
# whole layer: prefill (store + attention) then decode steps through Context
# ---------------------------------------------------------------------------


def gen_layer():
    out = {}
    cases = [
        ("bf16_qwen3", 4, 2, 128, [70, 130, 64], 3, torch.bfloat16),
        ("f32_opt", 4, 4, 64, [20, 65], 2, torch.float32),
    ]
    bs = 64
    for idx, (name, h, kvh, d, seqlens, steps, dt) in enumerate(cases):
        g = torch.Generator().manual_seed(400 + idx)
        scale = 1.0 / (d**0.5)
        attn = Attention(h, d, scale, kvh)
        final = [s + steps for s in seqlens]
        nblk = [(s + bs - 1) // bs for s in final]
        total = sum(nblk) + 2
        perm = torch.randperm(total, generator=g).tolist()
        tables, c = [], 0
        for nb in nblk:
            tables.append(perm[c : c + nb])
            c += nb
        attn.k_cache = randn(g, total, bs, kvh, d, dtype=dt) * 2
        attn.v_cache = randn(g, total, bs, kvh, d, dtype=dt) * 2
        out[f"{name}.k_cache_in"] = to_np(attn.k_cache)
        out[f"{name}.v_cache_in"] = to_np(attn.v_cache)
        # ---- prefill
        n = sum(seqlens)
        q = randn(g, n, h, d, dtype=dt)
        k = randn(g, n, kvh, d, dtype=dt)
        v = randn(g, n, kvh, d, dtype=dt)
        cu = torch.tensor([0] + list(np.cumsum(seqlens)), dtype=torch.int32)
        slots = []
        for sl, tb in zip(seqlens, tables):
            for p in range(sl):
                slots.append(tb[p // bs] * bs + p % bs)
        slots_t = torch.tensor(slots, dtype=torch.int32)
        width = max((s + bs - 1) // bs for s in seqlens)
        bt = torch.tensor(
            [tb[: (s + bs - 1) // bs] + [-1] * (width - (s + bs - 1) // bs) for s, tb in zip(seqlens, tables)],
            dtype=torch.int32,
        )
        set_context(
            is_prefill=True,
            max_seqlen_q=max(seqlens),
            max_seqlen_k=max(seqlens),
            cum_seqlens_q=cu,
            cum_seqlens_k=cu,
            slot_mapping=slots_t,
            block_tables=bt,
        )
        try:
            with torch.no_grad():
                o = attn(q, k, v)
        finally:
            reset_context()
        out[f"{name}.p.q"] = to_np(q)
        out[f"{name}.p.k"] = to_np(k)
        out[f"{name}.p.v"] = to_np(v)
        out[f"{name}.p.cu_seqlens"] = to_np(cu)
        out[f"{name}.p.slots"] = to_np(slots_t)
        out[f"{name}.p.block_tables"] = to_np(bt)
        out[f"{name}.p.out"] = to_np(o)
        out[f"{name}.p.k_cache_out"] = to_np(attn.k_cache)
        out[f"{name}.p.v_cache_out"] = to_np(attn.v_cache)
        # ---- decode steps
        lens = list(seqlens)
        b = len(lens)
        for st in range(steps):
            lens = [x + 1 for x in lens]
            q = randn(g, b, h, d, dtype=dt)
            k = randn(g, b, kvh, d, dtype=dt)
            v = randn(g, b, kvh, d, dtype=dt)
            slots_t = torch.tensor(
                [tb[(x - 1) // bs] * bs + (x - 1) % bs for x, tb in zip(lens, tables)],
                dtype=torch.int32,
            )
            width = max((x + bs - 1) // bs for x in lens)
            bt = torch.tensor(
                [tb[: (x + bs - 1) // bs] + [-1] * (width - (x + bs - 1) // bs) for x, tb in zip(lens, tables)],
                dtype=torch.int32,
            )
            ctx = torch.tensor(lens, dtype=torch.int32)
            set_context(is_prefill=False, slot_mapping=slots_t, context_lens=ctx, block_tables=bt)
            try:
                with torch.no_grad():
                    o = attn(q, k, v)
            finally:
                reset_context()
            pre = f"{name}.d{st}"
            out[f"{pre}.q"] = to_np(q)
            out[f"{pre}.k"] = to_np(k)
            out[f"{pre}.v"] = to_np(v)
            out[f"{pre}.slots"] = to_np(slots_t)
            out[f"{pre}.block_tables"] = to_np(bt)
            out[f"{pre}.context_lens"] = to_np(ctx)
            out[f"{pre}.out"] = to_np(o)
        out[f"{name}.k_cache_final"] = to_np(attn.k_cache)
        out[f"{name}.v_cache_final"] = to_np(attn.v_cache)
        out[f"{name}.meta"] = np.array([h, kvh, d, steps, len(seqlens)])
        out[f"{name}.scale"] = np.array(scale)
        out[f"{name}.dtype"] = np.array(dtype_name(dt))
    out["cases"] = np.array([c[0] for c in cases])
    save("layer", **out)


# ---------------------------------------------------------------------------
# attention metadata: the reference's own Sequence + BlockManager + executor prep
# ---------------------------------------------------------------------------


def gen_metadata():
    import types

    from minivllm.engine.block_manager import BlockManager
    from minivllm.engine.inference_executor import InferenceExecutor
    from minivllm.engine.sequence import Sequence
    from minivllm.utils.context import get_context

    bs, max_seqs, max_model_len = 64, 8, 1024
    fake = types.SimpleNamespace(block_size=bs, device=torch.device("cpu"))
    fake.config = types.SimpleNamespace(max_num_seqs=max_seqs, max_model_len=max_model_len, kvcache_block_size=bs)
    for name in ("_list_to_device_tensor", "_prepare_block_tables", "_prepare_block_tables_decode"):
        setattr(fake, name, types.MethodType(getattr(InferenceExecutor, name), fake))
    mb = (max_model_len + bs - 1) // bs
    fake._decode_input_ids_cpu = torch.zeros(max_seqs, dtype=torch.long)
    fake._decode_positions_cpu = torch.zeros(max_seqs, dtype=torch.long)
    fake._decode_slot_mapping_cpu = torch.zeros(max_seqs, dtype=torch.int32)
    fake._decode_context_lens_cpu = torch.zeros(max_seqs, dtype=torch.int32)
    fake._decode_input_ids = torch.zeros(max_seqs, dtype=torch.long)
    fake._decode_positions = torch.zeros(max_seqs, dtype=torch.long)
    fake._decode_slot_mapping = torch.zeros(max_seqs, dtype=torch.int32)
    fake._decode_context_lens = torch.zeros(max_seqs, dtype=torch.int32)
    fake._decode_block_tables_cpu = torch.full((max_seqs, mb), -1, dtype=torch.int32)
    fake._decode_block_tables = torch.full((max_seqs, mb), -1, dtype=torch.int32)
    fake._max_blocks_per_seq = mb

    g = torch.Generator().manual_seed(500)
    rnd = lambda n: torch.randint(10, 30000, (n,), generator=g).tolist()  # noqa: E731
    bm = BlockManager(40, bs)
    out = {}
    steps = []

    def record(tag, seqs, prefill):
        if prefill:
            ids, pos = InferenceExecutor._prepare_prefill_input(fake, seqs)
        else:
            ids, pos = InferenceExecutor._prepare_decode_input(fake, seqs)
        ctx = get_context()
        out[f"{tag}.n"] = np.array(len(seqs))
        for i, s in enumerate(seqs):
            out[f"{tag}.seq{i}.token_ids"] = np.array(s.token_ids, dtype=np.int64)
            out[f"{tag}.seq{i}.block_table"] = np.array(s.block_table, dtype=np.int64)
            out[f"{tag}.seq{i}.num_cached"] = np.array(s.num_cached_tokens)
        out[f"{tag}.input_ids"] = to_np(ids)
        out[f"{tag}.positions"] = to_np(pos)
        out[f"{tag}.is_prefill"] = np.array(int(prefill))
        for f in ("cum_seqlens_q", "cum_seqlens_k", "slot_mapping", "context_lens", "block_tables"):
            v = getattr(ctx, f)
            if v is not None:
                out[f"{tag}.{f}"] = to_np(v.clone())
        out[f"{tag}.max_seqlen_q"] = np.array(ctx.max_seqlen_q)
        out[f"{tag}.max_seqlen_k"] = np.array(ctx.max_seqlen_k)
        steps.append(tag)
        reset_context()

    base = rnd(254)
    a = Sequence(base, block_size=bs)
    bm.allocate(a)
    record("p0", [a], True)
    b = Sequence(base[:128] + rnd(70), block_size=bs)   # two full blocks hit the prefix cache
    c = Sequence(rnd(64), block_size=bs)                # exactly one block
    d = Sequence(base[:192], block_size=bs)             # fully cached prompt: Lq = 0 (block_manager.py:324-327)
    for s in (b, c, d):
        bm.allocate(s)
    assert b.num_cached_tokens == 128 and d.num_cached_tokens == 192
    record("p1", [b, c, d], True)
    live = [a, b, c]
    for t in range(4):  # a: 255, 256 (hash sealed), 257 (new block), 258; c: 65 (new block) ...
        for s in live:
            s.append_token(int(torch.randint(10, 30000, (1,), generator=g)))
            bm.may_append(s)
        record(f"d{t}", live, False)
    # warm-up shaped batches: no block tables at all (inference_executor.py:390-407)
    w = [Sequence([0] * 16, block_size=bs) for _ in range(2)]
    record("w_prefill", w, True)
    for s in w:
        s.append_token(0)
    record("w_decode", w, False)
    out["steps"] = np.array(steps)
    save("metadata", **out)


# ---------------------------------------------------------------------------
# end to end: the reference's LLM.generate on CPU (fp32, eager, random-init tiny Qwen3 / Qwen2)
# ---------------------------------------------------------------------------


def gen_engine():
    import json
    import tempfile
    from unittest import mock

    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from harness.model import PRESETS, seeded_tensor

    import minivllm.models.manager as manager
    from minivllm import LLM, SamplingParams
    from minivllm.config import Config

    def seeded_load(model, path):  # stands in for the safetensors loader (utils/loader.py:90-239)
        with torch.no_grad():
            for name, param in model.named_parameters():
                param.copy_(seeded_tensor(name, tuple(param.shape)).to(param.dtype))

    class Tok:  # manager.py:97-106, llm_engine.py:107-108,359-361
        eos_token_id = 1
        pad_token = "<pad>"
        pad_token_id = 0

        def batch_decode(self, ids, **kw):
            return ["" for _ in ids]

        def decode(self, ids, **kw):
            return ""

    out = {}
    names = []
    for preset, arch, mtype in (("tiny-qwen3", "Qwen3ForCausalLM", "qwen3"), ("tiny-qwen2", "Qwen2ForCausalLM", "qwen2")):
        cfg = PRESETS[preset]
        tmp = tempfile.mkdtemp()
        json.dump({
            "architectures": [arch], "model_type": mtype, "hidden_size": cfg.hidden_size,
            "num_hidden_layers": cfg.num_layers, "num_attention_heads": cfg.num_heads,
            "num_key_value_heads": cfg.num_kv_heads, "head_dim": cfg.head_dim,
            "intermediate_size": cfg.intermediate_size, "vocab_size": cfg.vocab_size,
            "rms_norm_eps": cfg.rms_eps, "rope_theta": cfg.rope_theta,
            "max_position_embeddings": cfg.max_position, "tie_word_embeddings": cfg.tie_embeddings,
            "hidden_act": "silu", "attention_bias": cfg.qkv_bias, "torch_dtype": "float32",
        }, open(os.path.join(tmp, "config.json"), "w"))
        g = torch.Generator().manual_seed(600)
        # distinct prompts, ONE call: the reference's eager prefill mis-slices k/v as soon as any
        # sequence in the batch has a prefix hit, so hits are kept out of the
        # golden run; prefix-hit generation is checked by self-consistency in tests/test_harness.py.
        lens = [165, 13, 203, 64, 250]  # 250 crosses a block boundary while decoding
        prompts = [torch.randint(10, cfg.vocab_size - 10, (n,), generator=g).tolist() for n in lens]
        with mock.patch.object(manager, "load_model", seeded_load), \
                mock.patch.object(manager.AutoTokenizer, "from_pretrained", lambda *a, **k: Tok()):
            llm = LLM(Config(model=tmp, max_num_seqs=8, max_model_len=1024, enforce_eager=True, dtype="float32",
                             num_kvcache_blocks=64))
            res = llm.generate(prompts, SamplingParams(temperature=0.0, max_tokens=32, ignore_eos=True), use_tqdm=False)
            assert llm.scheduler.block_manager.stats["cache_hits"] == 0
        out[f"{preset}.n"] = np.array(len(prompts))
        for i, (p, r) in enumerate(zip(prompts, res)):
            out[f"{preset}.prompt{i}"] = np.array(p, dtype=np.int64)
            out[f"{preset}.tokens{i}"] = np.array(r["token_ids"], dtype=np.int64)
        names.append(preset)
        print("   ", preset, [r["token_ids"][:6] for r in res])
    out["presets"] = np.array(names)
    save("engine", **out)


# ---------------------------------------------------------------------------
# end to end on the NAMED architectures: the reference's LLM.generate, eager, on CPU
# ---------------------------------------------------------------------------

ENGINE_NAMED_CASES = [
    # tag, preset, dtype, prompt lengths, config overrides
    ("qwen3-0.6b.float32", "qwen3-0.6b", "float32", [165, 13, 203, 64, 250], {}),
    # same dims with an untied lm_head: with tied random embeddings the argmax is mostly the input token
    # (the stream is still recorded above, it is the named config); untied, every token depends on the layers
    ("qwen3-0.6b-untied.float32", "qwen3-0.6b", "float32", [165, 13, 203, 64, 250], {"tie_word_embeddings": False}),
    ("qwen3-0.6b-untied.bfloat16", "qwen3-0.6b", "bfloat16", [165, 13, 203, 64, 250], {"tie_word_embeddings": False}),
    # cfg1: examples/cpu_inference_opt.py:143-151 (max_num_seqs=8, max_model_len=1024, eager, float32), OPT-125M dims
    ("opt-125m.float32", "opt-125m", "float32", [165, 13, 203, 64, 250, 511], {}),
]


def engine_named_prompts(vocab: int, lens):
    """Call A: distinct prompts.  Call B: two prompts that share 128 / 192 leading tokens with A[0] / A[2]
    (two / three full blocks) — issued after A on one engine they hit the prefix cache."""
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from harness.ref_engine import make_prompts

    a = make_prompts(vocab, lens, seed=600)
    tails = make_prompts(vocab, [70, 30], seed=601)
    b = [a[0][:128] + tails[0], a[2][:192] + tails[1]]
    return a, b


def gen_engine_named():
    """``LLM.generate`` (llm_engine.py:261-367) with ``enforce_eager=True`` on CPU.  Call B is recorded on a
    FRESH engine (no prefix hit): the reference's eager prefill mis-slices K/V under a hit,
    and a prompt's greedy stream does not depend on whether its prefix was cached — so the no-hit stream is
    the truth the GPU run (which does hit) is compared with."""
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from harness import ref_engine
    from oracle.reference_engine import PRESETS

    out = {}
    tags = []
    for tag, preset, dtype, lens, overrides in ENGINE_NAMED_CASES:
        a, b = engine_named_prompts(PRESETS[preset]["vocab_size"], lens)
        kw = dict(device="cpu", dtype=dtype, max_tokens=32, record_logits=True, config_overrides=overrides,
                  max_num_seqs=8, max_model_len=1024, num_kvcache_blocks=64)
        for call, prompts in (("a", a), ("b", b)):
            info = ref_engine.run(preset, calls=[prompts], **kw)
            res = info["calls"][0]
            assert res["cache_hits"] == 0
            out[f"{tag}.{call}.n"] = np.array(len(prompts))
            for i, p in enumerate(prompts):
                out[f"{tag}.{call}.prompt{i}"] = np.array(p, dtype=np.int64)
                out[f"{tag}.{call}.tokens{i}"] = np.array(res["tokens"][i], dtype=np.int64)
                out[f"{tag}.{call}.top2_{i}"] = np.array(res["top2"][i], dtype=np.float64)  # [steps, (v1, v2, id1, id2)]
            print("   ", tag, call, [t[:6] for t in res["tokens"]])
        tags.append(tag)
    out["cases"] = np.array(tags)
    save("engine_named", **out)


# ---------------------------------------------------------------------------
# caller side: q/k RMSNorm -> RoPE -> store, as QwenAttention.forward does it
# ---------------------------------------------------------------------------


def gen_qkrope():
    """``QwenAttention.forward`` between the qkv GEMM and ``self.attn`` (qwen_base.py:175-186):
    ``RMSNorm`` on q and k (layernorm.py:81-93), ``RotaryEmbedding.forward`` (rotary_embedding.py:166-193),
    then the store of the new K/V rows that ``Attention.forward`` performs first (attention.py:228-236)."""
    from minivllm.models.layers.layernorm import RMSNorm
    from minivllm.models.layers.rotary_embedding import RotaryEmbedding

    backend = StandardAttentionBackend()
    out = {}
    cases = [
        # name, N, H, KVH, D, dtype, qk-norm, max_pos, rope base, num_slots
        ("qwen3_bf16", 37, 4, 2, 128, torch.bfloat16, True, 256, 1000000.0, 128),
        ("qwen2_bf16_nonorm", 20, 7, 1, 128, torch.bfloat16, False, 256, 1000000.0, 64),
        ("f16_norm", 11, 4, 2, 128, torch.float16, True, 128, 10000.0, 64),
        ("f32_d64", 9, 3, 3, 64, torch.float32, True, 64, 10000.0, 32),
    ]
    for idx, (name, n, h, kvh, d, dt, norm, max_pos, base, slots_n) in enumerate(cases):
        g = torch.Generator().manual_seed(700 + idx)
        fused = randn(g, n, (h + 2 * kvh) * d, dtype=dt) * 2.0
        q, k, v = fused.split([h * d, kvh * d, kvh * d], dim=-1)
        q, k, v = q.view(n, h, d), k.view(n, kvh, d), v.view(n, kvh, d)
        positions = torch.randint(0, max_pos, (n,), generator=g, dtype=torch.int64)
        slots = torch.randperm(slots_n, generator=g)[:n].to(torch.int32)
        slots[1] = -1
        rope = RotaryEmbedding(d, d, max_pos, base)
        qn, kn = q, k
        if norm:
            q_norm, k_norm = RMSNorm(d, eps=1e-6), RMSNorm(d, eps=1e-6)
            q_norm.weight.data = (1.0 + 0.1 * randn(g, d)).to(dt)
            k_norm.weight.data = (1.0 + 0.1 * randn(g, d)).to(dt)
            qn, _ = q_norm(q)
            kn, _ = k_norm(k)
            out[f"{name}.q_weight"] = to_np(q_norm.weight.data)
            out[f"{name}.k_weight"] = to_np(k_norm.weight.data)
        qr, kr = rope(positions, qn, kn)
        k_cache = randn(g, slots_n, kvh * d, dtype=dt)
        v_cache = randn(g, slots_n, kvh * d, dtype=dt)
        out[f"{name}.k_cache_in"] = to_np(k_cache)
        out[f"{name}.v_cache_in"] = to_np(v_cache)
        backend.store_kv_cache(kr, v, k_cache, v_cache, slots)
        out[f"{name}.fused"] = to_np(fused)
        out[f"{name}.positions"] = to_np(positions)
        out[f"{name}.slots"] = to_np(slots)
        out[f"{name}.cos"] = to_np(rope.cos_cache)
        out[f"{name}.sin"] = to_np(rope.sin_cache)
        out[f"{name}.q_out"] = to_np(qr)
        out[f"{name}.k_out"] = to_np(kr)
        out[f"{name}.k_cache_out"] = to_np(k_cache)
        out[f"{name}.v_cache_out"] = to_np(v_cache)
        out[f"{name}.meta"] = np.array([n, h, kvh, d, int(norm), max_pos, slots_n])
        out[f"{name}.base"] = np.array(base)
        out[f"{name}.eps"] = np.array(1e-6)
        out[f"{name}.dtype"] = np.array(dtype_name(dt))
    out["cases"] = np.array([c[0] for c in cases])
    save("qkrope", **out)


# ---------------------------------------------------------------------------
# caller side: all-reduce after the row-parallel GEMM, then RMSNorm with residual
# ---------------------------------------------------------------------------


def gen_arnorm():
    """``RowParallelLinear.forward``'s ``dist.all_reduce(y)`` (linear.py:314-319) followed by the
    reference's own ``RMSNorm.forward(x, residual)`` (layernorm.py:81-131).  The collective is a SUM in
    the model dtype: recorded here as the in-dtype elementwise accumulation over ranks (what gloo does on
    CPU; for world 2 every implementation gives T(a+b) exactly)."""
    from minivllm.models.layers.layernorm import RMSNorm

    out = {}
    cases = [
        # name, world, rows, hidden, dtype, residual?, norm?
        ("tp2_bf16", 2, 8, 2560, torch.bfloat16, True, True),
        ("tp2_bf16_first", 2, 5, 1024, torch.bfloat16, False, True),
        ("tp4_bf16", 4, 3, 3584, torch.bfloat16, True, True),
        ("tp8_f16", 8, 2, 1024, torch.float16, True, True),
        ("tp2_f32", 2, 4, 256, torch.float32, True, True),
        ("tp2_bf16_plain", 2, 6, 512, torch.bfloat16, False, False),
    ]
    for idx, (name, world, rows, hidden, dt, has_res, has_norm) in enumerate(cases):
        g = torch.Generator().manual_seed(900 + idx)
        xs = [randn(g, rows, hidden, dtype=dt) for _ in range(world)]
        total = xs[0].clone()
        for x in xs[1:]:
            total += x
        residual = randn(g, rows, hidden, dtype=dt) if has_res else None
        if has_norm:
            norm = RMSNorm(hidden, eps=1e-6)
            norm.weight.data = (1.0 + 0.1 * randn(g, hidden)).to(dt)
            y, new_res = norm(total, residual)
            out[f"{name}.weight"] = to_np(norm.weight.data)
        else:
            y, new_res = total, None
        for r, x in enumerate(xs):
            out[f"{name}.x{r}"] = to_np(x)
        if has_res:
            out[f"{name}.residual_in"] = to_np(residual)
            out[f"{name}.residual_out"] = to_np(new_res)
        out[f"{name}.sum"] = to_np(total)
        out[f"{name}.y"] = to_np(y)
        out[f"{name}.meta"] = np.array([world, rows, hidden, int(has_res), int(has_norm)])
        out[f"{name}.dtype"] = np.array(dtype_name(dt))
    out["cases"] = np.array([c[0] for c in cases])
    save("arnorm", **out)


if __name__ == "__main__":
    print("generating golden vectors from", REFERENCE)
    only = set(sys.argv[1:])
    for fn in (gen_store, gen_decode, gen_prefill, gen_layer, gen_metadata, gen_engine, gen_engine_named, gen_qkrope, gen_arnorm):
        if not only and fn is gen_engine_named and os.environ.get('MVB200_SKIP_SLOW_GOLDEN'):
            continue
        if not only or fn.__name__[4:] in only:
            fn()
