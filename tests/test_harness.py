"""The harness (model + block pool + executor) against the reference engine's token streams.

CPU part: harness with the oracle attention double == reference ``LLM.generate`` (fp32, eager),
recorded in tests/golden/engine.npz.  GPU part (``-m gpu``): the same harness with the CUDA layer.
"""

import numpy as np
import pytest
import torch

from tests.conftest import load_golden


def _prompts(g, preset):
    n = int(g[f"{preset}.n"])
    return [g[f"{preset}.prompt{i}"].tolist() for i in range(n)], [g[f"{preset}.tokens{i}"].tolist() for i in range(n)]


def _engine(preset, device, dtype, attention_cls=None, **kw):
    from harness.engine import Engine
    from harness.model import PRESETS

    args = dict(device=device, dtype=dtype, max_num_seqs=8, max_model_len=1024, num_blocks=64,
                enforce_eager=True, init="seeded", attention_cls=attention_cls)
    args.update(kw)
    return Engine(PRESETS[preset], **args)


@pytest.mark.parametrize("preset", ["tiny-qwen3", "tiny-qwen2"])
def test_harness_reproduces_reference_token_streams_on_cpu(preset):
    from tests.oracle_attention import OracleAttention

    g = load_golden("engine")
    prompts, want = _prompts(g, preset)
    eng = _engine(preset, "cpu", torch.float32, OracleAttention)
    outs, _ = eng.generate(prompts, 32)
    assert outs == want
    assert eng.stats["prefix_hits"] == 0


def test_block_pool_follows_reference_layout():
    """block_manager.py:252-339, 395-450: prefix hit shares blocks, new block at len % 64 == 1."""
    from harness.engine import BlockPool, Seq

    pool = BlockPool(16)
    base = list(range(100, 300))
    a = Seq(base, 4)
    pool.allocate(a)
    assert a.block_table == [0, 1, 2, 3] and a.num_cached_tokens == 0
    b = Seq(base[:128] + [7] * 10, 4)
    pool.allocate(b)
    assert b.block_table[:2] == [0, 1] and b.num_cached_tokens == 128 and b.block_table[2] == 4
    c = Seq(list(range(64)), 4)
    pool.allocate(c)
    n_before = len(c.block_table)
    c.token_ids.append(5)
    pool.may_append(c)            # len 65 -> new block
    assert len(c.block_table) == n_before + 1
    pool.deallocate(a)
    d = Seq(base[:64], 4)        # block 0 is still referenced by b -> shared again
    pool.allocate(d)
    assert d.block_table == [0] and d.num_cached_tokens == 64
    # same hash function and chaining as the reference (xxh64 over prefix || int32 tokens)
    try:
        import xxhash, struct
        h = xxhash.xxh64()
        h.update(struct.pack("<3i", 1, 2, 3))
        assert BlockPool.compute_hash([1, 2, 3]) == h.intdigest()
    except ImportError:
        pass


def test_prefix_hits_do_not_change_generation_on_cpu():
    """Generation with prefix-cache hits == generation from a cold cache (same engine math)."""
    from tests.oracle_attention import OracleAttention

    g = torch.Generator().manual_seed(3)
    shared = torch.randint(10, 500, (128,), generator=g).tolist()
    prompts = [shared + torch.randint(10, 500, (n,), generator=g).tolist() for n in (40, 90)]
    cold = _engine("tiny-qwen2", "cpu", torch.float32, OracleAttention)
    ref, _ = cold.generate(prompts[1:], 12)
    warm = _engine("tiny-qwen2", "cpu", torch.float32, OracleAttention)
    warm.generate(prompts[:1], 3)
    got, _ = warm.generate(prompts[1:], 12)
    assert warm.stats["prefix_hits"] >= 2
    assert got == ref


# ------------------------------------------------------------------------------------------------
def test_prefill_batch_split_does_not_fake_prefix_hits():
    """A sequence put back because the token budget is full must not leave hashes of never-computed blocks
    behind (it would 'hit' them when it is scheduled again and skip its own prefill)."""
    from tests.oracle_attention import OracleAttention

    g = torch.Generator().manual_seed(21)
    prompts = [torch.randint(10, 500, (n,), generator=g).tolist() for n in (200, 150, 130, 70)]
    whole = _engine("tiny-qwen3", "cpu", torch.float32, OracleAttention)
    want, _ = whole.generate(prompts, 6)
    split = _engine("tiny-qwen3", "cpu", torch.float32, OracleAttention, max_num_batched_tokens=256)
    got, _ = split.generate(prompts, 6)
    assert split.stats["prefix_hits"] == 0 and whole.stats["prefix_hits"] == 0
    assert got == want


# GPU: the same harness around the CUDA layer
# ------------------------------------------------------------------------------------------------


@pytest.mark.gpu
@pytest.mark.parametrize("preset", ["tiny-qwen3", "tiny-qwen2"])
def test_gpu_fp32_token_streams_match_reference(preset, cuda_device):
    """fp32 model + generic CUDA kernels vs the reference's CPU fp32 stream: first 32 greedy tokens."""
    g = load_golden("engine")
    prompts, want = _prompts(g, preset)
    eng = _engine(preset, cuda_device, torch.float32)
    outs, picked, logits = eng.generate(prompts, 32, forced=want, return_logits=True)
    _assert_streams(picked, want, logits, margin=1e-3, min_exact=0.97)


@pytest.mark.gpu
@pytest.mark.parametrize("graphs", [False, True])
def test_gpu_bf16_token_streams_match_reference(graphs, cuda_device):
    """bf16 model + tcgen05 prefill / TMA decode kernels (head_dim 128), eager and CUDA-Graph decode.
    Teacher-forced on the reference stream; a disagreement is accepted only when the reference's token
    is a near-tie of the top logit (bf16 model arithmetic vs the reference's fp32)."""
    import minivllm_b200 as m

    g = load_golden("engine")
    prompts, want = _prompts(g, "tiny-qwen3")
    eng = _engine("tiny-qwen3", cuda_device, torch.bfloat16, enforce_eager=not graphs)
    outs, picked, logits = eng.generate(prompts, 32, forced=want, return_logits=True)
    assert m.ops.last_kernel() in ("decode_tma_mma", "decode_tma_mma_fused_reduce", "store_vec16", "prefill_tcgen05", "prefill_tcgen05_2q", "prefill_tcgen05_dbs")
    _assert_streams(picked, want, logits, margin=0.05, min_exact=0.9)


@pytest.mark.gpu
def test_gpu_prefix_hits_and_graphs_do_not_change_generation(cuda_device):
    g = torch.Generator().manual_seed(3)
    shared = torch.randint(10, 500, (192,), generator=g).tolist()
    prompts = [shared + torch.randint(10, 500, (n,), generator=g).tolist() for n in (40, 90, 300)]
    cold = _engine("tiny-qwen3", cuda_device, torch.bfloat16)
    ref, _ = cold.generate(prompts[1:], 40)
    warm = _engine("tiny-qwen3", cuda_device, torch.bfloat16, enforce_eager=False)
    warm.generate(prompts[:1], 3)
    got, _ = warm.generate(prompts[1:], 40)
    assert warm.stats["prefix_hits"] >= 6
    same = sum(a == b for x, y in zip(got, ref) for a, b in zip(x, y))
    assert same >= 0.9 * sum(len(x) for x in ref), (got, ref)


def _assert_streams(picked, want, logits, margin, min_exact=1.0):
    total = exact = 0
    for seq_pick, seq_want, seq_logits in zip(picked, want, logits):
        for t, (a, b) in enumerate(zip(seq_pick, seq_want)):
            total += 1
            if a == b:
                exact += 1
                continue
            lg = seq_logits[t]
            gap = float(lg.max() - lg[b])
            spread = float(lg.max() - lg.median())
            assert gap <= margin * max(spread, 1e-6), f"token {t}: picked {a}, reference {b}, logit gap {gap} (spread {spread})"
    assert exact >= min_exact * total, f"only {exact}/{total} tokens identical"
