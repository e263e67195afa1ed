"""Host logic of the executor (T2-T5) against metadata recorded from the reference's own
Sequence + BlockManager + InferenceExecutor._prepare_* (tests/golden/metadata.npz).  CPU only."""

import types

import numpy as np
import pytest
import torch

from tests.conftest import load_golden


def _executor(**kw):
    import minivllm_b200  # noqa: F401
    from minivllm_b200.executor import HotPathExecutor

    args = dict(num_layers=2, num_kv_heads=2, head_dim=64, hidden_size=128, dtype=torch.float32,
                device="cpu", max_num_seqs=8, max_model_len=1024)
    args.update(kw)
    return HotPathExecutor(None, **args)


class _Seq:
    def __init__(self, token_ids, block_table, num_cached):
        self.token_ids = list(token_ids)
        self.block_table = list(block_table)
        self.num_cached_tokens = int(num_cached)

    def __len__(self):
        return len(self.token_ids)


def _seqs(g, tag):
    return [_Seq(g[f"{tag}.seq{i}.token_ids"], g[f"{tag}.seq{i}.block_table"], g[f"{tag}.seq{i}.num_cached"])
            for i in range(int(g[f"{tag}.n"]))]


@pytest.mark.parametrize("tag", [str(t) for t in load_golden("metadata")["steps"]])
def test_metadata_matches_reference_executor(tag):
    import minivllm_b200 as m

    g = load_golden("metadata")
    ex = _executor()
    seqs = _seqs(g, tag)
    prefill = bool(g[f"{tag}.is_prefill"])
    ids, pos = ex.prepare_prefill(seqs) if prefill else ex.prepare_decode(seqs)
    ctx = m.get_context()
    try:
        assert ids.dtype == torch.int64 and pos.dtype == torch.int64
        assert np.array_equal(ids.numpy(), g[f"{tag}.input_ids"])
        assert np.array_equal(pos.numpy(), g[f"{tag}.positions"])
        assert ctx.is_prefill == prefill
        assert ctx.max_seqlen_q == int(g[f"{tag}.max_seqlen_q"]) and ctx.max_seqlen_k == int(g[f"{tag}.max_seqlen_k"])
        for f in ("cum_seqlens_q", "cum_seqlens_k", "slot_mapping", "context_lens", "block_tables"):
            mine = getattr(ctx, f)
            key = f"{tag}.{f}"
            if key in g.files:
                assert mine is not None, f
                assert mine.dtype == torch.int32, (f, mine.dtype)
                assert np.array_equal(mine.numpy(), g[key]), f
            else:
                assert mine is None, f
    finally:
        m.reset_context()


def test_prefix_hit_shapes_in_golden():
    """Sanity on the recorded scenario itself: 50 % style prefix hits and a fully cached prompt."""
    g = load_golden("metadata")
    assert int(g["p1.seq0.num_cached"]) == 128 and int(g["p1.seq2.num_cached"]) == 192
    cu_q, cu_k = g["p1.cum_seqlens_q"], g["p1.cum_seqlens_k"]
    assert list(np.diff(cu_q)) == [70, 64, 0] and list(np.diff(cu_k)) == [198, 64, 192]
    # the shared prefix blocks are physically shared (block_manager.py:324-336)
    assert list(g["p1.seq0.block_table"][:2]) == list(g["p0.seq0.block_table"][:2])


def test_graph_ladder_and_cache_geometry():
    from minivllm_b200.executor import graph_batch_sizes

    assert graph_batch_sizes(512)[:6] == [1, 2, 4, 8, 12, 16] and graph_batch_sizes(512)[-1] == 512
    assert graph_batch_sizes(256)[-1] == 256 and graph_batch_sizes(8) == [1, 2, 4, 8]
    assert graph_batch_sizes(20) == [1, 2, 4, 8, 12, 16, 20]
    ex = _executor(num_layers=28, num_kv_heads=8, head_dim=128, dtype=torch.bfloat16, max_num_seqs=256, max_model_len=4096)
    assert ex.kv_block_bytes() == 2 * 28 * 64 * 8 * 128 * 2
    assert ex.max_blocks_per_seq == 64
    assert ex.num_blocks_for_budget(10**15) == 256 * 64          # capped at max_num_seqs * blocks/seq
    assert ex.num_blocks_for_budget(ex.kv_block_bytes() * 10 + 5) == 10
    with pytest.raises(ValueError, match="multiple of 64"):
        _executor(block_size=48)


def test_cache_allocation_and_layer_assignment():
    import minivllm_b200 as m

    class Block(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.attn = m.Attention(4, 64, 0.0, 2)

    model = torch.nn.Sequential(Block(), Block())
    from minivllm_b200.executor import HotPathExecutor

    ex = HotPathExecutor(model, num_layers=2, num_kv_heads=2, head_dim=64, hidden_size=256, dtype=torch.float32,
                         device="cpu", max_num_seqs=4, max_model_len=256)
    kv = ex.allocate_kv_cache(6)
    assert kv.shape == (2, 2, 6, 64, 2, 64) and not kv.any()
    assert model[0].attn.k_cache.data_ptr() == kv[0, 0].data_ptr()
    assert model[1].attn.v_cache.data_ptr() == kv[1, 1].data_ptr()
    assert model[1].attn.k_cache.shape == (6, 64, 2, 64)
    with pytest.raises(ValueError):
        ex.allocate_kv_cache(0)


def test_decode_batch_larger_than_staging_raises():
    ex = _executor(max_num_seqs=2)
    seqs = [_Seq([1, 2, 3], [0], 0)] * 3
    with pytest.raises(ValueError, match="exceeds max_num_seqs"):
        ex.prepare_decode(seqs)
