"""Device-side decode bookkeeping: ``decode_advance`` and ``greedy_argmax``.

CPU part pins the oracle (oracle/decode_meta_oracle.py) on the metadata recorded from the reference's
own executor (tests/golden/metadata.npz, decode steps d0..d3 with a block-boundary crossing) and on
``torch.argmax``.  GPU part (``-m gpu``): the CUDA kernels, through the C ABI, bit-exact vs the oracle.
"""

import numpy as np
import pytest
import torch

from oracle import decode_meta_oracle as O
from tests.conftest import load_golden

DECODE_TAGS = ["d0", "d1", "d2", "d3"]


def _golden_seqs(g, tag):
    return [O.HostSeq(g[f"{tag}.seq{i}.token_ids"], g[f"{tag}.seq{i}.block_table"]) for i in range(int(g[f"{tag}.n"]))]


# ------------------------------------------------------------------------------------------------ CPU
@pytest.mark.parametrize("tag", DECODE_TAGS + ["w_decode"])
def test_oracle_prepare_decode_matches_reference_executor(tag):
    g = load_golden("metadata")
    got = O.prepare_decode(_golden_seqs(g, tag))
    for f in ("input_ids", "positions", "slot_mapping", "context_lens", "block_tables"):
        want = g[f"{tag}.{f}"]
        assert got[f].dtype == want.dtype, f
        assert np.array_equal(got[f], want), f


def test_oracle_greedy_is_torch_argmax():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(6, 1000, generator=g)
    x[1, 17] = x[1, 900] = 9.0            # tie -> first
    x[2, 5] = float("nan")                # NaN is the maximum
    x[3] = 0.0
    x[3, 10] = -0.0                       # -0.0 == 0.0 -> index 0
    x[4, 999] = float("inf")
    x[5] = -float("inf")
    assert O.greedy(x.numpy()).tolist() == torch.argmax(x, dim=-1).tolist()


def test_golden_decode_steps_cross_a_block_boundary():
    g = load_golden("metadata")
    assert len(g["d1.seq0.block_table"]) == 4 and len(g["d2.seq0.block_table"]) == 5
    assert int(g["d2.context_lens"][0]) % 64 == 1


# ------------------------------------------------------------------------------------------------ GPU
def _device_state(bs, mb, dev):
    return {
        "next_tokens": torch.zeros(bs, dtype=torch.int64, device=dev),
        "appends": torch.full((bs,), -1, dtype=torch.int32, device=dev),
        "input_ids": torch.full((bs,), 77, dtype=torch.int64, device=dev),
        "positions": torch.full((bs,), 77, dtype=torch.int64, device=dev),
        "slot_mapping": torch.full((bs,), 77, dtype=torch.int32, device=dev),
        "context_lens": torch.zeros(bs, dtype=torch.int32, device=dev),
        "block_tables": torch.full((bs, mb), -1, dtype=torch.int32, device=dev),
    }


def _advance(m, st, num_blocks):
    m.ops.decode_advance(st["next_tokens"], st["appends"], st["input_ids"], st["positions"], st["slot_mapping"],
                         st["context_lens"], st["block_tables"], num_blocks)


def _check(st, want, b):
    for f in ("input_ids", "positions", "slot_mapping", "context_lens"):
        assert np.array_equal(st[f][:b].cpu().numpy(), want[f]), f
    w = want["block_tables"].shape[1]
    assert np.array_equal(st["block_tables"][:b, :w].cpu().numpy(), want["block_tables"])
    assert (st["block_tables"][:b, w:] == -1).all()
    # bucket padding rows (inference_executor.py:757-760)
    assert (st["slot_mapping"][b:] == -1).all() and (st["context_lens"][b:] == 0).all()
    assert (st["input_ids"][b:] == 0).all() and (st["positions"][b:] == 0).all()
    assert (st["appends"] == -1).all()


@pytest.fixture(scope="module")
def m():
    import minivllm_b200

    return minivllm_b200


@pytest.mark.gpu
def test_advance_reproduces_reference_decode_steps(m, cuda_device):
    """From the state before step d0, four advances give exactly the tensors the reference executor built
    on the host for d0..d3 (d2 appends a block)."""
    g = load_golden("metadata")
    bs, mb, b = 4, 8, 3
    st = _device_state(bs, mb, cuda_device)
    seqs = _golden_seqs(g, "d0")
    for i, s in enumerate(seqs):
        st["block_tables"][i, : len(s.block_table)] = torch.tensor(s.block_table, dtype=torch.int32)
        st["context_lens"][i] = len(s) - 1
    prev_tables = [list(s.block_table) for s in seqs]
    for tag in DECODE_TAGS:
        seqs = _golden_seqs(g, tag)
        st["next_tokens"][:b] = torch.tensor([s.token_ids[-1] for s in seqs], device=cuda_device)
        for i, s in enumerate(seqs):
            if len(s.block_table) > len(prev_tables[i]):
                st["appends"][i] = s.block_table[-1]
        prev_tables = [list(s.block_table) for s in seqs]
        _advance(m, st, 1024)
        want = {f: g[f"{tag}.{f}"] for f in ("input_ids", "positions", "slot_mapping", "context_lens", "block_tables")}
        _check(st, want, b)


@pytest.mark.gpu
@pytest.mark.parametrize("batch,bucket,steps", [(37, 48, 150), (256, 256, 70), (1, 1, 130)])
def test_advance_long_run_against_oracle(m, cuda_device, batch, bucket, steps):
    rng = np.random.default_rng(batch)
    bs_tok, mb, nb = 64, 16, 4096
    free = list(rng.permutation(nb))
    seqs = []
    for _ in range(batch):
        n = int(rng.integers(2, 600))
        blocks = [int(free.pop(0)) for _ in range((n + bs_tok - 1) // bs_tok)]
        seqs.append(O.HostSeq(rng.integers(10, 1000, n), blocks))
    st = _device_state(bucket, mb, cuda_device)
    table = np.full((bucket, mb), -1, np.int32)
    ctx = np.zeros(bucket, np.int32)
    tok = np.zeros(bucket, np.int64)
    for i, s in enumerate(seqs):
        table[i, : len(s.block_table)] = s.block_table
        ctx[i] = len(s) - 1
        tok[i] = s.token_ids[-1]
    st["block_tables"].copy_(torch.from_numpy(table))
    st["context_lens"].copy_(torch.from_numpy(ctx))
    st["next_tokens"].copy_(torch.from_numpy(tok))
    for step in range(steps):
        _advance(m, st, nb)
        _check(st, O.prepare_decode(seqs, width=max(len(s.block_table) for s in seqs)), batch)
        # host side of the next step: append the sampled token, maybe take a block
        new = rng.integers(10, 1000, batch)
        app = np.full(bucket, -1, np.int32)
        for i, s in enumerate(seqs):
            s.token_ids.append(int(new[i]))
            bid = O.may_append(s, free)
            if bid is not None:
                app[i] = bid
        tok[:batch] = new
        st["next_tokens"].copy_(torch.from_numpy(tok))
        if (app >= 0).any():
            st["appends"].copy_(torch.from_numpy(app))


@pytest.mark.gpu
def test_advance_out_of_range_entries_give_no_slot(m, cuda_device):
    st = _device_state(4, 2, cuda_device)
    st["context_lens"].copy_(torch.tensor([5, 127, 128, 70], dtype=torch.int32))
    st["block_tables"].copy_(torch.tensor([[3, -1], [3, 9999], [3, 4], [3, -1]], dtype=torch.int32))
    _advance(m, st, 100)
    # row 1: block id beyond the cache; row 2: column 2 beyond the table; row 3: missing entry
    assert st["slot_mapping"].tolist() == [3 * 64 + 5, -1, -1, -1]
    assert st["context_lens"].tolist() == [6, 128, 129, 71]
    with pytest.raises(ValueError):
        m.ops.decode_advance(st["next_tokens"][:3], st["appends"], st["input_ids"], st["positions"],
                             st["slot_mapping"], st["context_lens"], st["block_tables"], 100)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m.ops.decode_advance(*(t.cpu() for t in (st["next_tokens"], st["appends"], st["input_ids"], st["positions"],
                                                 st["slot_mapping"], st["context_lens"], st["block_tables"])), 100)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("rows,vocab", [(1, 151936), (8, 151936), (256, 151936), (512, 32000), (3, 1000), (5, 7), (2, 75968)])
def test_greedy_argmax_matches_oracle(m, cuda_device, dtype, rows, vocab):
    g = torch.Generator().manual_seed(rows * 131 + vocab)
    x = torch.randn(rows, vocab, generator=g).to(dtype)  # bf16 rounding produces plenty of exact ties
    if vocab > 100:
        x[0, 17] = x[0, vocab - 3] = 30.0
        if rows > 1:
            x[1, vocab // 2] = float("nan")
        if rows > 2:
            x[2] = 0.0
            x[2, 9] = -0.0
        if rows > 3:
            x[3] = -float("inf")          # everything -inf -> index 0
        if rows > 4:
            x[4] = -float("inf")
            x[4, vocab - 2] = -1e30
    want = O.greedy(x.float().numpy())
    val = torch.empty(rows, dtype=torch.float32, device=cuda_device)
    got = m.ops.greedy_argmax(x.to(cuda_device), out_value=val, index_offset=1000)
    assert got.dtype == torch.int64
    assert np.array_equal(got.cpu().numpy(), want + 1000)
    picked = x.float()[torch.arange(rows), torch.from_numpy(want)]
    assert torch.equal(torch.nan_to_num(val.cpu(), nan=-1.0), torch.nan_to_num(picked, nan=-1.0))


@pytest.mark.gpu
def test_greedy_argmax_strided_rows_and_capture(m, cuda_device):
    big = torch.randn(16, 5003, device=cuda_device, dtype=torch.bfloat16)
    view = big[:, 1:4998]  # unaligned row starts -> scalar path
    assert np.array_equal(m.ops.greedy_argmax(view).cpu().numpy(), O.greedy(view.float().cpu().numpy()))
    logits = torch.randn(8, 151936, device=cuda_device, dtype=torch.bfloat16)
    out = torch.zeros(8, dtype=torch.int64, device=cuda_device)
    ws = torch.empty(m.ops.greedy_argmax_workspace_numel(8, 151936), dtype=torch.int64, device=cuda_device)
    m.ops.greedy_argmax(logits, out, workspace=ws)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        m.ops.greedy_argmax(logits, out, workspace=ws)
    logits.copy_(torch.randn(8, 151936, device=cuda_device).to(torch.bfloat16))
    graph.replay()
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), O.greedy(logits.float().cpu().numpy()))
    with pytest.raises(ValueError):
        m.ops.greedy_argmax(logits[0])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m.ops.greedy_argmax(logits.cpu())


@pytest.mark.gpu
@pytest.mark.parametrize("graphs", [False, True])
def test_device_loop_generates_the_same_tokens_as_the_host_loop(cuda_device, graphs):
    """Engine-level: decode_advance + greedy_argmax inside the step (eager and captured) vs the host
    bookkeeping of the reference (prepare_decode every step), across block boundaries and with
    sequences finishing at different times."""
    from harness.engine import Engine
    from harness.model import PRESETS

    g = torch.Generator().manual_seed(11)
    prompts = [torch.randint(10, 500, (n,), generator=g).tolist() for n in (60, 127, 5, 190, 64)]
    kw = dict(device=cuda_device, dtype=torch.float32, max_num_seqs=8, max_model_len=512, num_blocks=64,
              enforce_eager=not graphs, init="seeded")
    host = Engine(PRESETS["tiny-qwen2"], device_loop=False, **kw)
    loop = Engine(PRESETS["tiny-qwen2"], device_loop=True, **kw)
    assert loop.device_loop and not host.device_loop
    want, _ = host.generate(prompts, 70)
    got, _ = loop.generate(prompts, 70)
    assert got == want
    # different lengths -> the live set shrinks and the loop state is reloaded
    seqs_want, _ = host.generate(prompts, [9, 3, 70, 1, 20])
    seqs_got, _ = loop.generate(prompts, [9, 3, 70, 1, 20])
    assert seqs_got == seqs_want and [len(x) for x in seqs_got] == [9, 3, 70, 1, 20]
