"""Parity of the CUDA path (through the C ABI) against the golden vectors and the CPU oracle.

Run on the B200 box: ``python -m pytest tests -m gpu``.
Tolerances: store is bit-exact; attention max-abs <= 2e-2 for bf16 (BASELINE.json
north_star), 4e-3 for fp16, 2e-5 for fp32 (the reference's own tests use
rtol 1e-4 / atol 1e-5 on fp32, tests/test_page_attention.py:120).
"""

import numpy as np
import pytest
import torch

from oracle import paged_attention_oracle as orc
from tests.conftest import golden_to_float, golden_to_torch, load_golden

pytestmark = pytest.mark.gpu

TOL = {"f32": 2e-5, "f16": 4e-3, "bf16": 2e-2}
TDT = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}


def _cases(name):
    return [str(c) for c in load_golden(name)["cases"]]


def _f32(t: torch.Tensor) -> np.ndarray:
    return t.detach().float().cpu().numpy()


def _bits(t: torch.Tensor) -> np.ndarray:
    return t.detach().contiguous().cpu().view(torch.uint8).numpy()


@pytest.fixture(scope="module")
def m(cuda_device):
    import minivllm_b200

    return minivllm_b200


# ---------------------------------------------------------------------------
# store_kvcache: bit-exact
# ---------------------------------------------------------------------------


@pytest.mark.parametrize("case", _cases("store"))
def test_store_golden_bit_exact(m, cuda_device, case):
    g = load_golden("store")
    dt = str(g[f"{case}.dtype"])
    n, kvh, d, slots_n, strided = (int(x) for x in g[f"{case}.meta"])
    key = golden_to_torch(g[f"{case}.key"], dt, cuda_device)
    value = golden_to_torch(g[f"{case}.value"], dt, cuda_device)
    if strided:  # rebuild row-strided views of a fused buffer, as the model produces them
        fused = torch.zeros(n, 3 * kvh * d + 8, dtype=key.dtype, device=cuda_device)
        fused[:, kvh * d : 2 * kvh * d] = key.view(n, -1)
        fused[:, 2 * kvh * d : 3 * kvh * d] = value.view(n, -1)
        key = fused[:, kvh * d : 2 * kvh * d].view(n, kvh, d)
        value = fused[:, 2 * kvh * d : 3 * kvh * d].view(n, kvh, d)
        assert not key.is_contiguous() or n == 1
    kc = golden_to_torch(g[f"{case}.k_cache_in"], dt, cuda_device)
    vc = golden_to_torch(g[f"{case}.v_cache_in"], dt, cuda_device)
    slots = torch.from_numpy(g[f"{case}.slots"]).to(cuda_device)
    m.store_kvcache(key, value, kc, vc, slots)
    torch.cuda.synchronize()
    assert np.array_equal(_bits(kc).ravel(), g[f"{case}.k_cache_out"].view(np.uint8).ravel())
    assert np.array_equal(_bits(vc).ravel(), g[f"{case}.v_cache_out"].view(np.uint8).ravel())
    assert m.ops.last_kernel().startswith("store_vec")


def test_store_mismatch_raises(m, cuda_device):
    k = torch.zeros(3, 2, 8, device=cuda_device)
    c = torch.zeros(8, 16, device=cuda_device)
    with pytest.raises(ValueError, match="slot_mapping size"):
        m.store_kvcache(k, k, c, c.clone(), torch.tensor([1, 2], device=cuda_device, dtype=torch.int32))


@pytest.mark.parametrize("n,kvh", [(256, 8), (8192, 8), (16384, 1)])
def test_store_full_size_roundtrip(m, cuda_device, n, kvh):
    """BASELINE sizes (cfg2 decode N=256, cfg3 prefill N=8192, cfg5 N=16384/rank): scatter then
    gather must return the inputs bit-for-bit and leave every other slot untouched."""
    d, bs = 128, 64
    g = torch.Generator(device="cpu").manual_seed(n + kvh)
    nblocks = n // bs + 40
    key = torch.randn(n, kvh, d, generator=g).to(torch.bfloat16).to(cuda_device)
    fused = torch.randn(n, 2 * kvh * d, generator=g).to(torch.bfloat16).to(cuda_device)
    value = fused[:, kvh * d :].view(n, kvh, d)  # row-strided like the model's v
    perm = torch.randperm(nblocks * bs, generator=g)[:n].to(torch.int32)
    perm[::97] = -1
    slots = perm.to(cuda_device)
    kc = torch.full((nblocks, bs, kvh, d), 7.0, dtype=torch.bfloat16, device=cuda_device)
    vc = torch.full((nblocks, bs, kvh, d), -3.0, dtype=torch.bfloat16, device=cuda_device)
    m.store_kvcache(key, value, kc, vc, slots)
    valid = slots >= 0
    idx = slots[valid].long()
    assert torch.equal(kc.view(-1, kvh, d)[idx], key[valid])
    assert torch.equal(vc.view(-1, kvh, d)[idx], value[valid])
    untouched = torch.ones(nblocks * bs, dtype=torch.bool, device=cuda_device)
    untouched[idx] = False
    assert (kc.view(-1, kvh * d)[untouched] == 7.0).all()
    assert (vc.view(-1, kvh * d)[untouched] == -3.0).all()


# ---------------------------------------------------------------------------
# paged decode
# ---------------------------------------------------------------------------


def _decode_inputs(g, case, device):
    dt = str(g[f"{case}.dtype"])
    return (
        dt,
        golden_to_torch(g[f"{case}.q"], dt, device),
        golden_to_torch(g[f"{case}.k_cache"], dt, device),
        golden_to_torch(g[f"{case}.v_cache"], dt, device),
        torch.from_numpy(g[f"{case}.block_tables"]).to(device),
        torch.from_numpy(g[f"{case}.context_lens"]).to(device),
        float(g[f"{case}.scale"]),
    )


@pytest.mark.parametrize("case", _cases("decode"))
def test_decode_golden(m, cuda_device, case):
    g = load_golden("decode")
    dt, q, kc, vc, bt, ctx, scale = _decode_inputs(g, case, cuda_device)
    out = m.paged_decode_attention(q, kc, vc, bt, ctx, scale)
    torch.cuda.synchronize()
    kern = m.ops.last_kernel()
    fast = dt != "f32" and kc.size(1) % 64 == 0 and kc.size(3) in (64, 128)
    assert kern.startswith("decode_tma_mma") if fast else kern == "decode_generic", kern
    assert out.dtype == q.dtype and out.shape == q.shape
    ref = golden_to_float(g[f"{case}.out"], dt)
    err = np.abs(_f32(out) - ref)
    assert err.max() <= TOL[dt], (case, err.max(), err.mean())
    # and against the float64 oracle on the same inputs
    o = orc.paged_decode(_f32(q), _f32(kc), _f32(vc), g[f"{case}.block_tables"], g[f"{case}.context_lens"], scale)
    assert np.abs(_f32(out) - o).max() <= TOL[dt]


def test_torch_ops_namespace_runs_the_same_kernels(m, cuda_device):
    """torch.ops.mvb200.* (custom torch ops) == the direct wrappers, bit for bit."""
    g = load_golden("decode")
    case = _cases("decode")[0]
    dt, q, kc, vc, bt, ctx, scale = _decode_inputs(g, case, cuda_device)
    want = m.paged_decode_attention(q, kc, vc, bt, ctx, scale)
    got = torch.ops.mvb200.paged_decode(q, kc, vc, bt, ctx, scale)
    assert torch.equal(got, want)
    n, kvh, d = 5, kc.size(2), kc.size(3)
    key = torch.randn(n, kvh, d, device=cuda_device).to(kc.dtype)
    val = torch.randn(n, kvh, d, device=cuda_device).to(kc.dtype)
    slots = torch.tensor([3, -1, 0, 7, 1], dtype=torch.int32, device=cuda_device)
    k1, v1, k2, v2 = kc.clone(), vc.clone(), kc.clone(), vc.clone()
    m.store_kvcache(key, val, k1, v1, slots)
    torch.ops.mvb200.store_kvcache(key, val, k2, v2, slots)
    assert torch.equal(k1, k2) and torch.equal(v1, v2) and not torch.equal(k1, kc)
    logits = torch.randn(4, 1000, device=cuda_device)
    assert torch.equal(torch.ops.mvb200.greedy_argmax(logits), logits.argmax(dim=-1))


@pytest.mark.parametrize("case", ["qwen3_g2", "qwen3_g4", "qwen2_g7", "qwen2_g8", "opt_mha"])
@pytest.mark.parametrize("splits", [1, 2, 3, 9])
def test_decode_split_kv_invariance(m, cuda_device, case, splits):
    g = load_golden("decode")
    dt, q, kc, vc, bt, ctx, scale = _decode_inputs(g, case, cuda_device)
    out = m.paged_decode_attention(q, kc, vc, bt, ctx, scale, num_splits=splits)
    assert m.ops.last_kernel().startswith("decode_tma_mma")
    ref = golden_to_float(g[f"{case}.out"], dt)
    assert np.abs(_f32(out) - ref).max() <= TOL[dt]


def test_decode_padding_rows_and_stale_columns(m, cuda_device):
    """CUDA-Graph replay shape: full-width table with stale ids beyond the live columns,
    padded rows with context_len == 0 (inference_executor.py:751-768) -> zeros, no fault."""
    g = load_golden("decode")
    dt, q, kc, vc, bt, ctx, scale = _decode_inputs(g, "qwen3_g2", cuda_device)
    b, mb = bt.shape
    wide = torch.full((b + 3, mb + 20), 10**6, dtype=torch.int32, device=cuda_device)  # garbage ids
    wide[:b, :mb] = bt
    ctx2 = torch.zeros(b + 3, dtype=torch.int32, device=cuda_device)
    ctx2[:b] = ctx
    q2 = torch.randn(b + 3, q.size(1), q.size(2), device=cuda_device).to(q.dtype)
    q2[:b] = q
    for splits in (1, 4):
        out = m.paged_decode_attention(q2, kc, vc, wide, ctx2, scale, num_splits=splits)
        ref = golden_to_float(g["qwen3_g2.out"], dt)
        assert np.abs(_f32(out[:b]) - ref).max() <= TOL[dt]
        assert not out[b:].float().abs().any()


def test_decode_nan_garbage_beyond_context_is_ignored(m, cuda_device):
    """torch.empty caches can hold NaN bit patterns in never-written slots."""
    g = load_golden("decode")
    dt, q, kc, vc, bt, ctx, scale = _decode_inputs(g, "qwen3_g4", cuda_device)
    kc, vc = kc.clone(), vc.clone()
    bs = kc.size(1)
    for b in range(bt.size(0)):
        c = int(ctx[b])
        last = int(bt[b, (c - 1) // bs])
        if c % bs:
            kc[last, c % bs :] = float("nan")
            vc[last, c % bs :] = float("nan")
    out = m.paged_decode_attention(q, kc, vc, bt, ctx, scale)
    ref = golden_to_float(g["qwen3_g4.out"], dt)
    assert np.abs(_f32(out) - ref).max() <= TOL[dt]


def test_decode_error_contracts(m, cuda_device):
    q = torch.randn(2, 8, 64, device=cuda_device)
    empty = torch.tensor([], device=cuda_device)
    bt = torch.zeros(2, 2, dtype=torch.int32, device=cuda_device)
    ctx = torch.tensor([8, 8], dtype=torch.int32, device=cuda_device)
    with pytest.raises(ValueError, match="empty"):
        m.PagedDecodeAttention()(q, empty, empty, bt, ctx, 0.125)
    kc = torch.randn(4, 16, 8, 64, device=cuda_device)
    z = m.PagedDecodeAttention()(q, kc, kc, bt[:, :0], ctx, 0.125)
    assert z.shape == (2, 8, 64) and not z.any()
    # int64 index tensors are accepted like the reference's tests pass them
    out = m.paged_decode_attention(q, kc, kc, bt.long(), ctx.long(), 0.125)
    assert out.shape == (2, 8, 64)


def _random_paged(device, batch, ctxs, kvh, d, bs, dtype, seed, extra_blocks=8):
    g = torch.Generator(device="cpu").manual_seed(seed)
    nblk = [(c + bs - 1) // bs for c in ctxs]
    total = sum(nblk) + extra_blocks
    perm = torch.randperm(total, generator=g)
    mb = max(nblk)
    bt = torch.full((batch, mb), -1, dtype=torch.int32)
    c0 = 0
    for b, nb in enumerate(nblk):
        bt[b, :nb] = perm[c0 : c0 + nb]
        c0 += nb
    kc = torch.empty(total, bs, kvh, d, dtype=dtype, device=device).normal_(generator=None)
    vc = torch.empty(total, bs, kvh, d, dtype=dtype, device=device).normal_(generator=None)
    return kc, vc, bt.to(device), torch.tensor(ctxs, dtype=torch.int32, device=device)


@pytest.mark.parametrize(
    "name,batch,ctx,h,kvh",
    [
        ("cfg2_qwen3_0.6b", 256, 2048, 16, 8),
        ("cfg4_qwen2_7b_32k", 2, 32768, 28, 4),
        ("cfg5_qwen2_72b_rank", 512, 1024, 8, 1),
        ("cfg2_tp8_rank", 256, 2048, 2, 1),
        ("cfg4_tp4_rank", 1, 32768, 7, 1),
    ],
)
def test_decode_full_size_against_oracle_samples(m, cuda_device, name, batch, ctx, h, kvh):
    """BASELINE.json sizes: oracle on sampled sequences + split-count invariance on all rows."""
    torch.manual_seed(batch + ctx)
    d, bs = 128, 64
    rng = np.random.default_rng(3)
    ctxs = [int(x) for x in rng.integers(max(1, ctx // 2), ctx + 1, size=batch)]
    ctxs[0] = ctx
    kc, vc, bt, cl = _random_paged(cuda_device, batch, ctxs, kvh, d, bs, torch.bfloat16, 11)
    q = torch.randn(batch, h, d, device=cuda_device).to(torch.bfloat16)
    scale = d**-0.5
    out = m.paged_decode_attention(q, kc, vc, bt, cl, scale)
    assert m.ops.last_kernel().startswith("decode_tma_mma")
    out_s = m.paged_decode_attention(q, kc, vc, bt, cl, scale, num_splits=5)
    assert (out.float() - out_s.float()).abs().max().item() <= 1e-2
    for b in (0, batch // 2, batch - 1):
        blocks = bt[b, : (ctxs[b] + bs - 1) // bs].long()
        kk = _f32(kc[blocks])
        vv = _f32(vc[blocks])
        local_bt = np.arange(blocks.numel(), dtype=np.int32)[None, :]
        o = orc.paged_decode(_f32(q[b : b + 1]), kk, vv, local_bt, np.array([ctxs[b]]), scale)
        err = np.abs(_f32(out[b : b + 1]) - o)
        assert err.max() <= TOL["bf16"], (name, b, err.max())


def test_decode_cuda_graph_capture_and_replay(m, cuda_device):
    """Capture once with static buffers, replay with different lengths/tables (T5)."""
    d, bs, h, kvh, batch, mb = 128, 64, 16, 8, 8, 16
    ctxs = [100, 513, 64, 1, 700, 1000, 65, 300]
    kc, vc, bt, cl = _random_paged(cuda_device, batch, ctxs, kvh, d, bs, torch.bfloat16, 5)
    q = torch.randn(batch, h, d, device=cuda_device).to(torch.bfloat16)
    st_q = torch.zeros_like(q)
    st_bt = torch.zeros(batch, mb, dtype=torch.int32, device=cuda_device)
    st_cl = torch.zeros(batch, dtype=torch.int32, device=cuda_device)
    st_out = torch.zeros_like(q)
    splits = 3
    ws = m.ops.decode_workspace(m.ops.decode_workspace_numel(batch, h, d, splits), cuda_device)
    scale = d**-0.5
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        m.paged_decode_attention(st_q, kc, vc, st_bt, st_cl, scale, num_splits=splits, workspace=ws, out=st_out)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        m.paged_decode_attention(st_q, kc, vc, st_bt, st_cl, scale, num_splits=splits, workspace=ws, out=st_out)
    for shift in (0, 37):
        ctx_now = [max(1, c - shift) for c in ctxs]
        st_q.copy_(q)
        st_bt.zero_()
        st_bt[:, : bt.size(1)] = bt.clamp(min=0)
        st_cl.copy_(torch.tensor(ctx_now, dtype=torch.int32))
        graph.replay()
        torch.cuda.synchronize()
        eager = m.paged_decode_attention(q, kc, vc, bt, torch.tensor(ctx_now, dtype=torch.int32, device=cuda_device), scale, num_splits=1)
        assert (st_out.float() - eager.float()).abs().max().item() <= 1e-2
        o = orc.paged_decode(_f32(q[:2]), _f32(kc), _f32(vc), bt[:2].cpu().numpy(), np.array(ctx_now[:2]), scale)
        assert np.abs(_f32(st_out[:2]) - o).max() <= TOL["bf16"]


def test_decode_persistent_kernel_reuses_one_workspace(m, cuda_device):
    """The persistent kernel draws (sequence, split, kv head) units from a counter in the workspace and fuses
    the split reduction through per-(sequence, kv head) tickets; both are re-armed by the kernel itself.  One
    workspace therefore serves back-to-back launches of DIFFERENT shapes and split counts — more units than
    resident CTAs, empty splits, padded rows, ragged lengths — and the results equal the oracle / the
    split-free launch every time."""
    d, bs = 128, 64
    rng = np.random.default_rng(11)
    shapes = [
        # batch, kv heads, group, context lengths, split counts tried
        (300, 2, 4, rng.integers(1, 600, 300).tolist(), (1, 3)),
        (3, 1, 8, [4000, 65, 1], (7, 2)),
        (40, 8, 2, [0] * 5 + rng.integers(1, 300, 35).tolist(), (4, 1)),
        (1, 4, 7, [9000], (37, 128)),
    ]
    need = max(m.ops.decode_workspace_numel(b, kvh * g, d, max(sp)) for b, kvh, g, _, sp in shapes)
    ws = m.ops.decode_workspace(need, cuda_device)
    for rep in range(2):
        for b, kvh, g, ctxs, split_counts in shapes:
            h = kvh * g
            kc, vc, bt, cl = _random_paged(cuda_device, b, [max(c, 1) for c in ctxs], kvh, d, bs, torch.bfloat16, 50 + b)
            cl = torch.tensor(ctxs, dtype=torch.int32, device=cuda_device)
            q = torch.randn(b, h, d, device=cuda_device).to(torch.bfloat16)
            base = m.paged_decode_attention(q, kc, vc, bt, cl, d**-0.5, num_splits=1, workspace=ws)
            rows = sorted({0, b // 2, b - 1})
            o = orc.paged_decode(_f32(q[rows]), _f32(kc), _f32(vc), bt[rows].cpu().numpy(), np.array(ctxs)[rows], d**-0.5)
            assert np.abs(_f32(base[rows]) - o).max() <= TOL["bf16"], (b, kvh, g)
            for splits in split_counts:
                out = m.paged_decode_attention(q, kc, vc, bt, cl, d**-0.5, num_splits=splits, workspace=ws)
                assert torch.isfinite(out.float()).all()
                assert (out.float() - base.float()).abs().max().item() <= 1e-2, (b, kvh, g, splits, rep)
    # the persistent state is back to zero after every launch (tickets + both counters)
    tail = ws.view(torch.int32)[-(65536 + 16):]
    assert not tail[:65536].any() and not tail[65536:65538].any()
    # no workspace under capture -> the statically scheduled, split-free variant
    b, kvh, g, ctxs, _ = shapes[0]
    kc, vc, bt, cl = _random_paged(cuda_device, b, ctxs, kvh, d, bs, torch.bfloat16, 50 + b)
    q = torch.randn(b, kvh * g, d, device=cuda_device).to(torch.bfloat16)
    want = m.paged_decode_attention(q, kc, vc, bt, cl, d**-0.5, num_splits=1, workspace=ws)
    out = torch.empty_like(q)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        m.paged_decode_attention(q, kc, vc, bt, cl, d**-0.5, out=out)
    graph.replay()
    torch.cuda.synchronize()
    assert (out.float() - want.float()).abs().max().item() <= 1e-2


@pytest.mark.parametrize("batch,kvh,g,splits", [(1, 4, 7, 37), (8, 8, 4, 4), (64, 2, 8, 1), (3, 1, 2, 2)])
def test_decode_run_ahead_under_dependent_launch(m, cuda_device, batch, kvh, g, splits):
    """``kv_prefix_stable=True`` (MVB200_DECODE_KV_PREFIX_STABLE): the producer warp fetches the older K/V tiles of
    its first unit before the dependency wait, i.e. while the store of the step's new tokens (the launch before
    it) is still running.  Several captured [store -> decode] steps back to back, the new token always in each
    sequence's last tile (block boundaries crossed on the way): every step must equal the same chain run without
    the flag, bit for bit, and the oracle."""
    d, bs, steps = 128, 64, 6
    h = kvh * g
    rng = np.random.default_rng(batch * 7 + splits)
    ctx0 = rng.integers(60, 700, batch).tolist()
    ctx0[0] = 63  # crosses into a new block at step 2
    kc, vc, bt, _ = _random_paged(cuda_device, batch, [c + steps + 1 for c in ctx0], kvh, d, bs, torch.bfloat16, 90 + batch)
    ws = m.ops.decode_workspace(m.ops.decode_workspace_numel(batch, h, d, splits), cuda_device)
    g_cpu = torch.Generator().manual_seed(batch)
    qs = torch.randn(steps, batch, h, d, generator=g_cpu).to(torch.bfloat16).to(cuda_device)
    ks = torch.randn(steps, batch, kvh, d, generator=g_cpu).to(torch.bfloat16).to(cuda_device)
    vs = torch.randn(steps, batch, kvh, d, generator=g_cpu).to(torch.bfloat16).to(cuda_device)
    lens = [torch.tensor([c + s + 1 for c in ctx0], dtype=torch.int32, device=cuda_device) for s in range(steps)]
    rows = torch.arange(batch, device=cuda_device)
    slots = [(bt[rows, (ln.long() - 1) // bs].long() * bs + (ln.long() - 1) % bs).to(torch.int32) for ln in lens]

    def chain(flag, kcache, vcache):
        outs = torch.empty(steps, batch, h, d, dtype=torch.bfloat16, device=cuda_device)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        graph = torch.cuda.CUDAGraph()

        def body():
            for s in range(steps):
                m.store_kvcache(ks[s], vs[s], kcache, vcache, slots[s])
                m.paged_decode_attention(qs[s], kcache, vcache, bt, lens[s], d**-0.5, num_splits=splits, workspace=ws,
                                         out=outs[s], kv_prefix_stable=flag)

        k0, v0 = kcache.clone(), vcache.clone()
        with torch.cuda.stream(side):
            body()  # warm-up (eager launches are chained with dependent launch as well)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        eager = outs.clone()
        kcache.copy_(k0)
        vcache.copy_(v0)
        with torch.cuda.graph(graph):
            body()
        for _ in range(3):
            kcache.copy_(k0)
            vcache.copy_(v0)
            graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(outs, eager)
        return outs

    plain = chain(False, kc.clone(), vc.clone())
    kc2, vc2 = kc.clone(), vc.clone()
    ahead = chain(True, kc2, vc2)
    assert torch.equal(ahead, plain)
    pick = sorted({0, batch // 2, batch - 1})
    o = orc.paged_decode(_f32(qs[-1][pick]), _f32(kc2), _f32(vc2), bt[pick].cpu().numpy(),
                         lens[-1][pick].cpu().numpy(), d**-0.5)
    assert np.abs(_f32(ahead[-1][pick]) - o).max() <= TOL["bf16"]


@pytest.mark.parametrize("dtype,batch,kvh,g,d,splits", [("bf16", 5, 2, 4, 128, 1), ("bf16", 1, 4, 7, 128, 9), ("f16", 33, 1, 8, 64, 2),
                                                        ("bf16", 256, 8, 2, 128, 1), ("f32", 4, 2, 2, 64, 1)])
def test_decode_with_fused_store(m, cuda_device, dtype, batch, kvh, g, d, splits):
    """``paged_decode_attention(new_key=..., new_value=..., slot_mapping=...)``: the decode kernel stores the step's
    new tokens itself.  Cache contents must be bit-identical to ``store_kvcache`` followed by the plain decode, and so
    must the attention output (same arithmetic); padded rows (slot -1, context 0) and row-strided k/v (views of a
    fused qkv buffer) included.  fp32 takes the generic kernel: the library stores with a separate launch."""
    bs, h = 64, kvh * g
    rng = np.random.default_rng(batch + splits)
    ctxs = rng.integers(1, 400, batch).tolist()
    if batch > 2:
        ctxs[1] = 64   # new token is the last row of a block
        ctxs[2] = 65   # ... the first row of a fresh block
    kc, vc, bt, cl = _random_paged(cuda_device, batch, ctxs, kvh, d, bs, TDT[dtype], 70 + batch)
    slots = (bt[torch.arange(batch), (cl.long() - 1) // bs].long() * bs + (cl.long() - 1) % bs).to(torch.int32)
    if batch > 4:  # CUDA-graph padding rows
        cl[-2:] = 0
        slots[-2:] = -1
    fused = torch.randn(batch, (h + 2 * kvh) * d, device=cuda_device).to(TDT[dtype])
    q = fused[:, : h * d].view(batch, h, d)
    k = fused[:, h * d : (h + kvh) * d].view(batch, kvh, d)
    v = fused[:, (h + kvh) * d :].view(batch, kvh, d)
    ws = m.ops.decode_workspace(m.ops.decode_workspace_numel(batch, h, d, splits), cuda_device)
    k1, v1 = kc.clone(), vc.clone()
    m.store_kvcache(k, v, k1, v1, slots)
    want = m.paged_decode_attention(q, k1, v1, bt, cl, d**-0.5, num_splits=splits, workspace=ws)
    k2, v2 = kc.clone(), vc.clone()
    got = m.paged_decode_attention(q, k2, v2, bt, cl, d**-0.5, num_splits=splits, workspace=ws,
                                   new_key=k, new_value=v, slot_mapping=slots, kv_prefix_stable=True)
    torch.cuda.synchronize()
    if dtype != "f32":
        assert "store_fused_in_decode" in m.ops.kernels_seen()
    assert np.array_equal(_bits(k2), _bits(k1)) and np.array_equal(_bits(v2), _bits(v1))
    assert not np.array_equal(_bits(k2), _bits(kc))
    assert torch.equal(got, want)
    with pytest.raises(ValueError, match="slot_mapping size"):
        m.paged_decode_attention(q, k2, v2, bt, cl, d**-0.5, new_key=k, new_value=v, slot_mapping=slots[:-1])


def test_decode_error_leaves_the_cache_untouched(m, cuda_device):
    """A ``mvb200_paged_decode_ex`` call that is refused must not have stored the step's tokens first:
    head_dim 320 has no kernel (EUNSUPPORTED) and its store is not fusable, so the separate store launch used to
    run before the refusal.  Called through the C ABI; ``ops`` would stop the shape earlier."""
    import ctypes

    native = m._native
    b, kvh, h, d, bs, nb = 2, 2, 4, 320, 64, 4
    kc = torch.randn(nb, bs, kvh, d, device=cuda_device)
    vc = torch.randn(nb, bs, kvh, d, device=cuda_device)
    k0, v0 = kc.clone(), vc.clone()
    q = torch.randn(b, h, d, device=cuda_device)
    out = torch.zeros_like(q)
    k = torch.randn(b, kvh, d, device=cuda_device)
    v = torch.randn(b, kvh, d, device=cuda_device)
    bt = torch.tensor([[0, 1], [2, 3]], dtype=torch.int32, device=cuda_device)
    cl = torch.tensor([5, 70], dtype=torch.int32, device=cuda_device)
    slots = torch.tensor([4, 3 * bs + 5], dtype=torch.int32, device=cuda_device)
    store = native.DecodeStore(k.data_ptr(), v.data_ptr(), kvh * d, kvh * d, slots.data_ptr())
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def call(head_dim, dtype_code):
        return native.lib().mvb200_paged_decode_ex(
            q.data_ptr(), h * d, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), 2, cl.data_ptr(), out.data_ptr(), h * d,
            None, 0, b, h, kvh, head_dim, bs, nb, 2, 1, d**-0.5, dtype_code, 0, ctypes.byref(store), stream)

    assert call(d, 0) == 2 and "head_dim" in native.last_error()
    assert call(d, 7) == 1 and "dtype" in native.last_error()
    torch.cuda.synchronize()
    assert torch.equal(kc, k0) and torch.equal(vc, v0)


# ---------------------------------------------------------------------------
# prefill
# ---------------------------------------------------------------------------


@pytest.mark.parametrize("case", _cases("prefill"))
def test_prefill_dense_golden(m, cuda_device, case):
    g = load_golden("prefill")
    dt = str(g[f"{case}.dtype"])
    q = golden_to_torch(g[f"{case}.q"], dt, cuda_device)
    k = golden_to_torch(g[f"{case}.k"], dt, cuda_device)
    v = golden_to_torch(g[f"{case}.v"], dt, cuda_device)
    cu = torch.from_numpy(g[f"{case}.cu_seqlens"]).to(cuda_device)
    lens = np.diff(g[f"{case}.cu_seqlens"])
    scale = float(g[f"{case}.scale"])
    out = m.prefill_varlen_attention(q, k, v, cu, cu, int(lens.max()), int(lens.max()), scale)
    ref = golden_to_float(g[f"{case}.out"], dt)
    o = orc.prefill_dense(_f32(q), _f32(k), _f32(v), g[f"{case}.cu_seqlens"], g[f"{case}.cu_seqlens"], scale)
    err = np.abs(_f32(out) - o)
    assert err.max() <= TOL[dt], (case, m.ops.last_kernel(), err.max(), err.mean())
    assert np.abs(_f32(out) - ref).max() <= 2 * TOL[dt]  # the reference rounds its scores to bf16


@pytest.mark.parametrize("dtype,h,kvh,d", [("bf16", 8, 2, 128), ("bf16", 7, 1, 128), ("f16", 4, 4, 64), ("f32", 4, 2, 64)])
def test_prefill_prefix_hits_equal_suffix_of_full_prefill(m, cuda_device, dtype, h, kvh, d):
    """P1: queries of the uncached suffix, keys read through block_tables (50 % hits)."""
    bs = 64
    tdt = TDT[dtype]
    lens = [300, 128, 65, 1024]
    cached = [128, 64, 64, 512]
    g = torch.Generator().manual_seed(9)
    n = sum(lens)
    q = torch.randn(n, h, d, generator=g).to(tdt).to(cuda_device)
    k = torch.randn(n, kvh, d, generator=g).to(tdt).to(cuda_device)
    v = torch.randn(n, kvh, d, generator=g).to(tdt).to(cuda_device)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    nblk = [(x + bs - 1) // bs for x in lens]
    perm = torch.randperm(sum(nblk) + 5, generator=g)
    kc = torch.full((perm.numel(), bs, kvh, d), float("nan"), dtype=tdt, device=cuda_device)
    vc = torch.full((perm.numel(), bs, kvh, d), float("nan"), dtype=tdt, device=cuda_device)
    bt = torch.full((len(lens), max(nblk)), -1, dtype=torch.int32)
    slots, c0 = [], 0
    for s, nb in enumerate(nblk):
        bt[s, :nb] = perm[c0 : c0 + nb]
        c0 += nb
        slots += [int(bt[s, p // bs]) * bs + p % bs for p in range(lens[s])]
    m.store_kvcache(k, v, kc, vc, torch.tensor(slots, dtype=torch.int32, device=cuda_device))
    rows = np.concatenate([np.arange(cu[s] + cached[s], cu[s + 1]) for s in range(len(lens))])
    lq = [l - c for l, c in zip(lens, cached)]
    cu_q = torch.tensor(np.concatenate([[0], np.cumsum(lq)]).astype(np.int32), device=cuda_device)
    cu_k = torch.tensor(cu, device=cuda_device)
    scale = d**-0.5
    out = m.prefill_varlen_attention(
        q[torch.from_numpy(rows).to(cuda_device)], None, None, cu_q, cu_k, max(lq), max(lens), scale,
        k_cache=kc, v_cache=vc, block_tables=bt.to(cuda_device),
    )
    full = orc.prefill_dense(_f32(q), _f32(k), _f32(v), cu, cu, scale)
    err = np.abs(_f32(out) - full[rows])
    assert np.isfinite(_f32(out)).all()
    assert err.max() <= TOL[dtype], (m.ops.last_kernel(), err.max(), err.mean())


# ---------------------------------------------------------------------------
# the layer: Attention.forward through the Context, prefill then decode steps
# ---------------------------------------------------------------------------


@pytest.mark.parametrize("case", _cases("layer"))
def test_attention_layer_flow_golden(m, cuda_device, case):
    g = load_golden("layer")
    dt = str(g[f"{case}.dtype"])
    h, kvh, d, steps, _ = (int(x) for x in g[f"{case}.meta"])
    scale = float(g[f"{case}.scale"])
    attn = m.Attention(h, d, scale, kvh)
    attn.k_cache = golden_to_torch(g[f"{case}.k_cache_in"], dt, cuda_device)
    attn.v_cache = golden_to_torch(g[f"{case}.v_cache_in"], dt, cuda_device)
    lens = np.diff(g[f"{case}.p.cu_seqlens"])
    cu = torch.from_numpy(g[f"{case}.p.cu_seqlens"]).to(cuda_device)
    m.set_context(
        is_prefill=True, max_seqlen_q=int(lens.max()), max_seqlen_k=int(lens.max()),
        cum_seqlens_q=cu, cum_seqlens_k=cu,
        slot_mapping=torch.from_numpy(g[f"{case}.p.slots"]).to(cuda_device),
        block_tables=torch.from_numpy(g[f"{case}.p.block_tables"]).to(cuda_device),
    )
    try:
        out = attn(*(golden_to_torch(g[f"{case}.p.{x}"], dt, cuda_device) for x in "qkv"))
    finally:
        m.reset_context()
    assert np.array_equal(_bits(attn.k_cache).ravel(), g[f"{case}.p.k_cache_out"].view(np.uint8).ravel())
    assert np.array_equal(_bits(attn.v_cache).ravel(), g[f"{case}.p.v_cache_out"].view(np.uint8).ravel())
    assert np.abs(_f32(out) - golden_to_float(g[f"{case}.p.out"], dt)).max() <= 2 * TOL[dt]
    for st in range(steps):
        pre = f"{case}.d{st}"
        m.set_context(
            is_prefill=False,
            slot_mapping=torch.from_numpy(g[f"{pre}.slots"]).to(cuda_device),
            context_lens=torch.from_numpy(g[f"{pre}.context_lens"]).to(cuda_device),
            block_tables=torch.from_numpy(g[f"{pre}.block_tables"]).to(cuda_device),
        )
        try:
            out = attn(*(golden_to_torch(g[f"{pre}.{x}"], dt, cuda_device) for x in "qkv"))
        finally:
            m.reset_context()
        assert np.abs(_f32(out) - golden_to_float(g[f"{pre}.out"], dt)).max() <= TOL[dt]
    assert np.array_equal(_bits(attn.k_cache).ravel(), g[f"{case}.k_cache_final"].view(np.uint8).ravel())
    assert np.array_equal(_bits(attn.v_cache).ravel(), g[f"{case}.v_cache_final"].view(np.uint8).ravel())


def test_attention_decode_before_cache_raises(m, cuda_device):
    attn = m.Attention(4, 64, 0.0, 4)
    assert abs(attn.scale - 64**-0.5) < 1e-12
    m.set_context(is_prefill=False)
    try:
        with pytest.raises(RuntimeError, match="KV cache has not been initialized"):
            attn(*(torch.zeros(1, 4, 64, device=cuda_device),) * 3)
    finally:
        m.reset_context()


# ---------------------------------------------------------------------------
# tcgen05 prefill kernel: selection, ragged shapes, BASELINE sizes
# ---------------------------------------------------------------------------


def _paged_prefill_problem(device, lens, cached, h, kvh, d, bs, dtype, seed, nan_fill=True):
    """Full K/V for every sequence stored through block tables; queries are the uncached suffix."""
    g = torch.Generator().manual_seed(seed)
    n = sum(lens)
    q = torch.randn(n, h, d, generator=g).to(dtype).to(device)
    k = torch.randn(n, kvh, d, generator=g).to(dtype).to(device)
    v = torch.randn(n, kvh, d, generator=g).to(dtype).to(device)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    nblk = [(x + bs - 1) // bs for x in lens]
    perm = torch.randperm(sum(nblk) + 3, generator=g)
    fill = float("nan") if nan_fill else 0.0
    kc = torch.full((perm.numel(), bs, kvh, d), fill, dtype=dtype, device=device)
    vc = torch.full((perm.numel(), bs, kvh, d), fill, dtype=dtype, device=device)
    bt = torch.full((len(lens), max(max(nblk), 1)), -1, dtype=torch.int32)
    slots, c0 = [], 0
    for s, nb in enumerate(nblk):
        bt[s, :nb] = perm[c0 : c0 + nb]
        c0 += nb
        slots += [int(bt[s, p // bs]) * bs + p % bs for p in range(lens[s])]
    rows = np.concatenate([np.arange(cu[s] + cached[s], cu[s + 1]) for s in range(len(lens))]).astype(np.int64)
    lq = [l - c for l, c in zip(lens, cached)]
    cu_q = np.concatenate([[0], np.cumsum(lq)]).astype(np.int32)
    return dict(q=q, k=k, v=v, cu=cu, cu_q=cu_q, rows=rows, lq=lq, kc=kc, vc=vc, bt=bt.to(device),
                slots=torch.tensor(slots, dtype=torch.int32, device=device))


def _oracle_rows(pr, seq, heads, h, kvh, scale):
    """float64 oracle for one sequence and a few heads (keeps the CPU time bounded)."""
    cu, group = pr["cu"], h // kvh
    a, b = int(cu[seq]), int(cu[seq + 1])
    outs = []
    for hd in heads:
        o = orc.prefill_dense(_f32(pr["q"][a:b, hd : hd + 1]), _f32(pr["k"][a:b, hd // group : hd // group + 1]),
                              _f32(pr["v"][a:b, hd // group : hd // group + 1]), [0, b - a], [0, b - a], scale)
        outs.append(o[:, 0])
    return np.stack(outs, axis=1)  # [len, heads, D] for ALL positions of the sequence


@pytest.mark.parametrize("bs", [64, 128, 256])
@pytest.mark.parametrize("dtype", ["bf16", "f16"])
@pytest.mark.parametrize("d", [128, 64])
def test_prefill_tcgen05_ragged_paged(m, cuda_device, bs, dtype, d):
    """Lq/Lk off the 128/64 grid, a fully cached sequence (Lq = 0), single-token suffix,
    never-written slots full of NaN, block sizes 64/128/256, head_dim 128 and 64 (GPT-2 / OPT heads)."""
    h, kvh = 8, 2
    lens = [300, 2 * bs, 65, 1000, 129, 64]
    cached = [0, 2 * bs, 64, 3 * bs if 3 * bs < 1000 else bs, 128, 0]
    pr = _paged_prefill_problem(cuda_device, lens, cached, h, kvh, d, bs, TDT[dtype], 21)
    m.store_kvcache(pr["k"], pr["v"], pr["kc"], pr["vc"], pr["slots"])
    scale = d**-0.5
    out = m.prefill_varlen_attention(
        pr["q"][torch.from_numpy(pr["rows"]).to(cuda_device)], None, None,
        torch.from_numpy(pr["cu_q"]).to(cuda_device), torch.from_numpy(pr["cu"]).to(cuda_device),
        max(pr["lq"]), max(lens), scale, k_cache=pr["kc"], v_cache=pr["vc"], block_tables=pr["bt"],
    )
    assert m.ops.last_kernel().startswith("prefill_tcgen05")
    full = orc.prefill_dense(_f32(pr["q"]), _f32(pr["k"]), _f32(pr["v"]), pr["cu"], pr["cu"], scale)
    got = _f32(out)
    assert np.isfinite(got).all()
    err = np.abs(got - full[pr["rows"]])
    assert err.max() <= TOL[dtype], (err.max(), err.mean())


def test_prefill_tcgen05_dense_strided_source(m, cuda_device):
    """No block tables (engine warm-up / layer tests): K/V are row-strided views of a fused qkv."""
    h, kvh, d = 4, 2, 128
    lens = [200, 1, 130, 64]
    n = sum(lens)
    g = torch.Generator().manual_seed(4)
    fused = torch.randn(n, (h + 2 * kvh) * d, generator=g).to(torch.bfloat16).to(cuda_device)
    q = fused[:, : h * d].view(n, h, d)
    k = fused[:, h * d : (h + kvh) * d].view(n, kvh, d)
    v = fused[:, (h + kvh) * d :].view(n, kvh, d)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    cut = torch.from_numpy(cu).to(cuda_device)
    out = m.prefill_varlen_attention(q, k, v, cut, cut, max(lens), max(lens), d**-0.5)
    assert m.ops.last_kernel().startswith("prefill_tcgen05")
    ref = orc.prefill_dense(_f32(q), _f32(k), _f32(v), cu, cu, d**-0.5)
    assert np.abs(_f32(out) - ref).max() <= TOL["bf16"]


@pytest.mark.parametrize("dtype,d", [("bf16", 128), ("f16", 128), ("bf16", 64), ("f32", 128)])
def test_prefill_fewer_keys_than_queries(m, cuda_device, dtype, d):
    """Lk < Lq (bottom-right causal mask, attention.py:584-587): the first Lq - Lk query rows see no key at all and
    must come out as zeros (documented deviation: torch gives NaN there), the others match the oracle.  The rows
    without keys sit in tiles that do have keys, i.e. they run through the softmax (FMA-pipe exponentials included)."""
    h, kvh = 4, 2
    lq = [300, 70, 129, 5]
    lk = [200, 69, 1, 5]
    g = torch.Generator().manual_seed(8)
    q = torch.randn(sum(lq), h, d, generator=g).to(TDT[dtype]).to(cuda_device)
    k = torch.randn(sum(lk), kvh, d, generator=g).to(TDT[dtype]).to(cuda_device)
    v = torch.randn(sum(lk), kvh, d, generator=g).to(TDT[dtype]).to(cuda_device)
    cu_q = np.concatenate([[0], np.cumsum(lq)]).astype(np.int32)
    cu_k = np.concatenate([[0], np.cumsum(lk)]).astype(np.int32)
    out = _f32(m.prefill_varlen_attention(q, k, v, torch.from_numpy(cu_q).to(cuda_device), torch.from_numpy(cu_k).to(cuda_device),
                                          max(lq), max(lk), d**-0.5))
    if dtype != "f32" and d == 128:
        assert m.ops.last_kernel().startswith("prefill_tcgen05")
    with np.errstate(invalid="ignore"):
        ref = orc.prefill_dense(_f32(q), _f32(k), _f32(v), cu_q, cu_k, d**-0.5)
    for s in range(len(lq)):
        blind = max(0, lq[s] - lk[s])
        a = int(cu_q[s])
        assert (out[a : a + blind] == 0).all(), (s, np.abs(out[a : a + blind]).max())
        err = np.abs(out[a + blind : a + lq[s]] - ref[a + blind : a + lq[s]])
        assert np.isfinite(out[a : a + lq[s]]).all() and (err.size == 0 or err.max() <= TOL[dtype]), (s, err.max())


@pytest.mark.parametrize(
    "name,lens,cached,h,kvh",
    [
        ("cfg3_8x(1024+1024)_tp1", [2048] * 8, [1024] * 8, 32, 8),
        ("cfg3_1x(4096+4096)_tp1", [8192], [4096], 32, 8),
        ("cfg3_tp8_rank", [2048] * 8, [1024] * 8, 4, 1),
        ("cfg5_rank_16k_tokens", [1024] * 16, [0] * 16, 8, 1),
    ],
)
def test_prefill_full_size_against_oracle_samples(m, cuda_device, name, lens, cached, h, kvh):
    """BASELINE.json prefill sizes: 8k new tokens with 50 % prefix hits (cfg3), 16k tokens (cfg5)."""
    d, bs = 128, 64
    pr = _paged_prefill_problem(cuda_device, lens, cached, h, kvh, d, bs, torch.bfloat16, 33)
    m.store_kvcache(pr["k"], pr["v"], pr["kc"], pr["vc"], pr["slots"])
    scale = d**-0.5
    out = m.prefill_varlen_attention(
        pr["q"][torch.from_numpy(pr["rows"]).to(cuda_device)], None, None,
        torch.from_numpy(pr["cu_q"]).to(cuda_device), torch.from_numpy(pr["cu"]).to(cuda_device),
        max(pr["lq"]), max(lens), scale, k_cache=pr["kc"], v_cache=pr["vc"], block_tables=pr["bt"],
    )
    assert m.ops.last_kernel().startswith("prefill_tcgen05")
    got = _f32(out)
    assert np.isfinite(got).all()
    heads = sorted({0, h // 2, h - 1})
    worst = 0.0
    for seq in sorted({0, len(lens) - 1}):
        ref = _oracle_rows(pr, seq, heads, h, kvh, scale)  # all positions of the sequence
        a = int(pr["cu_q"][seq])
        mine = got[a : a + pr["lq"][seq]][:, heads]
        err = np.abs(mine - ref[cached[seq] :])
        worst = max(worst, float(err.max()))
        assert err.max() <= TOL["bf16"], (name, seq, err.max(), err.mean())
    # linearity in V (size-independent property): out(V) + out(V2) == out(V + V2) up to rounding
    # is implied by the oracle check on sampled rows; here check row-sums of P == 1 via V = ones
    ones = torch.ones_like(pr["vc"])
    out1 = m.prefill_varlen_attention(
        pr["q"][torch.from_numpy(pr["rows"]).to(cuda_device)], None, None,
        torch.from_numpy(pr["cu_q"]).to(cuda_device), torch.from_numpy(pr["cu"]).to(cuda_device),
        max(pr["lq"]), max(lens), scale, k_cache=pr["kc"], v_cache=ones, block_tables=pr["bt"],
    )
    assert (out1.float() - 1.0).abs().max().item() <= 1e-2


def test_prefill_kernel_selection(m, cuda_device):
    """bf16/fp16 with head_dim 128 or 64 -> tcgen05; everything else -> the generic kernel."""
    for dt, d, want in [(torch.bfloat16, 128, "prefill_tcgen05"), (torch.float16, 128, "prefill_tcgen05"),
                        (torch.float32, 128, "prefill_generic"), (torch.bfloat16, 64, "prefill_tcgen05"),
                        (torch.float16, 64, "prefill_tcgen05"), (torch.bfloat16, 96, "prefill_generic")]:
        q = torch.randn(10, 2, d, device=cuda_device).to(dt)
        cu = torch.tensor([0, 10], dtype=torch.int32, device=cuda_device)
        m.prefill_varlen_attention(q, q, q, cu, cu, 10, 10, 0.1)
        assert m.ops.last_kernel().startswith(want)


def test_second_device_in_one_process(m, cuda_device):
    """One process driving two GPUs (a single-process multi-GPU harness): the per-device caches of the library
    (dynamic shared-memory opt-in, SM count) must not carry over from the first device — the decode and prefill
    kernels launched on cuda:1 after cuda:0 give the same results as on cuda:0."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    d, bs, kvh, g = 128, 64, 2, 4
    h = kvh * g
    ctxs = [300, 65, 1000]
    outs = []
    for dev in (torch.device("cuda", 0), torch.device("cuda", 1)):
        with torch.cuda.device(dev):
            kc, vc, bt, cl = _random_paged(dev, len(ctxs), ctxs, kvh, d, bs, torch.bfloat16, 5)
            gen = torch.Generator().manual_seed(3)
            kc.copy_(torch.randn(kc.shape, generator=gen).to(torch.bfloat16))
            vc.copy_(torch.randn(vc.shape, generator=gen).to(torch.bfloat16))
            q = torch.randn(len(ctxs), h, d, generator=gen).to(torch.bfloat16).to(dev)
            dec = m.paged_decode_attention(q, kc, vc, bt, cl, d**-0.5, num_splits=3)
            n = sum(ctxs)
            qp = torch.randn(n, h, d, generator=gen).to(torch.bfloat16).to(dev)
            kp = torch.randn(n, kvh, d, generator=gen).to(torch.bfloat16).to(dev)
            cu = torch.tensor(np.concatenate([[0], np.cumsum(ctxs)]).astype(np.int32), device=dev)
            pre = m.prefill_varlen_attention(qp, kp, kp, cu, cu, max(ctxs), max(ctxs), d**-0.5)
            assert m.ops.last_kernel().startswith("prefill_tcgen05")
            torch.cuda.synchronize(dev)
            outs.append((dec.cpu(), pre.cpu()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
