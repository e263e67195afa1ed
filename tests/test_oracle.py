"""The CPU oracle against the golden vectors recorded from the reference (CPU only)."""

import numpy as np
import pytest

from oracle import paged_attention_oracle as orc
from tests.conftest import golden_to_float, load_golden

# The reference computes in the tensor dtype (bf16 matmul + bf16 softmax for bf16
# inputs, attention.py:575-594); the oracle is float64.  Tolerance per dtype:
TOL = {"f32": 2e-5, "f16": 4e-3, "bf16": 2e-2}


def _cases(g):
    return [str(c) for c in g["cases"]]


@pytest.mark.parametrize("case", _cases(load_golden("store")))
def test_store_bit_exact(case):
    g = load_golden("store")
    kc = g[f"{case}.k_cache_in"].copy()
    vc = g[f"{case}.v_cache_in"].copy()
    orc.store_kvcache(g[f"{case}.key"], g[f"{case}.value"], kc, vc, g[f"{case}.slots"])
    assert np.array_equal(kc.view(np.uint8), g[f"{case}.k_cache_out"].view(np.uint8))
    assert np.array_equal(vc.view(np.uint8), g[f"{case}.v_cache_out"].view(np.uint8))


def test_store_size_mismatch_raises():
    k = np.zeros((3, 2, 4), np.float32)
    c = np.zeros((8, 8), np.float32)
    with pytest.raises(ValueError, match="slot_mapping size"):
        orc.store_kvcache(k, k, c, c.copy(), np.array([1, 2]))


@pytest.mark.parametrize("case", _cases(load_golden("decode")))
def test_decode_matches_reference(case):
    g = load_golden("decode")
    dt = str(g[f"{case}.dtype"])
    out = orc.paged_decode(
        golden_to_float(g[f"{case}.q"], dt),
        golden_to_float(g[f"{case}.k_cache"], dt),
        golden_to_float(g[f"{case}.v_cache"], dt),
        g[f"{case}.block_tables"],
        g[f"{case}.context_lens"],
        float(g[f"{case}.scale"]),
    )
    ref = golden_to_float(g[f"{case}.out"], dt)
    assert np.abs(out - ref).max() <= TOL[dt]


def test_decode_error_contracts():
    # tests/test_page_attention.py:301-329 of the reference
    q = np.zeros((2, 8, 64), np.float32)
    with pytest.raises(ValueError, match="empty"):
        orc.paged_decode(q, np.zeros((0,)), np.zeros((0,)), np.zeros((2, 2), np.int32), np.array([8, 8]), 0.125)
    kc = np.ones((4, 16, 8, 64), np.float32)
    with pytest.raises(RuntimeError, match="No valid KV cache"):
        orc.paged_decode(q, kc, kc, np.full((2, 2), -1, np.int32), np.array([8, 8]), 0.125)
    z = orc.paged_decode(q, kc, kc, np.zeros((2, 0), np.int32), np.array([8, 8]), 0.125)
    assert z.shape == (2, 8, 64) and not z.any()


@pytest.mark.parametrize("case", _cases(load_golden("prefill")))
def test_prefill_dense_matches_reference(case):
    g = load_golden("prefill")
    dt = str(g[f"{case}.dtype"])
    cu = g[f"{case}.cu_seqlens"]
    out = orc.prefill_dense(
        golden_to_float(g[f"{case}.q"], dt),
        golden_to_float(g[f"{case}.k"], dt),
        golden_to_float(g[f"{case}.v"], dt),
        cu,
        cu,
        float(g[f"{case}.scale"]),
    )
    ref = golden_to_float(g[f"{case}.out"], dt)
    assert np.abs(out - ref).max() <= TOL[dt]


@pytest.mark.parametrize("case", _cases(load_golden("layer")))
def test_layer_flow_matches_reference(case):
    """store -> prefill -> (store -> decode) x steps, caches bit-exact, outputs in tolerance."""
    g = load_golden("layer")
    dt = str(g[f"{case}.dtype"])
    h, kvh, d, steps, _ = (int(x) for x in g[f"{case}.meta"])
    scale = float(g[f"{case}.scale"])
    kc = g[f"{case}.k_cache_in"].copy()
    vc = g[f"{case}.v_cache_in"].copy()
    orc.store_kvcache(g[f"{case}.p.k"], g[f"{case}.p.v"], kc, vc, g[f"{case}.p.slots"])
    assert np.array_equal(kc, g[f"{case}.p.k_cache_out"])
    assert np.array_equal(vc, g[f"{case}.p.v_cache_out"])
    cu = g[f"{case}.p.cu_seqlens"]
    qf = golden_to_float(g[f"{case}.p.q"], dt)
    dense = orc.prefill_dense(qf, golden_to_float(g[f"{case}.p.k"], dt), golden_to_float(g[f"{case}.p.v"], dt), cu, cu, scale)
    ref = golden_to_float(g[f"{case}.p.out"], dt)
    assert np.abs(dense - ref).max() <= TOL[dt]
    # the same prefill read back through the block tables (flash call-site semantics)
    paged = orc.prefill_paged(qf, golden_to_float(kc, dt), golden_to_float(vc, dt), g[f"{case}.p.block_tables"], cu, cu, scale)
    assert np.abs(paged - dense).max() <= 1e-12
    for st in range(steps):
        pre = f"{case}.d{st}"
        orc.store_kvcache(g[f"{pre}.k"], g[f"{pre}.v"], kc, vc, g[f"{pre}.slots"])
        out = orc.paged_decode(
            golden_to_float(g[f"{pre}.q"], dt), golden_to_float(kc, dt), golden_to_float(vc, dt),
            g[f"{pre}.block_tables"], g[f"{pre}.context_lens"], scale,
        )
        assert np.abs(out - golden_to_float(g[f"{pre}.out"], dt)).max() <= TOL[dt]
    assert np.array_equal(kc, g[f"{case}.k_cache_final"])
    assert np.array_equal(vc, g[f"{case}.v_cache_final"])


def test_prefix_prefill_equals_suffix_of_dense():
    """Pins the restated prefix-hit prefill to the golden-pinned dense prefill."""
    rng = np.random.default_rng(7)
    h, kvh, d, bs = 4, 2, 32, 16
    lens, cached = [50, 33, 64], [32, 0, 48]
    n_full = sum(lens)
    q = rng.standard_normal((n_full, h, d)).astype(np.float32)
    k = rng.standard_normal((n_full, kvh, d)).astype(np.float32)
    v = rng.standard_normal((n_full, kvh, d)).astype(np.float32)
    cu = np.concatenate([[0], np.cumsum(lens)])
    dense = orc.prefill_dense(q, k, v, cu, cu, 0.2)
    nblk = [(x + bs - 1) // bs for x in lens]
    perm = rng.permutation(sum(nblk) + 2)
    kc = rng.standard_normal((perm.size, bs, kvh, d)).astype(np.float32)
    vc = rng.standard_normal((perm.size, bs, kvh, d)).astype(np.float32)
    bt = np.full((len(lens), max(nblk)), -1, np.int32)
    slots, c = [], 0
    for s, nb in enumerate(nblk):
        bt[s, :nb] = perm[c : c + nb]
        c += nb
        slots += [int(bt[s, p // bs]) * bs + p % bs for p in range(lens[s])]
    orc.store_kvcache(k, v, kc, vc, np.array(slots))
    rows = np.concatenate([np.arange(cu[s] + cached[s], cu[s + 1]) for s in range(len(lens))])
    cu_q = np.concatenate([[0], np.cumsum([l - c_ for l, c_ in zip(lens, cached)])])
    paged = orc.prefill_paged(q[rows], kc, vc, bt, cu_q, cu, 0.2)
    assert np.abs(paged - dense[rows]).max() < 1e-12


def test_slot_helpers_follow_executor_layout():
    # tests/test_integration.py:105-113 of the reference: 64-token blocks
    bt = [7, 3, 9]
    assert orc.prefill_slots(bt, 130, 0) == list(range(448, 512)) + list(range(192, 256)) + [576, 577]
    assert orc.prefill_slots(bt, 130, 64) == list(range(192, 256)) + [576, 577]
    assert orc.decode_slot(bt, 130) == 577
    assert orc.decode_slot([7, 3], 128) == 3 * 64 + 63
    assert orc.decode_slot([7, 3, 9], 129) == 576


def test_bf16_roundtrip_matches_torch():
    import torch

    x = torch.randn(4096, generator=torch.Generator().manual_seed(0)) * 50
    bits = orc.f32_to_bf16_bits(x.numpy())
    tb = x.to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(bits, tb)
    assert np.array_equal(orc.bf16_bits_to_f32(bits), x.to(torch.bfloat16).float().numpy())


# ---------------------------------------------------------------------------
# the torch-CPU port used for cpu_baseline timing: same golden vectors
# ---------------------------------------------------------------------------
from oracle import eager_port as port  # noqa: E402
from tests.conftest import golden_to_torch  # noqa: E402


@pytest.mark.parametrize("case", _cases(load_golden("store")))
def test_eager_port_store(case):
    import torch

    g = load_golden("store")
    dt = str(g[f"{case}.dtype"])
    kc = golden_to_torch(g[f"{case}.k_cache_in"], dt).clone()
    vc = golden_to_torch(g[f"{case}.v_cache_in"], dt).clone()
    port.store_kvcache(golden_to_torch(g[f"{case}.key"], dt), golden_to_torch(g[f"{case}.value"], dt),
                       kc, vc, torch.from_numpy(g[f"{case}.slots"]))
    assert torch.equal(kc, golden_to_torch(g[f"{case}.k_cache_out"], dt))
    assert torch.equal(vc, golden_to_torch(g[f"{case}.v_cache_out"], dt))


@pytest.mark.parametrize("case", _cases(load_golden("decode")))
def test_eager_port_decode(case):
    import torch

    g = load_golden("decode")
    dt = str(g[f"{case}.dtype"])
    out = port.paged_decode(
        golden_to_torch(g[f"{case}.q"], dt), golden_to_torch(g[f"{case}.k_cache"], dt),
        golden_to_torch(g[f"{case}.v_cache"], dt), torch.from_numpy(g[f"{case}.block_tables"]),
        torch.from_numpy(g[f"{case}.context_lens"]), float(g[f"{case}.scale"]),
    )
    # same operators on the same inputs as the reference: identical up to thread-count reassociation
    assert (out.float() - golden_to_torch(g[f"{case}.out"], dt).float()).abs().max().item() <= TOL[dt] / 4


@pytest.mark.parametrize("case", _cases(load_golden("prefill")))
def test_eager_port_prefill(case):
    import torch

    g = load_golden("prefill")
    dt = str(g[f"{case}.dtype"])
    cu = torch.from_numpy(g[f"{case}.cu_seqlens"])
    out = port.prefill_dense(golden_to_torch(g[f"{case}.q"], dt), golden_to_torch(g[f"{case}.k"], dt),
                             golden_to_torch(g[f"{case}.v"], dt), cu, cu, float(g[f"{case}.scale"]))
    assert (out.float() - golden_to_torch(g[f"{case}.out"], dt).float()).abs().max().item() <= TOL[dt] / 4
