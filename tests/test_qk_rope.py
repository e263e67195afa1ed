"""q/k RMSNorm + RoPE + store: oracle pinned to the reference, CUDA kernel to the oracle.

Golden vectors: tests/golden/qkrope.npz, recorded from the unmodified reference's ``RMSNorm``,
``RotaryEmbedding`` and ``store_kv_cache`` on CPU (tests/golden/gen_golden.py qkrope).

Bars.  V cache: bit-exact (bits are moved).  q, k and the K cache: the reference's fp32 operation
order and rounding points are reproduced, the one freedom is the summation order inside ``mean``;
so bf16/fp16 results may differ from the reference by one unit in the last place on a small
fraction of elements (bar: <= 1 ulp everywhere, >= 99 % identical); without q/k norm there is no
sum and the result is bit-exact.  fp32: |diff| <= 2e-6 * max(1, |x|).
"""

import numpy as np
import pytest
import torch

from oracle import qk_rope_oracle as qro
from tests.conftest import golden_to_torch, load_golden

TDT = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}


def _cases():
    return [str(c) for c in load_golden("qkrope")["cases"]]


def _split(g, case):
    n, h, kvh, d, norm, max_pos, slots_n = (int(x) for x in g[f"{case}.meta"])
    fused = g[f"{case}.fused"]
    q = fused[:, : h * d].reshape(n, h, d)
    k = fused[:, h * d : (h + kvh) * d].reshape(n, kvh, d)
    v = fused[:, (h + kvh) * d :].reshape(n, kvh, d)
    return (n, h, kvh, d, norm, max_pos, slots_n), q, k, v


def _ulp_diff(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """Distance in representable values between two arrays of 16-bit float bit patterns."""
    def key(x):
        x = x.view(np.uint16).astype(np.int32)
        return np.where(x & 0x8000, -(x & 0x7FFF), x)
    return np.abs(key(a) - key(b))


def _close(got: np.ndarray, want: np.ndarray, dt: str, exact: bool):
    if exact:
        assert np.array_equal(got.view(np.uint8), want.view(np.uint8))
    elif dt == "f32":
        assert np.all(np.abs(got - want) <= 2e-6 * np.maximum(1.0, np.abs(want)))
    else:
        ulp = _ulp_diff(got, want)
        assert ulp.max() <= 1, f"max ulp distance {ulp.max()}"
        assert (ulp == 0).mean() >= 0.99


# ---------------------------------------------------------------------------
# CPU: the oracle against the reference's recorded outputs
# ---------------------------------------------------------------------------


@pytest.mark.parametrize("case", _cases())
def test_oracle_matches_reference_golden(case):
    g = load_golden("qkrope")
    (n, h, kvh, d, norm, max_pos, slots_n), q, k, v = _split(g, case)
    dt = str(g[f"{case}.dtype"])
    kc, vc = g[f"{case}.k_cache_in"].copy(), g[f"{case}.v_cache_in"].copy()
    qw = g[f"{case}.q_weight"] if norm else None
    kw = g[f"{case}.k_weight"] if norm else None
    q_out, k_out = qro.qk_norm_rope_store(q, k, v, g[f"{case}.positions"], g[f"{case}.cos"], g[f"{case}.sin"],
                                          qw, kw, float(g[f"{case}.eps"]), g[f"{case}.slots"], kc, vc, dt)
    exact = not norm
    _close(q_out, g[f"{case}.q_out"], dt, exact)
    _close(k_out, g[f"{case}.k_out"], dt, exact)
    _close(kc, g[f"{case}.k_cache_out"], dt, exact)
    assert np.array_equal(vc.view(np.uint8), g[f"{case}.v_cache_out"].view(np.uint8))


def test_oracle_rope_tables_follow_the_reference_formula():
    g = load_golden("qkrope")
    for case in _cases():
        d, max_pos = int(g[f"{case}.meta"][3]), int(g[f"{case}.meta"][5])
        cos, sin = qro.rope_cache(d, max_pos, float(g[f"{case}.base"]))
        # libm vs numpy pow/cos differ in the last bits at large angles: not a bit-level pin
        assert np.abs(cos - g[f"{case}.cos"]).max() < 5e-5 and np.abs(sin - g[f"{case}.sin"]).max() < 5e-5


def test_op_refuses_cpu_tensors():
    import minivllm_b200 as m

    z = torch.zeros(2, 1, 8)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m.ops.qk_norm_rope_store(z, z, z, torch.zeros(2, dtype=torch.int64), torch.zeros(4, 4), torch.zeros(4, 4))


# ---------------------------------------------------------------------------
# GPU: the kernel (through the C ABI) against golden vectors and the oracle
# ---------------------------------------------------------------------------


def _np_bits(t: torch.Tensor, dt: str) -> np.ndarray:
    t = t.detach().contiguous().cpu()
    if dt == "bf16":
        return t.view(torch.int16).numpy().view(np.uint16)
    return t.numpy()


@pytest.mark.gpu
@pytest.mark.parametrize("inplace", [True, False])
@pytest.mark.parametrize("case", _cases())
def test_kernel_matches_reference_golden(cuda_device, case, inplace):
    import minivllm_b200 as m

    g = load_golden("qkrope")
    (n, h, kvh, d, norm, max_pos, slots_n), _, _, _ = _split(g, case)
    dt = str(g[f"{case}.dtype"])
    fused = golden_to_torch(g[f"{case}.fused"], dt, cuda_device)  # q, k, v stay row-strided views, as in the model
    q, k, v = fused.split([h * d, kvh * d, kvh * d], dim=-1)
    q, k, v = q.view(n, h, d), k.view(n, kvh, d), v.view(n, kvh, d)
    v_before = v.clone()
    kc = golden_to_torch(g[f"{case}.k_cache_in"], dt, cuda_device)
    vc = golden_to_torch(g[f"{case}.v_cache_in"], dt, cuda_device)
    qw = golden_to_torch(g[f"{case}.q_weight"], dt, cuda_device) if norm else None
    kw = golden_to_torch(g[f"{case}.k_weight"], dt, cuda_device) if norm else None
    q_out, k_out = m.ops.qk_norm_rope_store(
        q, k, v, torch.from_numpy(g[f"{case}.positions"]).to(cuda_device),
        torch.from_numpy(g[f"{case}.cos"]).to(cuda_device), torch.from_numpy(g[f"{case}.sin"]).to(cuda_device),
        qw, kw, float(g[f"{case}.eps"]), torch.from_numpy(g[f"{case}.slots"]).to(cuda_device), kc, vc, inplace=inplace)
    torch.cuda.synchronize()
    assert m.ops.last_kernel() == "qk_norm_rope_store"
    assert (q_out.data_ptr() == q.data_ptr()) == inplace
    exact = not norm
    _close(_np_bits(q_out, dt), g[f"{case}.q_out"], dt, exact)
    _close(_np_bits(k_out, dt), g[f"{case}.k_out"], dt, exact)
    _close(_np_bits(kc, dt), g[f"{case}.k_cache_out"], dt, exact)
    assert np.array_equal(_np_bits(vc, dt).view(np.uint8), g[f"{case}.v_cache_out"].view(np.uint8))
    assert torch.equal(v, v_before)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(256, 16, 8, 128, True), (1000, 28, 4, 128, False), (513, 4, 1, 128, True),
                                   (64, 12, 12, 64, True), (33, 2, 2, 256, True)])
def test_kernel_matches_oracle_seeded(cuda_device, shape):
    """Model-sized rows (cfg2 decode batch, Qwen2-7B group 7, TP8 slice, OPT-like D=64, D=256)."""
    import minivllm_b200 as m

    n, h, kvh, d, norm = shape
    gen = torch.Generator().manual_seed(n * 7 + h)
    fused = (torch.randn(n, (h + 2 * kvh) * d, generator=gen) * 1.5).to(torch.bfloat16)
    max_pos, slots_n = 4096, 2 * n + 64
    pos = torch.randint(0, max_pos, (n,), generator=gen)
    slots = torch.randperm(slots_n, generator=gen)[:n].to(torch.int32)
    slots[::7] = -1
    cos, sin = qro.rope_cache(d, max_pos, 1e6)
    qw = (1 + 0.1 * torch.randn(d, generator=gen)).to(torch.bfloat16) if norm else None
    kw = (1 + 0.1 * torch.randn(d, generator=gen)).to(torch.bfloat16) if norm else None
    kc0 = torch.randn(slots_n, kvh * d, generator=gen).to(torch.bfloat16)
    vc0 = torch.randn(slots_n, kvh * d, generator=gen).to(torch.bfloat16)

    bits = lambda t: t.contiguous().view(torch.int16).numpy().view(np.uint16)  # noqa: E731
    qn, kn, vn = (bits(x).reshape(n, -1, d) for x in fused.split([h * d, kvh * d, kvh * d], dim=-1))
    kc_ref, vc_ref = bits(kc0).copy(), bits(vc0).copy()
    q_ref, k_ref = qro.qk_norm_rope_store(qn, kn, vn, pos.numpy(), cos, sin, bits(qw) if norm else None,
                                          bits(kw) if norm else None, 1e-6, slots.numpy(), kc_ref, vc_ref, "bf16")

    dev = cuda_device
    f = fused.to(dev)
    q, k, v = f.split([h * d, kvh * d, kvh * d], dim=-1)
    kc, vc = kc0.to(dev), vc0.to(dev)
    q_out, k_out = m.ops.qk_norm_rope_store(
        q.view(n, h, d), k.view(n, kvh, d), v.view(n, kvh, d), pos.to(dev), torch.from_numpy(cos).to(dev),
        torch.from_numpy(sin).to(dev), qw.to(dev) if norm else None, kw.to(dev) if norm else None, 1e-6,
        slots.to(dev), kc, vc)
    torch.cuda.synchronize()
    exact = not norm
    _close(_np_bits(q_out, "bf16"), q_ref, "bf16", exact)
    _close(_np_bits(k_out, "bf16"), k_ref, "bf16", exact)
    _close(_np_bits(kc, "bf16"), kc_ref, "bf16", exact)
    assert np.array_equal(_np_bits(vc, "bf16"), vc_ref)


@pytest.mark.gpu
def test_kernel_without_store_and_argument_errors(cuda_device):
    import minivllm_b200 as m

    dev = cuda_device
    n, h, kvh, d = 5, 2, 1, 128
    q = torch.randn(n, h, d, device=dev, dtype=torch.bfloat16)
    k = torch.randn(n, kvh, d, device=dev, dtype=torch.bfloat16)
    v = torch.randn(n, kvh, d, device=dev, dtype=torch.bfloat16)
    cos, sin = (torch.from_numpy(x).to(dev) for x in qro.rope_cache(d, 64, 1e4))
    pos = torch.arange(n, device=dev)
    q_out, k_out = m.ops.qk_norm_rope_store(q, k, v, pos, cos, sin, inplace=False)  # warm-up path: no cache
    torch.cuda.synchronize()
    assert torch.equal(q_out[0], q[0]) and torch.equal(k_out[0], k[0])  # position 0: cos 1, sin 0
    assert not torch.equal(q_out[1], q[1])
    with pytest.raises(ValueError, match="positions size"):
        m.ops.qk_norm_rope_store(q, k, v, pos[:2], cos, sin)
    with pytest.raises(ValueError, match="given together"):
        m.ops.qk_norm_rope_store(q, k, v, pos, cos, sin, torch.ones(d, device=dev, dtype=torch.bfloat16), None)
    with pytest.raises(ValueError, match="slot_mapping size"):
        c = torch.zeros(8, kvh * d, device=dev, dtype=torch.bfloat16)
        m.ops.qk_norm_rope_store(q, k, v, pos, cos, sin, None, None, 1e-6, torch.zeros(2, dtype=torch.int32, device=dev), c, c.clone())
    with pytest.raises(ValueError, match="float32"):
        m.ops.qk_norm_rope_store(q, k, v, pos, cos.half(), sin.half())
