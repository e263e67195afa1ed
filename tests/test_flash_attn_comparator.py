"""The reference's real CUDA path as an independent comparator (tests only — the product never calls it).

On CUDA the reference's ``Attention.forward`` calls flash-attn 2 (attention.py:422-433 prefill with
``block_table=``, :439-447 decode ``flash_attn_with_kvcache``).  Whether the wheel installed in this image
(flash-attn 2.8.x) runs on sm_100 is probed here, not assumed; the outcome is written to
``gpurun_out/flash_attn_probe.json`` and summarised in DESIGN.md section 6.

If it runs it is (i) the only independent numerical check of prefix-hit prefill (otherwise pinned only through
"rows equal the suffix of a dense prefill") and (ii) a second check of paged decode.  flash-attn 2 pages K/V
in blocks that must be a multiple of 256 tokens; the reference's ``kvcache_block_size`` is 64, so the same
logical K/V is laid out twice: 64-token pages for ``mvb200`` and 256-token pages for flash-attn.
"""

import json
import os

import numpy as np
import pytest
import torch

from tests.conftest import ROOT

pytestmark = pytest.mark.gpu

PROBE_OUT = os.path.join(ROOT, "gpurun_out", "flash_attn_probe.json")


def _record(key, value):
    os.makedirs(os.path.dirname(PROBE_OUT), exist_ok=True)
    data = {}
    if os.path.exists(PROBE_OUT):
        try:
            data = json.load(open(PROBE_OUT))
        except Exception:
            data = {}
    data[key] = value
    json.dump(data, open(PROBE_OUT, "w"), indent=1)


@pytest.fixture(scope="module")
def m():
    import minivllm_b200

    return minivllm_b200


_CHILD_PROBE = r"""
import torch, flash_attn
from flash_attn import flash_attn_varlen_func, flash_attn_with_kvcache
dev = torch.device('cuda:0')
q = torch.randn(2, 1, 4, 128, device=dev, dtype=torch.bfloat16)
kc = torch.randn(4, 256, 2, 128, device=dev, dtype=torch.bfloat16)
bt = torch.tensor([[0, 1], [2, 3]], dtype=torch.int32, device=dev)
o = flash_attn_with_kvcache(q, kc, kc, cache_seqlens=torch.tensor([300, 17], dtype=torch.int32, device=dev),
                            block_table=bt, softmax_scale=0.1, causal=True)
cu = torch.tensor([0, 40, 100], dtype=torch.int32, device=dev)
cuk = torch.tensor([0, 300, 317], dtype=torch.int32, device=dev)
o2 = flash_attn_varlen_func(torch.randn(100, 4, 128, device=dev, dtype=torch.bfloat16), kc, kc, cu_seqlens_q=cu,
                            cu_seqlens_k=cuk, max_seqlen_q=60, max_seqlen_k=300, softmax_scale=0.1, causal=True,
                            block_table=bt)
torch.cuda.synchronize()
assert torch.isfinite(o.float()).all() and torch.isfinite(o2.float()).all()
print('FA_OK', flash_attn.__version__)
"""


@pytest.fixture(scope="module")
def fa():
    """flash-attn is first exercised in a CHILD process: a wheel without sm_100 code could fault the CUDA
    context, which must not take the rest of the GPU suite down with it."""
    import subprocess
    import sys

    r = subprocess.run([sys.executable, "-c", _CHILD_PROBE], capture_output=True, text=True, timeout=600)
    if "FA_OK" not in r.stdout:
        _record("child_probe", {"ok": False, "rc": r.returncode, "stderr_tail": r.stderr[-1500:]})
        pytest.skip("flash-attn does not run on this GPU (see gpurun_out/flash_attn_probe.json)")
    _record("child_probe", {"ok": True, "version": r.stdout.split("FA_OK")[-1].strip()})
    from flash_attn import flash_attn_varlen_func, flash_attn_with_kvcache

    return flash_attn_varlen_func, flash_attn_with_kvcache


def _paged(k, v, lens, page, gen, device, fill=float("nan")):
    """Lay dense per-sequence K/V rows [sum(lens), KVH, D] out in shuffled pages of ``page`` tokens."""
    kvh, d = k.shape[1:]
    nblk = [(x + page - 1) // page for x in lens]
    perm = torch.randperm(sum(nblk) + 3, generator=gen)
    kc = torch.full((perm.numel(), page, kvh, d), fill, dtype=k.dtype, device=device)
    vc = torch.full_like(kc, fill)
    bt = torch.zeros((len(lens), max(nblk)), dtype=torch.int32)
    c0 = r0 = 0
    for s, (n, nb) in enumerate(zip(lens, nblk)):
        bt[s, :nb] = perm[c0:c0 + nb]
        c0 += nb
        for j in range(nb):
            rows = slice(r0 + j * page, r0 + min((j + 1) * page, n))
            cnt = rows.stop - rows.start
            kc[int(bt[s, j]), :cnt] = k[rows]
            vc[int(bt[s, j]), :cnt] = v[rows]
        r0 += n
    return kc, vc, bt.to(device)


def _try(name, fn):
    try:
        out = fn()
        torch.cuda.synchronize()
        _record(name, {"ok": True})
        return out
    except Exception as e:  # noqa: BLE001
        torch.cuda.synchronize()
        _record(name, {"ok": False, "error": (type(e).__name__ + ": " + str(e))[:600]})
        return None


def test_probe_reference_block_size_64(fa, cuda_device):
    """The reference's own configuration (kvcache_block_size = 64) on its flash-attn decode call site."""
    _, with_kvcache = fa
    g = torch.Generator().manual_seed(1)
    k = torch.randn(200, 2, 128, generator=g).to(torch.bfloat16).to(cuda_device)
    kc, vc, bt = _paged(k, k, [200], 64, g, cuda_device, fill=0.0)
    q = torch.randn(1, 1, 4, 128, generator=g).to(torch.bfloat16).to(cuda_device)
    out = _try("decode_block64", lambda: with_kvcache(
        q, kc, vc, cache_seqlens=torch.tensor([200], dtype=torch.int32, device=cuda_device),
        block_table=bt, softmax_scale=128**-0.5, causal=True))
    if out is None:
        pytest.skip("flash_attn_with_kvcache rejects 64-token pages (recorded in flash_attn_probe.json)")


@pytest.mark.parametrize("h,kvh", [(8, 2), (7, 1), (32, 8)])
def test_prefix_hit_prefill_matches_flash_attn(m, fa, cuda_device, h, kvh):
    """P1 against attention.py:422-433: uncached-suffix queries, keys through block tables (50 % hits)."""
    varlen, _ = fa
    d = 128
    lens = [300, 128, 65, 1024, 2048]
    cached = [128, 64, 64, 512, 1024]
    g = torch.Generator().manual_seed(9)
    n = sum(lens)
    q, k, v = (torch.randn(n, x, d, generator=g).to(torch.bfloat16).to(cuda_device) for x in (h, kvh, kvh))
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    rows = np.concatenate([np.arange(cu[s] + cached[s], cu[s + 1]) for s in range(len(lens))])
    qs = q[torch.from_numpy(rows).to(cuda_device)].contiguous()
    lq = [a - c for a, c in zip(lens, cached)]
    cu_q = torch.tensor(np.concatenate([[0], np.cumsum(lq)]).astype(np.int32), device=cuda_device)
    cu_k = torch.tensor(cu, device=cuda_device)
    scale = d**-0.5
    kc64, vc64, bt64 = _paged(k, v, lens, 64, g, cuda_device)
    ours = m.prefill_varlen_attention(qs, None, None, cu_q, cu_k, max(lq), max(lens), scale,
                                      k_cache=kc64, v_cache=vc64, block_tables=bt64)
    kc256, vc256, bt256 = _paged(k, v, lens, 256, g, cuda_device, fill=0.0)
    theirs = _try(f"prefill_block256_h{h}", lambda: varlen(
        qs, kc256, vc256, max_seqlen_q=max(lq), cu_seqlens_q=cu_q, max_seqlen_k=max(lens), cu_seqlens_k=cu_k,
        softmax_scale=scale, causal=True, block_table=bt256))
    if theirs is None:
        pytest.skip("flash_attn_varlen_func(block_table=) does not run on this GPU (recorded)")
    diff = (ours.float() - theirs.float()).abs()
    _record(f"prefill_h{h}_kvh{kvh}", {"max_abs": diff.max().item(), "mean_abs": diff.mean().item(),
                                      "kernel": m.ops.last_kernel()})
    assert torch.isfinite(ours.float()).all()
    assert diff.max().item() <= 2e-2, (diff.max().item(), diff.mean().item())


@pytest.mark.parametrize("h,kvh,ctx", [(16, 8, [2048, 1, 64, 65, 1500, 777]), (28, 4, [8192, 4000]), (8, 1, [513] * 5)])
def test_paged_decode_matches_flash_attn(m, fa, cuda_device, h, kvh, ctx):
    """D1 against attention.py:439-447."""
    _, with_kvcache = fa
    d = 128
    g = torch.Generator().manual_seed(11)
    n = sum(ctx)
    k, v = (torch.randn(n, kvh, d, generator=g).to(torch.bfloat16).to(cuda_device) for _ in range(2))
    q = torch.randn(len(ctx), h, d, generator=g).to(torch.bfloat16).to(cuda_device)
    lens_t = torch.tensor(ctx, dtype=torch.int32, device=cuda_device)
    scale = d**-0.5
    kc64, vc64, bt64 = _paged(k, v, ctx, 64, g, cuda_device)
    ours = m.paged_decode_attention(q, kc64, vc64, bt64, lens_t, scale)
    kc256, vc256, bt256 = _paged(k, v, ctx, 256, g, cuda_device, fill=0.0)
    theirs = _try(f"decode_block256_h{h}", lambda: with_kvcache(
        q.unsqueeze(1), kc256, vc256, cache_seqlens=lens_t, block_table=bt256, softmax_scale=scale, causal=True))
    if theirs is None:
        pytest.skip("flash_attn_with_kvcache does not run on this GPU (recorded)")
    diff = (ours.float() - theirs.squeeze(1).float()).abs()
    _record(f"decode_h{h}_kvh{kvh}", {"max_abs": diff.max().item(), "mean_abs": diff.mean().item(),
                                     "kernel": m.ops.last_kernel()})
    assert diff.max().item() <= 2e-2, (diff.max().item(), diff.mean().item())
