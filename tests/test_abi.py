"""The C-ABI library: it loads on a CPU-only box and exports what include/mvb200.h declares."""

import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def native():
    import __graft_entry__ as entry

    entry.build()
    import minivllm_b200

    return minivllm_b200._native


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "mvb200.h")).read()
    return sorted(set(re.findall(r"MVB200_API[^;(]*?\b(mvb200_\w+)\s*\(", text)))


def test_header_declares_the_hot_path():
    syms = _declared_symbols()
    for want in ("mvb200_store_kvcache", "mvb200_paged_decode", "mvb200_paged_decode_ex", "mvb200_prefill_varlen",
                 "mvb200_decode_workspace_bytes", "mvb200_decode_suggest_splits",
                 "mvb200_version", "mvb200_last_error", "mvb200_last_kernel"):
        assert want in syms


def test_library_exports_every_declared_symbol(native):
    handle = ctypes.CDLL(native.LIB_PATH)
    for sym in _declared_symbols():
        assert hasattr(handle, sym), f"{sym} declared in mvb200.h but not exported"
    assert sorted(native.SIGNATURES) == _declared_symbols()
    assert native.lib().mvb200_version() == 130


def test_host_side_helpers_need_no_gpu(native):
    lib = native.lib()
    # workspace: B*H*S*(D+1) floats of partials when S > 1, plus the fixed tail (65536 tickets + 16 counters)
    tail = 65536 * 4 + 64
    assert lib.mvb200_decode_workspace_bytes(256, 16, 128, 1) == tail
    assert lib.mvb200_decode_workspace_bytes(256, 16, 128, 4) == 256 * 16 * 4 * 129 * 4 + tail
    assert lib.mvb200_decode_workspace_init(None, 0, None) == 1  # EINVAL without touching a GPU
    # plenty of (sequence, kv head) units -> no split; one long sequence -> many splits
    assert lib.mvb200_decode_suggest_splits(256, 8, 2048, 148) == 1
    s = lib.mvb200_decode_suggest_splits(1, 4, 32768, 148)
    assert 32 <= s <= 128
    assert lib.mvb200_decode_suggest_splits(1, 1, 64, 148) == 1
    # never more units than resident CTA slots (2 per SM): one wave, the fused reduction waits for nobody queued
    for batch, kvh, ctx in ((8, 8, 2304), (1, 4, 32768), (16, 2, 4096), (3, 1, 100000), (100, 1, 8192)):
        s = lib.mvb200_decode_suggest_splits(batch, kvh, ctx, 148)
        assert 1 <= s <= 128 and batch * kvh * s <= 2 * 148, (batch, kvh, ctx, s)
    assert lib.mvb200_decode_suggest_splits(8, 8, 2304, 148) == 4
    assert lib.mvb200_decode_suggest_splits(512, 1, 1024, 148) == 1  # splitting a multi-wave batch costs more than its tail


def test_argument_errors_are_reported_without_a_gpu(native):
    lib = native.lib()
    rc = lib.mvb200_store_kvcache(None, None, 8, 8, None, None, None, 4, 8, 16, 0, None)
    assert rc == 1 and "null" in native.last_error()
    rc = lib.mvb200_store_kvcache(None, None, 8, 8, None, None, None, 4, 8, 16, 9, None)
    assert rc == 1
    rc = lib.mvb200_paged_decode(None, 0, None, None, None, 0, None, None, 0, None, 0,
                                 2, 7, 2, 128, 64, 10, 4, 1, 1.0, 2, None)
    assert rc == 1 and "multiple" in native.last_error()
    # the _ex entry with a fused store whose pointers are missing
    store = native.DecodeStore(None, None, 0, 0, None)
    rc = lib.mvb200_paged_decode_ex(ctypes.c_void_p(16), 256, ctypes.c_void_p(16), ctypes.c_void_p(16), ctypes.c_void_p(16), 4,
                                    ctypes.c_void_p(16), ctypes.c_void_p(16), 256, None, 0, 2, 2, 2, 128, 64, 10, 4, 1, 1.0, 2, 0,
                                    ctypes.byref(store), None)
    assert rc == 1 and "fused store" in native.last_error()
    # NVLS all-reduce: world of one, null buffers, unaligned slot
    rc = lib.mvb200_nvls_allreduce_add_rms_norm(None, None, 1024, None, None, 0, 1, None, 0, None, 0, None, None, 0,
                                                4, 256, 1e-6, 2, 0, 1, None)
    assert rc == 1 and "rank/world" in native.last_error()
    rc = lib.mvb200_nvls_allreduce_add_rms_norm(None, None, 1 << 20, None, None, 0, 2, None, 0, None, 0, None, None, 0,
                                                4, 256, 1e-6, 2, 0, 1, None)
    assert rc == 1 and "null" in native.last_error()
    with pytest.raises(ValueError):
        native.check(1, "x")
    with pytest.raises(RuntimeError):
        native.check(3, "x")


def test_ops_refuse_cpu_tensors(native):
    import torch

    import minivllm_b200 as m

    z = torch.zeros(1, 1, 8)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m.store_kvcache(z, z, torch.zeros(4, 8), torch.zeros(4, 8), torch.zeros(1, dtype=torch.int32))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m.paged_decode_attention(z, torch.zeros(2, 4, 1, 8), torch.zeros(2, 4, 1, 8),
                                 torch.zeros(1, 1, dtype=torch.int32), torch.ones(1, dtype=torch.int32), 1.0)
    with pytest.raises(ValueError, match="empty"):
        m.paged_decode_attention(z, torch.tensor([]), torch.tensor([]),
                                 torch.zeros(1, 1, dtype=torch.int32), torch.ones(1, dtype=torch.int32), 1.0)


def test_custom_torch_ops_are_registered():
    """the operators are also exposed as torch.ops.mvb200.* with mutation annotated."""
    import torch

    import minivllm_b200 as m

    for name in m.torch_ops.REGISTERED:
        assert hasattr(torch.ops.mvb200, name), name
    schema = str(torch.ops.mvb200.store_kvcache.default._schema)
    assert "Tensor(a2!) k_cache" in schema and "Tensor(a3!) v_cache" in schema, schema
    # fake (shape-only) kernels run without a GPU
    from torch._subclasses.fake_tensor import FakeTensorMode

    with FakeTensorMode():
        q = torch.empty(3, 16, 128, dtype=torch.bfloat16)
        kc = torch.empty(8, 64, 8, 128, dtype=torch.bfloat16)
        bt = torch.empty(3, 4, dtype=torch.int32)
        cl = torch.empty(3, dtype=torch.int32)
        out = torch.ops.mvb200.paged_decode(q, kc, kc, bt, cl, 0.1)
        assert out.shape == q.shape and out.dtype == q.dtype
        tok = torch.ops.mvb200.greedy_argmax(torch.empty(3, 100, dtype=torch.bfloat16))
        assert tok.shape == (3,) and tok.dtype == torch.int64
    # real tensors on CPU: the op refuses, there is no fallback
    import pytest

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        torch.ops.mvb200.greedy_argmax(torch.zeros(2, 10))
