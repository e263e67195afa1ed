"""``mvb200_greedy_argmax_static``: the greedy token pick for callers that cannot allocate device memory.

Who needs it: a host that binds libmvb200.so from a language without a device allocator (cgo, JNI, a bare
ctypes shim), or a serving process that gave every spare byte of HBM to the KV cache at start-up and whose
step loop must not run ``cudaMalloc`` (it stalls the whole device).  Such a caller has no workspace to pass to
``mvb200_greedy_argmax``; this entry point keeps the per-chunk partial winners in an array that is part of
the library image instead (``MVB200_GREEDY_ARGMAX_STATIC_BYTES``, include/mvb200.h).

The array is shared by every call on the device and keeps the previous call's contents, so the test makes two
calls in a row with different shapes and data and checks each against ``mvb200_greedy_argmax`` (caller
workspace) on the same logits: indices and winning values must be identical — integer work, bit-exact, the bar
tests/test_decode_meta.py uses for the existing entry point.

This file sorts last on purpose: the driver runs ``pytest -x``, and a failure here must not hide other tests.
It passes on its own in a fresh process (the first call on a device resolves the array's address) and after
the rest of the suite.
"""

import ctypes

import pytest
import torch

CODES = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _with_workspace(native, logits, index_offset):
    """The existing entry point, called through the C ABI with a caller-provided workspace."""
    lib = native.lib()
    rows, vocab = logits.shape
    nbytes = int(lib.mvb200_greedy_argmax_workspace_bytes(rows, vocab))
    ws = torch.empty(max(nbytes // 8, 1), dtype=torch.int64, device=logits.device)
    idx = torch.full((rows,), -7, dtype=torch.int64, device=logits.device)
    val = torch.full((rows,), -7.0, dtype=torch.float32, device=logits.device)
    rc = lib.mvb200_greedy_argmax(logits.data_ptr(), logits.stride(0), rows, vocab, CODES[logits.dtype],
                                  ws.data_ptr(), ws.numel() * 8, idx.data_ptr(), val.data_ptr(), index_offset,
                                  _stream())
    assert rc == 0, native.last_error()
    return idx, val


def _static(native, logits, idx, val, index_offset):
    """The entry point under test: inputs and outputs are the caller's, there is no workspace argument."""
    rows, vocab = logits.shape
    rc = native.lib().mvb200_greedy_argmax_static(logits.data_ptr(), logits.stride(0), rows, vocab,
                                                  CODES[logits.dtype], idx.data_ptr(), val.data_ptr(), index_offset,
                                                  _stream())
    assert rc == 0, native.last_error()


def _logits(rows, vocab, dtype, seed, device):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(rows, vocab, generator=g).to(dtype)  # bf16 rounding gives plenty of exact ties
    x[0, 17] = x[0, vocab - 3] = 30.0                     # a tie across two chunks: the lower index wins
    if rows > 1:
        x[1, vocab // 2] = float("nan")                   # NaN is the maximum, as in torch.argmax
    if rows > 2:
        x[2] = -float("inf")                              # nothing but -inf: index 0
    return x.to(device)


@pytest.mark.gpu
def test_static_scratch_argmax_equals_the_workspace_entry_point_twice_in_a_row(cuda_device):
    import minivllm_b200

    native = minivllm_b200._native
    # Call 1: 8 rows x Qwen3 vocabulary in bf16 -> 19 chunks per row, 152 partial keys in the array.
    # Call 2: fewer rows, another dtype and vocabulary -> 3 rows x 10 chunks; it finds call 1's keys in the
    # array and must not be influenced by them.  Call 3 repeats call 1's shape on new data.
    cases = [(8, 151936, torch.bfloat16, 11, 1000), (3, 75968, torch.float32, 12, 0),
             (8, 151936, torch.bfloat16, 13, 5)]
    prepared = []
    for rows, vocab, dtype, seed, offset in cases:
        x = _logits(rows, vocab, dtype, seed, cuda_device)
        want_idx, want_val = _with_workspace(native, x, offset)
        idx = torch.full((rows,), -9, dtype=torch.int64, device=cuda_device)
        val = torch.full((rows,), -9.0, dtype=torch.float32, device=cuda_device)
        prepared.append((x, offset, idx, val, want_idx, want_val))
    torch.cuda.synchronize()

    # every buffer exists by now; the calls below must leave the device's free memory where it is
    x, offset, idx, val, _, _ = prepared[0]
    _static(native, x, idx, val, offset)  # first call on this device: also resolves the array's address
    torch.cuda.synchronize()
    assert native.last_kernel() == "greedy_argmax_static"
    free_before, _ = torch.cuda.mem_get_info(cuda_device)
    for x, offset, idx, val, _, _ in prepared[1:]:
        _static(native, x, idx, val, offset)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info(cuda_device)
    assert free_after == free_before, "mvb200_greedy_argmax_static changed the device's free memory"

    for x, offset, idx, val, want_idx, want_val in prepared:
        assert torch.equal(idx, want_idx)
        first_max = torch.argmax(torch.nan_to_num(x.float().cpu(), nan=float("inf")), dim=-1)  # CPU: first of equals
        assert torch.equal(idx.cpu() - offset, first_max)
        assert torch.equal(torch.nan_to_num(val, nan=-1.0), torch.nan_to_num(want_val, nan=-1.0))

    # argument errors come back as codes with a message, before anything is launched
    lib = native.lib()
    assert lib.mvb200_greedy_argmax_static(None, 10, 2, 10, 2, None, None, 0, None) == 1
    assert "null" in native.last_error()
    assert lib.mvb200_greedy_argmax_static(x.data_ptr(), 10, 2, 10, 9, idx.data_ptr(), None, 0, None) == 1
    assert "dtype" in native.last_error()
