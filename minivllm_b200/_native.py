"""ctypes binding of libmvb200.so (the C ABI declared in include/mvb200.h).

The library is built in-tree by ``minivllm_b200/build.py`` (``__graft_entry__.build()``).
There is no fallback: if the library is missing or a launcher reports an error the
caller gets an exception.
"""

from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int32, c_int64, c_size_t, c_uint32, c_void_p


class DecodeStore(ctypes.Structure):
    """``mvb200_decode_store`` (include/mvb200.h): the step's new tokens for the fused store of ``mvb200_paged_decode_ex``."""

    _fields_ = [("key", c_void_p), ("value", c_void_p), ("key_row_stride", c_int64), ("value_row_stride", c_int64),
                ("slot_mapping", c_void_p)]

_HERE = os.path.dirname(os.path.abspath(__file__))
# MVB200_LIB points debugging tools at an instrumented build of the same sources (tools/prefill_trace.py)
LIB_PATH = os.environ.get("MVB200_LIB") or os.path.join(_HERE, "libmvb200.so")

_lib = None

# name -> (restype, argtypes); mirrors include/mvb200.h one to one
SIGNATURES = {
    "mvb200_version": (c_int32, []),
    "mvb200_last_error": (c_char_p, []),
    "mvb200_last_kernel": (c_char_p, []),
    "mvb200_kernels_seen": (c_int32, [c_char_p, c_int32]),
    "mvb200_store_kvcache": (
        c_int32,
        [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p,
         c_int32, c_int32, c_int64, c_int32, c_void_p],
    ),
    "mvb200_decode_workspace_bytes": (c_size_t, [c_int32, c_int32, c_int32, c_int32]),
    "mvb200_decode_suggest_splits": (c_int32, [c_int32, c_int32, c_int32, c_int32]),
    "mvb200_decode_workspace_init": (c_int32, [c_void_p, c_size_t, c_void_p]),
    "mvb200_paged_decode": (
        c_int32,
        [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p,
         c_void_p, c_int64, c_void_p, c_size_t,
         c_int32, c_int32, c_int32, c_int32, c_int32, c_int64, c_int32,
         c_int32, c_float, c_int32, c_void_p],
    ),
    "mvb200_paged_decode_ex": (
        c_int32,
        [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p,
         c_void_p, c_int64, c_void_p, c_size_t,
         c_int32, c_int32, c_int32, c_int32, c_int32, c_int64, c_int32,
         c_int32, c_float, c_int32, c_uint32, c_void_p, c_void_p],
    ),
    "mvb200_prefill_varlen": (
        c_int32,
        [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int64, c_int64,
         c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p,
         c_void_p, c_int64,
         c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32,
         c_int32, c_int64, c_int32, c_float, c_int32, c_void_p],
    ),
    "mvb200_qk_norm_rope_store": (
        c_int32,
        [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64,
         c_void_p, c_int64, c_void_p, c_int64,
         c_void_p, c_void_p, c_void_p, c_int64,
         c_void_p, c_void_p, c_float,
         c_void_p, c_void_p, c_void_p, c_int64,
         c_int32, c_int32, c_int32, c_int32, c_int32, c_void_p],
    ),
    "mvb200_decode_advance": (
        c_int32,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
         c_int32, c_int32, c_int32, c_int64, c_void_p],
    ),
    "mvb200_greedy_argmax_workspace_bytes": (c_size_t, [c_int32, c_int32]),
    "mvb200_greedy_argmax": (
        c_int32,
        [c_void_p, c_int64, c_int32, c_int32, c_int32, c_void_p, c_size_t,
         c_void_p, c_void_p, c_int64, c_void_p],
    ),
    "mvb200_greedy_argmax_static": (
        c_int32,
        [c_void_p, c_int64, c_int32, c_int32, c_int32, c_void_p, c_void_p, c_int64, c_void_p],
    ),
    "mvb200_peer_flag_bytes": (c_int32, []),
    "mvb200_peer_epoch_bytes": (c_int32, []),
    "mvb200_peer_max_ctas": (c_int32, []),
    "mvb200_peer_alloc": (c_int32, [c_size_t, c_void_p]),
    "mvb200_peer_free": (c_int32, [c_void_p]),
    "mvb200_peer_export": (c_int32, [c_void_p, c_void_p]),
    "mvb200_peer_open": (c_int32, [c_void_p, c_void_p]),
    "mvb200_peer_close": (c_int32, [c_void_p]),
    "mvb200_peer_allreduce_add_rms_norm": (
        c_int32,
        [c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32,
         c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int64,
         c_int32, c_int32, c_float, c_int32, c_int32, c_int32, c_int32, c_void_p],
    ),
    "mvb200_nvls_allreduce_add_rms_norm": (
        c_int32,
        [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int32, c_int32,
         c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int64,
         c_int32, c_int32, c_float, c_int32, c_int32, c_int32, c_void_p],
    ),
}


class NativeLibraryError(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    """Load libmvb200.so once; raise loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NativeLibraryError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). The mvb200 ops have no CPU or PyTorch fallback."
            )
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the .so is stale
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def last_error() -> str:
    return (lib().mvb200_last_error() or b"").decode()


def last_kernel() -> str:
    return (lib().mvb200_last_kernel() or b"").decode()


def kernels_seen() -> list[str]:
    buf = ctypes.create_string_buffer(2048)
    lib().mvb200_kernels_seen(buf, len(buf))
    return [k for k in buf.value.decode().split(",") if k]


def check(rc: int, what: str) -> None:
    if rc == 0:
        return
    msg = f"{what} failed (code {rc}): {last_error()}"
    if rc == 1:
        raise ValueError(msg)
    raise RuntimeError(msg)
