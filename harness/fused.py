"""ctypes binding of harness/libharness_ops.so (fused add+RMSNorm and SiLU*mul for the harness decoder).

Harness support only — outside the attention hot path of SURVEY §8; see harness/csrc/fused_elementwise.cu.
No CPU fallback here: callers use these on CUDA tensors and the plain-torch formulas on CPU.
"""

from __future__ import annotations

import ctypes
import os
import subprocess

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "fused_elementwise.cu")
LIB_PATH = os.path.join(HERE, "libharness_ops.so")
_DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}
_lib = None


def build(force: bool = False) -> str:
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(SRC):
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
                        "-Xcompiler", "-fPIC", "-shared", "-cudart", "shared", "-o", LIB_PATH, SRC], check=True)
    return LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run __graft_entry__.build()")
        h = ctypes.CDLL(LIB_PATH)
        i64, i32, vp = ctypes.c_int64, ctypes.c_int32, ctypes.c_void_p
        h.harness_add_rms_norm.restype = i32
        h.harness_add_rms_norm.argtypes = [vp, i64, vp, i64, vp, vp, i64, i32, i32, ctypes.c_float, i32, vp]
        h.harness_silu_mul.restype = i32
        h.harness_silu_mul.argtypes = [vp, i64, vp, i64, i32, i32, i32, vp]
        _lib = h
    return _lib


def add_rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: torch.Tensor | None = None):
    """``RMSNorm.forward(x, residual)`` (layernorm.py:107-131): returns ``(y, residual)``; ``residual`` is
    updated in place with ``x + residual`` when given, and is ``x`` itself otherwise (first layer)."""
    n, hidden = x.shape
    assert x.stride(1) == 1 and weight.dtype == x.dtype
    y = torch.empty((n, hidden), dtype=x.dtype, device=x.device)
    if residual is not None:
        assert residual.stride(1) == 1 and residual.dtype == x.dtype
    rc = lib().harness_add_rms_norm(
        x.data_ptr(), x.stride(0), residual.data_ptr() if residual is not None else None,
        residual.stride(0) if residual is not None else 0, weight.data_ptr(), y.data_ptr(), y.stride(0),
        n, hidden, float(eps), _DT[x.dtype], torch.cuda.current_stream().cuda_stream)
    if rc:
        raise RuntimeError(f"harness_add_rms_norm failed ({rc}) for shape {tuple(x.shape)} {x.dtype}")
    return y, (residual if residual is not None else x)


def silu_mul(gate_up: torch.Tensor) -> torch.Tensor:
    """``SiluAndMul.forward`` (activation.py:70-95) on ``[N, 2*I]``."""
    n, two_i = gate_up.shape
    assert gate_up.stride(1) == 1 and two_i % 2 == 0
    out = torch.empty((n, two_i // 2), dtype=gate_up.dtype, device=gate_up.device)
    rc = lib().harness_silu_mul(gate_up.data_ptr(), gate_up.stride(0), out.data_ptr(), out.stride(0), n, two_i // 2,
                                _DT[gate_up.dtype], torch.cuda.current_stream().cuda_stream)
    if rc:
        raise RuntimeError(f"harness_silu_mul failed ({rc}) for shape {tuple(gate_up.shape)} {gate_up.dtype}")
    return out
