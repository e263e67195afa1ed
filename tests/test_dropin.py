"""Drop-in surface of ``minivllm_b200.Attention`` checked against the UNMODIFIED reference.

Runs only where ``/root/reference`` exists (the build container; the GPU box does
not have it) and in a child process, because importing the reference needs a
``colorama`` stand-in and a hidden ``flash_attn`` that must not leak into the
test session.  What is pinned:

* ``install_into_reference()`` rebinds ``Attention`` in every reference module
  that imported it by name (layers/__init__.py:22, qwen_base.py, opt.py, gpt2.py)
* constructor and ``forward`` signatures equal the reference's (attention.py:143-150,196-198)
* ``scale == 0 -> 1/sqrt(head_dim)`` (attention.py:154)
* the executor's attribute discovery finds the layer (inference_executor.py:325-331)
  and a reference ``QwenAttention`` built after the swap holds the new layer with
  no extra parameters or buffers in its ``state_dict``
* the reference's ``set_context`` is what the new ``forward`` reads
* there is no CPU fallback: CPU tensors raise instead of computing
"""

import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE = os.environ.get("MINIVLLM_REFERENCE", "/root/reference")

CHILD = textwrap.dedent(
    r"""
    import inspect, math, os, sys, types
    os.environ["MINIVLLM_DEVICE"] = "cpu"
    os.environ["CUDA_VISIBLE_DEVICES"] = ""
    shim = types.ModuleType("colorama")
    class _Blank:
        def __getattr__(self, _n):
            return ""
    shim.Fore = shim.Style = shim.Back = _Blank()
    shim.init = lambda *a, **k: None
    sys.modules.setdefault("colorama", shim)
    sys.modules["flash_attn"] = None
    sys.path.insert(0, sys.argv[1])   # reference
    sys.path.insert(0, sys.argv[2])   # this repo

    import torch
    import minivllm.models.layers as ref_layers
    from minivllm.models.layers.attention import Attention as RefAttention
    import minivllm_b200
    from minivllm_b200 import Attention

    # -- signatures -------------------------------------------------------------------------
    def params(fn):
        return [(p.name, p.default) for p in inspect.signature(fn).parameters.values()]
    assert params(Attention.__init__) == params(RefAttention.__init__), (
        params(Attention.__init__), params(RefAttention.__init__))
    assert [n for n, _ in params(Attention.forward)] == [n for n, _ in params(RefAttention.forward)]

    # -- rebinding --------------------------------------------------------------------------
    minivllm_b200.install_into_reference()
    import importlib
    assert ref_layers.Attention is Attention
    seen = 0
    for name in ("minivllm.models.qwen_base", "minivllm.models.opt", "minivllm.models.gpt2"):
        try:
            mod = importlib.import_module(name)
        except Exception:
            continue
        if hasattr(mod, "Attention"):
            assert mod.Attention is Attention, name
            seen += 1
    assert seen >= 1, "no reference model module imported Attention by name"

    # -- layer surface ----------------------------------------------------------------------
    ref = RefAttention(8, 64, 0, 2)
    new = Attention(8, 64, 0, 2)
    assert math.isclose(new.scale, 1.0 / math.sqrt(64)) and math.isclose(ref.scale, new.scale)
    for attr in ("num_heads", "head_dim", "num_kv_heads", "k_cache", "v_cache", "_cache_initialized"):
        assert hasattr(new, attr), attr
        assert hasattr(ref, attr), attr
    assert new.k_cache.numel() == 0 and new.v_cache.numel() == 0
    assert list(new.state_dict().keys()) == list(ref.state_dict().keys()) == []
    assert list(new.parameters()) == []

    # executor-style discovery (inference_executor.py:325-331)
    holder = torch.nn.Sequential(torch.nn.Linear(4, 4), new)
    found = [m for m in holder.modules() if hasattr(m, "k_cache") and hasattr(m, "v_cache")]
    assert found == [new]

    # -- the reference's ContextVar feeds the new forward --------------------------------------
    from minivllm.utils.context import set_context, reset_context, get_context
    from minivllm_b200 import attention as new_attention
    assert new_attention._get_context is get_context

    # -- no CPU fallback --------------------------------------------------------------------
    q = torch.zeros(3, 8, 64); k = torch.zeros(3, 2, 64); v = torch.zeros(3, 2, 64)
    new.k_cache = torch.zeros(2, 64, 2, 64); new.v_cache = torch.zeros(2, 64, 2, 64)
    new._cache_initialized = True
    # our set_context takes the reference's parameters, in order, plus optional extras (context.py:88-97)
    from minivllm_b200 import context as new_context
    ref_params = params(set_context)
    assert params(new_context.set_context)[:len(ref_params)] == ref_params
    assert all(d is None for _, d in params(new_context.set_context)[len(ref_params):])
    import dataclasses
    from minivllm.utils.context import Context as RefContext
    ref_fields = [(f.name, f.default) for f in dataclasses.fields(RefContext)]
    new_fields = [(f.name, f.default) for f in dataclasses.fields(new_context.Context)]
    assert new_fields[:len(ref_fields)] == ref_fields, (new_fields, ref_fields)
    assert all(d is None for _, d in new_fields[len(ref_fields):])
    set_context(True, max_seqlen_q=3, max_seqlen_k=3,
                cum_seqlens_q=torch.tensor([0, 3], dtype=torch.int32),
                cum_seqlens_k=torch.tensor([0, 3], dtype=torch.int32),
                slot_mapping=torch.tensor([0, 1, 2], dtype=torch.int32))
    try:
        new(q, k, v)
    except (RuntimeError, ValueError, OSError) as exc:
        msg = str(exc)
    else:
        raise AssertionError("CPU tensors were accepted: there must be no CPU fallback")
    finally:
        reset_context()

    # -- a reference model block built after the swap holds the new layer ----------------------
    try:
        import torch.distributed as dist
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29631")
            dist.init_process_group("gloo", rank=0, world_size=1)
        from minivllm.models import qwen_base
        sig = inspect.signature(qwen_base.QwenAttention.__init__).parameters
        kw = dict(hidden_size=256, num_heads=4, num_kv_heads=2)
        for name, val in (("max_position", 128), ("head_dim", 64), ("qkv_bias", False),
                          ("rms_norm_eps", 1e-6), ("rope_theta", 10000.0)):
            if name in sig:
                kw[name] = val
        block = qwen_base.QwenAttention(**kw)
        assert isinstance(block.attn, Attention), type(block.attn)
        assert not any(k.startswith("attn.") for k in block.state_dict().keys())
        print("QWEN_BLOCK_OK")
    finally:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    print("DROPIN_OK")
    """
)


@pytest.mark.skipif(not os.path.isdir(os.path.join(REFERENCE, "minivllm")),
                    reason="the reference tree is only present in the build container")
def test_dropin_surface_against_reference(tmp_path):
    script = tmp_path / "child.py"
    script.write_text(CHILD)
    env = dict(os.environ)
    env.pop("PYTHONPATH", None)
    proc = subprocess.run([sys.executable, str(script), REFERENCE, ROOT], capture_output=True,
                          text=True, timeout=300, env=env)
    assert proc.returncode == 0, proc.stdout[-2000:] + "\n" + proc.stderr[-4000:]
    assert "QWEN_BLOCK_OK" in proc.stdout
    assert "DROPIN_OK" in proc.stdout


def test_debug_check_of_the_fused_store_slots():
    """``ops._check_decode_slots`` (MVB200_DEBUG_CHECKS=1): slot b must be the slot of position context_lens[b]-1
    (inference_executor.py:568-572); padding rows are exempt.  Pure tensor logic, runs on CPU."""
    import torch

    from minivllm_b200 import ops

    bt = torch.tensor([[3, 5, -1], [7, -1, -1], [2, 4, 6]], dtype=torch.int32)
    cl = torch.tensor([70, 1, 0], dtype=torch.int32)
    good = torch.tensor([5 * 64 + 5, 7 * 64, -1], dtype=torch.int32)
    ops._check_decode_slots(good, bt, cl, 64)
    with pytest.raises(ValueError, match=r"slot_mapping\[0\]"):
        ops._check_decode_slots(torch.tensor([3 * 64 + 5, 7 * 64, -1], dtype=torch.int32), bt, cl, 64)
    with pytest.raises(ValueError, match=r"slot_mapping\[1\]"):
        ops._check_decode_slots(torch.tensor([5 * 64 + 5, 7 * 64 + 1, -1], dtype=torch.int32), bt, cl, 64)
