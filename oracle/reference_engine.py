"""Drive the UNMODIFIED reference engine (``minivllm.LLM``) — test / bench infrastructure, not product.

Only ``tests/``, ``bench.py`` (its reference and drop-in legs), ``tools/`` and
``__graft_entry__`` import this module; nothing under ``minivllm_b200/`` does.

Where the reference comes from: ``/root/reference`` in the build container, else the
copy that ``oracle/make_ref.py`` placed under the git-ignored ``oracle/_ref/`` (the
copy is what travels to the GPU box; reference sources are never committed).

The three harness-level patches of SURVEY.md 8(c) (no reference file is edited):

* ``minivllm.models.manager.load_model`` (utils/loader.py:90-239) -> seeded random
  init of every parameter, keyed by the parameter name (``seeded_tensor``);
* ``AutoTokenizer.from_pretrained`` (manager.py:94-106) -> a stub tokenizer
  (``eos_token_id``, ``pad_token``, ``batch_decode``; llm_engine.py:107-108,359-361);
* a synthetic model directory holding only ``config.json`` with the public
  HuggingFace dimensions of the named architecture.

Prompts are token-id lists (llm_engine.py:191-194).
"""

from __future__ import annotations

import contextlib
import json
import os
import sys
import tempfile
import types
import zlib

HERE = os.path.dirname(os.path.abspath(__file__))
REF_COPY = os.path.join(HERE, "_ref")

# Public HF config values.  ``tiny-*`` are the 2-layer parity presets of round 1.
PRESETS: dict[str, dict] = {
    "qwen3-0.6b": dict(architectures=["Qwen3ForCausalLM"], model_type="qwen3", vocab_size=151936, hidden_size=1024,
                       num_hidden_layers=28, num_attention_heads=16, num_key_value_heads=8, head_dim=128,
                       intermediate_size=3072, rms_norm_eps=1e-6, rope_theta=1e6, max_position_embeddings=40960,
                       tie_word_embeddings=True, hidden_act="silu", attention_bias=False),
    "qwen3-4b": dict(architectures=["Qwen3ForCausalLM"], model_type="qwen3", vocab_size=151936, hidden_size=2560,
                     num_hidden_layers=36, num_attention_heads=32, num_key_value_heads=8, head_dim=128,
                     intermediate_size=9728, rms_norm_eps=1e-6, rope_theta=1e6, max_position_embeddings=40960,
                     tie_word_embeddings=True, hidden_act="silu", attention_bias=False),
    "qwen2-7b": dict(architectures=["Qwen2ForCausalLM"], model_type="qwen2", vocab_size=152064, hidden_size=3584,
                     num_hidden_layers=28, num_attention_heads=28, num_key_value_heads=4,
                     intermediate_size=18944, rms_norm_eps=1e-6, rope_theta=1e6, max_position_embeddings=131072,
                     tie_word_embeddings=False, hidden_act="silu"),
    "opt-125m": dict(architectures=["OPTForCausalLM"], model_type="opt", vocab_size=50272, hidden_size=768,
                     num_hidden_layers=12, num_attention_heads=12, ffn_dim=3072, max_position_embeddings=2048,
                     do_layer_norm_before=True, word_embed_proj_dim=768, enable_bias=True,
                     layer_norm_elementwise_affine=True, activation_function="relu", tie_word_embeddings=True,
                     pad_token_id=1, bos_token_id=2, eos_token_id=2),
    "tiny-qwen3": dict(architectures=["Qwen3ForCausalLM"], model_type="qwen3", vocab_size=512, hidden_size=256,
                       num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2, head_dim=128,
                       intermediate_size=512, rms_norm_eps=1e-6, rope_theta=1e6, max_position_embeddings=2048,
                       tie_word_embeddings=False, hidden_act="silu", attention_bias=False),
    "tiny-qwen2": dict(architectures=["Qwen2ForCausalLM"], model_type="qwen2", vocab_size=512, hidden_size=256,
                       num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2, head_dim=64,
                       intermediate_size=512, rms_norm_eps=1e-6, rope_theta=1e6, max_position_embeddings=2048,
                       tie_word_embeddings=False, hidden_act="silu", attention_bias=True),
}


def locate_reference() -> str | None:
    """Directory that contains the reference's ``minivllm`` package, or ``None``."""
    for cand in (os.environ.get("MINIVLLM_REFERENCE"), "/root/reference", REF_COPY):
        if cand and os.path.isdir(os.path.join(cand, "minivllm")):
            return cand
    return None


def import_reference(device: str = "cpu", *, hide_flash_attn: bool = True, isolate_cuda: bool = True):
    """Import the reference package for ``device`` ('cpu' | 'cuda') and return the module.

    ``colorama`` is the reference's only dependency missing from the image: a blank stand-in is
    registered.  ``hide_flash_attn`` forces the reference's own eager attention branch
    (attention.py:80-87,450-453); it is irrelevant once ``install_into_reference()`` has swapped the layer.
    ``MINIVLLM_DEVICE`` is read by the reference at call time (utils/device.py:74-78), so one process may
    run a CPU engine and a CUDA engine one after the other.
    """
    root = locate_reference()
    if root is None:
        raise RuntimeError("reference not found: neither /root/reference nor oracle/_ref/ (run oracle/make_ref.py)")
    os.environ["MINIVLLM_DEVICE"] = device
    if device == "cpu" and isolate_cuda:  # bench.py times the CPU path inside a process that also drives the GPU
        os.environ["CUDA_VISIBLE_DEVICES"] = ""
    if "colorama" not in sys.modules:
        shim = types.ModuleType("colorama")

        class _Blank:
            def __getattr__(self, _name):
                return ""

        shim.Fore = shim.Style = shim.Back = _Blank()
        shim.init = lambda *a, **k: None
        sys.modules["colorama"] = shim
    if hide_flash_attn:
        sys.modules["flash_attn"] = None
    if root not in sys.path:
        sys.path.insert(0, root)
    import minivllm  # noqa: PLC0415

    return minivllm


def seeded_tensor(name: str, shape, seed: int = 1234):
    """Deterministic CPU init keyed by the parameter name (identical here, in the golden generator and on
    the GPU box): N(0, 0.02) for matrices and biases, 1 + 0.05 N(0, 1) for norm weights."""
    import torch  # noqa: PLC0415

    g = torch.Generator().manual_seed(seed + zlib.crc32(name.encode()))
    if name.endswith("norm.weight"):
        return 1.0 + 0.05 * torch.randn(shape, generator=g)
    return 0.02 * torch.randn(shape, generator=g)


def seeded_load(model, path=None) -> None:
    """Stands in for ``load_model`` (utils/loader.py:90-239).

    Under tensor parallelism every rank draws the FULL tensor of each parameter (same name -> same values as
    at TP=1) and hands it to the reference's own ``weight_loader`` of that parameter (linear.py:146-270,
    embed_head.py:106-119), which keeps this rank's shard — so a TP=N engine holds exactly the TP=1 weights."""
    import torch  # noqa: PLC0415
    import torch.distributed as dist  # noqa: PLC0415

    tp = dist.get_world_size() if dist.is_initialized() else 1
    with torch.no_grad():
        if tp == 1:
            for name, param in model.named_parameters():
                param.copy_(seeded_tensor(name, tuple(param.shape)).to(param.dtype))
            return
        seen: set[int] = set()
        for mod_name, mod in model.named_modules():
            for pname, param in mod.named_parameters(recurse=False):
                if id(param) in seen:  # tied lm_head / embedding
                    continue
                seen.add(id(param))
                name = f"{mod_name}.{pname}" if mod_name else pname
                cls = type(mod).__name__
                shape = list(param.shape)
                dt = param.dtype
                if cls == "QKVParallelLinear":
                    q, kv = mod.num_heads * tp * mod.head_size, mod.num_kv_heads * tp * mod.head_size
                    shape[0] = q + 2 * kv
                    for sid, part in zip("qkv", seeded_tensor(name, shape).split([q, kv, kv], 0)):
                        param.weight_loader(param, part.to(dt), sid)
                elif cls == "MergedColumnParallelLinear":
                    shape[0] = sum(mod.output_sizes)
                    for i, part in enumerate(seeded_tensor(name, shape).split(list(mod.output_sizes), 0)):
                        param.weight_loader(param, part.to(dt), i)
                elif cls in ("VocabParallelEmbedding", "ParallelLMHead"):
                    shape[0] *= tp
                    param.weight_loader(param, seeded_tensor(name, shape).to(dt))
                elif getattr(mod, "tp_dim", None) is not None and param.dim() > mod.tp_dim:
                    shape[mod.tp_dim] *= tp
                    param.weight_loader(param, seeded_tensor(name, shape).to(dt))
                else:  # replicated: norms, row-parallel biases
                    param.copy_(seeded_tensor(name, tuple(param.shape)).to(dt))


def fast_device_load(device: str, dtype_name: str, seed: int = 1234):
    """Loader for throughput runs (no parity): move the freshly built model to ``device`` in its final
    dtype and draw the weights there — 4 B-parameter models initialise in a second instead of minutes."""

    def load(model, path=None) -> None:
        import torch  # noqa: PLC0415

        dt = {"float32": torch.float32, "bfloat16": torch.bfloat16, "float16": torch.float16}[dtype_name]
        model.to(device=device, dtype=dt)
        g = torch.Generator(device=device).manual_seed(seed)
        with torch.no_grad():
            for name, param in model.named_parameters():
                if name.endswith("norm.weight"):
                    param.fill_(1.0)
                else:
                    param.normal_(0.0, 0.02, generator=g)

    return load


class StubTokenizer:
    """manager.py:94-106, llm_engine.py:107-108,359-361."""

    eos_token_id = 1
    pad_token = "<pad>"
    pad_token_id = 0

    def batch_decode(self, ids, **kw):
        return ["" for _ in ids]

    def decode(self, ids, **kw):
        return ""

    def encode(self, text, **kw):
        return [10 + (ord(c) % 200) for c in text]


def write_model_dir(preset: str, dtype: str = "float32", **overrides) -> str:
    """Synthetic model directory: ``config.json`` only."""
    cfg = dict(PRESETS[preset])
    cfg["torch_dtype"] = dtype
    cfg.update(overrides)
    tmp = tempfile.mkdtemp(prefix=f"mvb200_{preset}_")
    with open(os.path.join(tmp, "config.json"), "w") as f:
        json.dump(cfg, f)
    return tmp


@contextlib.contextmanager
def engine_patches(loader=seeded_load):
    """Apply the loader and tokenizer patches while a reference ``LLM`` is constructed."""
    from unittest import mock  # noqa: PLC0415

    import minivllm.models.manager as manager  # noqa: PLC0415

    with mock.patch.object(manager, "load_model", loader), \
            mock.patch.object(manager.AutoTokenizer, "from_pretrained", lambda *a, **k: StubTokenizer()):
        yield


def make_llm(preset: str, *, dtype: str = "float32", loader=seeded_load, config_overrides: dict | None = None, **config_kw):
    """``minivllm.LLM(Config(...))`` on a synthetic ``preset`` model.  ``import_reference`` must have run."""
    from minivllm import LLM  # noqa: PLC0415
    from minivllm.config import Config  # noqa: PLC0415

    path = write_model_dir(preset, dtype, **(config_overrides or {}))
    config_kw.setdefault("dtype", dtype)
    with engine_patches(loader):
        return LLM(Config(model=path, **config_kw))


def greedy_params(max_tokens: int = 32):
    """The reference's greedy setting: ``temperature=0`` still samples through ``torch.multinomial`` on a
    one-hot distribution (sampling/functional.py:434-436)."""
    from minivllm import SamplingParams  # noqa: PLC0415

    return SamplingParams(temperature=0.0, max_tokens=max_tokens, ignore_eos=True)
