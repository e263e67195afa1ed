"""Rebind the three executor subsystems north_star says change, on the reference's own classes.

``install_into_reference(executor=True)`` ends here.  Nothing in the reference is edited on disk: the
methods below are assigned onto ``minivllm.engine.inference_executor.InferenceExecutor`` at run time and
delegate to a ``HotPathExecutor`` (``executor.py``) that shares the reference executor's model, KV cache
and ``Context`` accessors.

| reference method (inference_executor.py)            | replaced by                                   | why |
|---|---|---|
| ``__init__`` (:59-95)                                | same + ``num_kv_heads //= tp``                | cache sized with the global KV-head count (defect 2) |
| ``_prepare_prefill_input`` (:468-549)                | ``HotPathExecutor.prepare_prefill``           | O(N) Python ``extend(range())`` per step -> numpy |
| ``_prepare_decode_input`` + ``_prepare_block_tables_decode`` (:551-688) | ``HotPathExecutor.prepare_decode`` | per-sequence ``torch.tensor`` per step -> one vectorised fill, pinned staging |
| ``capture_device_graphs`` (:823-920)                 | ``HotPathExecutor.capture_graphs``            | ``outputs.copy_(model(...)[0])`` row-0 broadcast, fp32 outputs buffer, frozen split count + workspace |
| ``_execute_with_cuda_graph`` (:728-774)              | ``HotPathExecutor.replay_decode`` + ``compute_logits`` | same replay protocol on the buffers above |
| ``_sample_tokens`` (:776-806), all-greedy batches only | ``ops.greedy_argmax``                         | the full sampling pipeline + multinomial on a one-hot distribution |
| ``initialize`` (:184-196), ``execute_batch`` (:412-449), TP > 1 only | same + capture on every rank; workers skip logits/sampling | graphs captured on rank 0 only (model_runner.py:113-115); ``ParallelLMHead`` returns ``None`` off rank 0 |

``install_fused_qk()`` (``install_into_reference(fuse_qk=True)``) additionally rebinds ``QwenAttention.forward``
(models/qwen_base.py:146-190) to the one-kernel q/k RMSNorm + RoPE + cache store.

The scheduler, block manager, sequences, the rest of the models, the sampler and ``LLM.generate`` are untouched.
"""

from __future__ import annotations

import importlib
from typing import Any

import torch

from .executor import HotPathExecutor

_PATCHED = "_mvb200_patched"


def _hot(self) -> HotPathExecutor:
    """The ``HotPathExecutor`` twin of a reference ``InferenceExecutor`` (created on first use)."""
    hot = getattr(self, "_mvb200_hot", None)
    if hot is not None and hot.kv_cache is self.kv_cache:
        return hot
    ref_ctx = importlib.import_module("minivllm.utils.context")

    def set_context(**kw: Any) -> None:
        extra = {k: kw.pop(k) for k in ("decode_num_splits", "decode_workspace") if k in kw}
        ref_ctx.set_context(**kw)
        if extra:  # the reference's Context is a plain dataclass: optional fields ride on the instance
            ctx = ref_ctx.get_context()
            for k, v in extra.items():
                setattr(ctx, k, v)

    hf = self.config.hf_config
    tp = torch.distributed.get_world_size() if torch.distributed.is_initialized() else 1
    # the first call comes from the reference's warm-up, which runs under torch.inference_mode()
    # (inference_executor.py:396-404): persistent staging buffers must be ordinary tensors
    with torch.inference_mode(False):
        hot = _make_hot(self, hf, tp, set_context, ref_ctx)
    if self.kv_cache is not None:
        hot.kv_cache = self.kv_cache
        hot.num_blocks = self.kv_cache.size(2)
    self._mvb200_hot = hot
    return hot


def _make_hot(self, hf, tp, set_context, ref_ctx) -> HotPathExecutor:
    return HotPathExecutor(
        self.model,
        num_layers=self.num_layers,
        num_kv_heads=self.num_kv_heads,
        head_dim=self.head_dim,
        hidden_size=hf.hidden_size,
        dtype=self.dtype,
        device=self.device,
        max_num_seqs=self.config.max_num_seqs,
        max_model_len=self.config.max_model_len,
        block_size=self.block_size,
        num_heads=max(hf.num_attention_heads // tp, 1),
        set_context=set_context,
        reset_context=ref_ctx.reset_context,
        get_context=ref_ctx.get_context,
    )


def install_executor() -> None:
    mod = importlib.import_module("minivllm.engine.inference_executor")
    cls = mod.InferenceExecutor
    if getattr(cls, _PATCHED, False):
        return
    orig_init = cls.__init__

    def __init__(self, config, model) -> None:
        orig_init(self, config, model)
        tp = torch.distributed.get_world_size() if torch.distributed.is_initialized() else 1
        if tp > 1:
            if self.num_kv_heads % tp:
                raise ValueError(f"num_kv_heads ({self.num_kv_heads}) must be divisible by tp_size ({tp})")
            self.num_kv_heads //= tp  # this rank's share: what the sharded layers hand to Attention

    def _prepare_prefill_input(self, sequences):
        return _hot(self).prepare_prefill(sequences)

    def _prepare_decode_input(self, sequences):
        return _hot(self).prepare_decode(sequences)

    orig_initialize = cls.initialize
    orig_execute_batch = cls.execute_batch

    def initialize(self, max_num_batched_tokens: int, max_num_seqs: int) -> None:
        orig_initialize(self, max_num_batched_tokens, max_num_seqs)
        # graphs on EVERY rank: the reference captures on rank 0 only (model_runner.py:113-115), which cannot
        # work once the captured forward holds collectives
        if torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1:
            self.capture_device_graphs(max_num_seqs)

    def execute_batch(self, sequences, prefill: bool = True):
        tp_worker = torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1 \
            and torch.distributed.get_rank() != 0
        if not tp_worker or not sequences:
            return orig_execute_batch(self, sequences, prefill)
        # ParallelLMHead gathers the logits on rank 0 and returns None elsewhere (embed_head.py:217-229): workers
        # run the same forward (their half of every collective) and sample nothing
        ref_ctx = importlib.import_module("minivllm.utils.context")
        try:
            input_ids, positions = self._prepare_batch_input(sequences, prefill)
            self._execute_model(input_ids, positions, prefill)
            self.inference_count += 1
            return None, []
        finally:
            ref_ctx.reset_context()

    def capture_device_graphs(self, max_batch_size: int) -> None:
        if self.enforce_eager or self.device.type != "cuda" or self.graphs:
            return
        hot = _hot(self)
        with torch.inference_mode(False):
            hot.capture_graphs(hot_graph_sizes(max_batch_size))
        self.graphs, self.graph_bs, self.graph_vars, self.graph_pool = hot.graphs, hot.graph_bs, hot.graph_vars, hot.graph_pool

    def _execute_with_cuda_graph(self, input_ids, positions):
        out = _hot(self).replay_decode(input_ids, positions)
        if out is None:
            return self._execute_eager(input_ids, positions)
        with torch.inference_mode():
            return self.model.compute_logits(out)  # None on TP workers

    orig_sample = cls._sample_tokens

    def _sample_tokens(self, logits, sequences):
        """Greedy fast path: when every sequence asks for ``temperature == 0`` the
        reference still builds four per-step tensors and runs its whole temperature / top-k / top-p / min-p
        pipeline before ``torch.multinomial`` on a one-hot distribution (inference_executor.py:776-806,
        sampling/functional.py:434-436) — i.e. the argmax its sampler documents for this case
        (sampler.py:104-108).  One ``greedy_argmax`` launch instead; ties go to the lowest index."""
        if logits is None:  # TP worker: the logits live on rank 0 (embed_head.py:217-229)
            return []
        if logits.is_cuda and logits.dim() == 2 and all(seq.temperature == 0 for seq in sequences):
            from . import ops

            return ops.greedy_argmax(logits).tolist()
        return orig_sample(self, logits, sequences)

    cls._sample_tokens = _sample_tokens
    cls.__init__ = __init__
    cls.initialize = initialize
    cls.execute_batch = execute_batch
    cls._prepare_prefill_input = _prepare_prefill_input
    cls._prepare_decode_input = _prepare_decode_input
    cls.capture_device_graphs = capture_device_graphs
    cls._execute_with_cuda_graph = _execute_with_cuda_graph
    setattr(cls, _PATCHED, True)


def install_fused_qk() -> None:
    """Rebind ``QwenAttention.forward`` (models/qwen_base.py:146-190, an attention layer of ``minivllm/models``)
    so that the flat-token path runs ``q_norm / k_norm -> rotary_emb -> store_kv_cache`` as ONE kernel
    (``Attention.forward_fused`` -> ``mvb200_qk_norm_rope_store``) instead of the ~25
    elementwise launches of ``RMSNorm.forward`` x 2 (layernorm.py:81-93) and ``RotaryEmbedding.forward``
    (rotary_embedding.py:166-193).  Same projections, same module attributes, same ``state_dict``; anything the
    fused kernel does not cover (3-D batches, CPU tensors, a rotary module without cos/sin caches, an ``Attention``
    that is not ours) takes the reference's own forward."""
    mod = importlib.import_module("minivllm.models.qwen_base")
    cls = mod.QwenAttention
    if getattr(cls, "_mvb200_fused_qk", False):
        return
    orig_forward = cls.forward

    def _f32_cache(rot, name):
        t = getattr(rot, name)
        if t.dtype == torch.float32:
            return t
        # a loader that casts buffers (model.to(bf16)) leaves rounded cos/sin: the reference then rotates with those
        # values; fp32 copies of them (exact) keep the fused kernel on the same numbers
        key = "_mvb200_" + name
        c = getattr(rot, key, None)
        if c is None or c.device != t.device or c.shape != t.shape:
            c = t.float()
            setattr(rot, key, c)
        return c

    def forward(self, positions, hidden_states):
        rot = self.rotary_emb
        if (hidden_states.dim() != 2 or not hidden_states.is_cuda or not hasattr(self.attn, "forward_fused")
                or not hasattr(rot, "cos_cache") or not hasattr(rot, "sin_cache")
                or (self.q_norm is None) != (self.k_norm is None)):
            return orig_forward(self, positions, hidden_states)
        qkv = self.qkv_proj(hidden_states)
        n = hidden_states.size(0)
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        q = q.view(n, self.num_heads, self.head_dim)
        k = k.view(n, self.num_kv_heads, self.head_dim)
        v = v.view(n, self.num_kv_heads, self.head_dim)
        qw = self.q_norm.weight if self.q_norm is not None else None
        kw = self.k_norm.weight if self.k_norm is not None else None
        eps = float(getattr(self.q_norm, "eps", 1e-6)) if self.q_norm is not None else 1e-6
        o = self.attn.forward_fused(positions, q, k, v, _f32_cache(rot, "cos_cache"), _f32_cache(rot, "sin_cache"), qw, kw, eps)
        return self.o_proj(o.reshape(n, self.num_heads * self.head_dim))

    cls.forward = forward
    cls._mvb200_fused_qk = True


_PEER_COMM: dict = {}


def _peer_comm(device: torch.device):
    """Process-wide ``PeerAllReduce`` endpoint of the tensor-parallel group (created on the first decode-size
    all-reduce, which every rank reaches in the same eager warm-up forward)."""
    comm = _PEER_COMM.get("comm")
    if comm is None and not _PEER_COMM.get("failed"):
        import torch.distributed as dist

        from .comm import PeerAllReduce

        from .comm import PeerSetupError

        gloo = dist.new_group(backend="gloo")  # handle exchange only; a collective every rank makes
        try:
            # The constructor agrees on the outcome across the group: it either succeeds on every rank or raises
            # PeerSetupError on every rank (e.g. no peer access between one pair of devices), so all ranks take the
            # same branch here and nobody is left waiting in a barrier or in the kernel for a rank that fell back.
            comm = PeerAllReduce(dist.get_rank(), dist.get_world_size(), PEER_ALLREDUCE_MAX_BYTES, group=gloo, device=device)
            _PEER_COMM["comm"] = comm
        except PeerSetupError as e:  # stay on NCCL, on all ranks
            _PEER_COMM["failed"] = str(e)
    return comm


def peer_allreduce_status() -> dict:
    """Which collective served the row-parallel all-reduces of this process so far (calls made from Python: a call
    captured into a CUDA graph counts once, at capture).  ``collective`` is ``"mvb200_peer_one_shot"`` when the peer
    kernel is in use, ``"nccl"`` when the set-up failed on some rank (``failed`` says why) or was never installed."""
    comm = _PEER_COMM.get("comm")
    return {
        "installed": bool(_PEER_COMM.get("installed")),
        "collective": "mvb200_peer_one_shot" if comm is not None else "nccl",
        "failed": _PEER_COMM.get("failed") or None,
        "peer_calls": int(_PEER_COMM.get("peer_calls", 0)),
        "nccl_calls": int(_PEER_COMM.get("nccl_calls", 0)),
    }


#: messages up to this size (decode steps: max_num_seqs x hidden elements) take the one-shot peer kernel, larger
#: ones (prefill) stay on NCCL
PEER_ALLREDUCE_MAX_BYTES = 256 * 1024


def install_peer_allreduce() -> None:
    """Rebind ``RowParallelLinear.forward`` (models/layers/linear.py:314-319): the ``dist.all_reduce`` after the
    row-parallel o_proj / down_proj goes through the one-shot NVLink peer kernel for decode-size messages
    (~10 us against ~30 us of NCCL at 8 x 2560 bf16 on 8 GPUs, and CUDA-Graph capturable with no
    per-replay argument).  Same sum (fp32 accumulation in rank order, one rounding); prefill-size messages, CPU
    tensors and 3-D activations stay on the reference's own path."""
    mod = importlib.import_module("minivllm.models.layers.linear")
    cls = mod.RowParallelLinear
    if getattr(cls, "_mvb200_peer", False):
        return
    import torch.distributed as dist
    import torch.nn.functional as F

    def forward(self, x):
        y = F.linear(x, self.weight, self.bias if self.tp_rank == 0 else None)
        if self.tp_size > 1:
            if y.is_cuda and y.dim() == 2 and y.is_contiguous() and y.numel() * y.element_size() <= PEER_ALLREDUCE_MAX_BYTES:
                comm = _peer_comm(y.device)
                if comm is not None and comm.fits(y):
                    _PEER_COMM["peer_calls"] = _PEER_COMM.get("peer_calls", 0) + 1
                    return comm.all_reduce(y, algo=1)
            _PEER_COMM["nccl_calls"] = _PEER_COMM.get("nccl_calls", 0) + 1
            dist.all_reduce(y)
        return y

    cls.forward = forward
    cls._mvb200_peer = True
    _PEER_COMM["installed"] = True


def hot_graph_sizes(max_batch_size: int) -> list[int]:
    from .executor import graph_batch_sizes

    return graph_batch_sizes(max_batch_size)
