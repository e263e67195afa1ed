"""N > 1 on CPU: two ``gloo`` ranks run the KV-head-sharded path and must reproduce the
single-rank results.

What is sharded (qwen_base.py:95-108, linear.py:229-270,316-319): rank r owns q heads
``[r*H/tp, (r+1)*H/tp)``, kv heads ``[r*KVH/tp, (r+1)*KVH/tp)`` and its own cache slice; metadata is
identical on every rank; the only collective is the all-reduce after the row-parallel projections
(plus the harness's vocab-sharded argmax all-gather).  The attention double on CPU is the oracle
(tests/oracle_attention.py); the CUDA layer takes its place under ``-m gpu``.
"""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests.conftest import load_golden


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, preset: str, prompts, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from harness.engine import Engine
        from harness.model import PRESETS
        from tests.oracle_attention import OracleAttention

        eng = Engine(PRESETS[preset], device="cpu", dtype=torch.float32, tp=world, rank=rank, max_num_seqs=8,
                     max_model_len=1024, num_blocks=64, enforce_eager=True, init="seeded", attention_cls=OracleAttention)
        # each rank holds its own slice of the KV heads
        assert eng.exe.kv_cache.shape[4] == PRESETS[preset].num_kv_heads // world
        outs, _ = eng.generate(prompts, 16)
        # metadata must be identical on all ranks: compare the block tables the pools produced
        tables = [None] * world
        dist.all_gather_object(tables, eng.pool.hits)
        assert len(set(tables)) == 1
        ret[rank] = outs
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("preset", ["tiny-qwen3", "tiny-qwen2"])
def test_tp2_gloo_reproduces_reference_token_streams(preset):
    g = load_golden("engine")
    n = int(g[f"{preset}.n"])
    prompts = [g[f"{preset}.prompt{i}"].tolist() for i in range(n)]
    want = [g[f"{preset}.tokens{i}"].tolist()[:16] for i in range(n)]
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, 2, port, preset, prompts, ret)) for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(300)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert ret[0] == ret[1] == want


def _shard_worker(rank: int, world: int, port: int, ret):
    """The path itself has no collective: each rank attends over its own heads; gathering the per-rank
    outputs along the head axis must equal the unsharded oracle."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import paged_attention_oracle as orc

        rng = np.random.default_rng(0)  # same inputs on every rank
        b, h, kvh, d, bs, nb = 3, 8, 4, 32, 64, 12
        q = rng.standard_normal((b, h, d)).astype(np.float32)
        kc = rng.standard_normal((nb, bs, kvh, d)).astype(np.float32)
        vc = rng.standard_normal((nb, bs, kvh, d)).astype(np.float32)
        bt = rng.permutation(nb).astype(np.int32).reshape(b, 4)
        ctx_lens = np.array([200, 64, 129])
        hs, ks = h // world, kvh // world
        mine = orc.paged_decode(q[:, rank * hs:(rank + 1) * hs], kc[:, :, rank * ks:(rank + 1) * ks],
                                vc[:, :, rank * ks:(rank + 1) * ks], bt, ctx_lens, d ** -0.5)
        parts = [torch.empty(b, hs, d, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(parts, torch.from_numpy(np.ascontiguousarray(mine, dtype=np.float64)))
        full = orc.paged_decode(q, kc, vc, bt, ctx_lens, d ** -0.5)
        ret[rank] = float(np.abs(torch.cat(parts, dim=1).numpy() - full).max())
    finally:
        dist.destroy_process_group()


def test_kv_head_sharding_needs_no_collective_inside_the_path():
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_shard_worker, args=(r, 2, port, ret)) for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert ret[0] < 1e-12 and ret[1] < 1e-12


def _ref_engine_worker(rank: int, world: int, port: int, prompts, ret):
    """The REFERENCE's own engine at TP=2 inside launcher-started ranks (what bench.py's engine leg does under
    torchrun): process group first, then the model (defect 1), per-rank KV heads (defect 2), no spawn."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), LOCAL_RANK=str(rank),
                      WORLD_SIZE=str(world), MINIVLLM_TP_BACKEND="gloo")
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from harness import ref_engine
    from oracle import reference_engine as R

    # the optional rebindings must leave everything they do not cover on the reference's own path: on CPU tensors
    # QwenAttention.forward and RowParallelLinear.forward (dist.all_reduce over gloo) stay the reference's
    R.import_reference("cpu", hide_flash_attn=True, isolate_cuda=True)
    from minivllm_b200 import reference_patch

    reference_patch.install_fused_qk()
    reference_patch.install_peer_allreduce()
    info = ref_engine.run("tiny-qwen3", device="cpu", dtype="float32", executor=True, calls=[prompts], max_tokens=16,
                          tensor_parallel_size=world, under_torchrun=True, max_num_seqs=8, max_model_len=1024,
                          num_kvcache_blocks=64)
    if rank == 0:
        ret["tokens"] = info["calls"][0]["tokens"]
        ret["kv_heads_rank0"] = info["kv_heads_per_rank"]
    else:
        ret["worker"] = bool(info.get("worker"))


@pytest.mark.skipif(not os.path.isdir("/root/reference/minivllm") and not os.path.isdir(
    os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "minivllm")),
    reason="needs the reference (or oracle/_ref)")
def test_reference_engine_tp2_gloo_matches_its_single_rank_stream():
    g = load_golden("engine")
    n = int(g["tiny-qwen3.n"])
    prompts = [g[f"tiny-qwen3.prompt{i}"].tolist() for i in range(n)]
    want = [g[f"tiny-qwen3.tokens{i}"].tolist()[:16] for i in range(n)]
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_ref_engine_worker, args=(r, 2, port, prompts, ret)) for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(600)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert ret["worker"] is True
        assert ret["kv_heads_rank0"] == 1  # tiny-qwen3 has 2 KV heads: one per rank in the cache
        assert ret["tokens"] == want


def _peer_setup_worker(rank: int, world: int, port: int, fail_rank: int, fail_at: str, ret):
    """PeerAllReduce set-up with one rank failing alone (device calls stubbed: no GPU here)."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import contextlib

        from minivllm_b200 import comm

        closed = []
        comm.PeerAllReduce._alloc = lambda self, n: (_ for _ in ()).throw(RuntimeError("out of memory")) \
            if (fail_at == "alloc" and rank == fail_rank) else 0x1000 * (rank + 1)
        comm.PeerAllReduce._export = lambda self, ptr: bytes([rank]) * 64
        comm.PeerAllReduce._open = lambda self, h: (_ for _ in ()).throw(RuntimeError("peer access unsupported")) \
            if (fail_at == "open" and rank == fail_rank) else 0x9000
        comm.PeerAllReduce.close = lambda self: closed.append(rank)
        comm.torch.cuda.device = lambda d: contextlib.nullcontext()
        try:
            comm.PeerAllReduce(rank, world, 4096, device="cpu")
            ret[rank] = "ok"
        except comm.PeerSetupError as e:
            ret[rank] = f"{e}|closed={closed}"
        dist.barrier()  # every rank got here: nobody is stuck in the set-up
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("fail_at", ["alloc", "open", "none"])
def test_peer_allreduce_setup_outcome_is_collective(fail_at):
    """a rank whose IPC open (or allocation) fails alone must not fall back to NCCL by itself while
    its peers wait for it — every rank raises PeerSetupError (and closes its endpoint) or every rank succeeds."""
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_peer_setup_worker, args=(r, 2, port, 1, fail_at, ret)) for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        if fail_at == "none":
            assert ret[0] == ret[1] == "ok"
        else:
            for r in range(2):
                assert "rank(s) [1]" in ret[r] and f"closed=[{r}]" in ret[r], ret[r]
