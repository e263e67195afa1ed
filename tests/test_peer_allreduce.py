"""One-shot peer all-reduce fused with residual add + RMSNorm.

CPU part pins oracle/allreduce_oracle.py on tests/golden/arnorm.npz (the reference's own RMSNorm module on
an in-dtype sum over ranks).  GPU part (``-m gpu``): the CUDA kernel through the C ABI.  On a single GPU
the ``world`` endpoints live in one process and run concurrently on ``world`` streams — the kernel only
sees peer pointers, so the flag/epoch protocol, the double buffering and the arithmetic are exactly what
runs across processes over NVLink (that transport itself is exercised by tests/mp_peer_allreduce_check.py
under torchrun, and by the test at the bottom when the box has two GPUs).
"""

import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import allreduce_oracle as O
from tests.conftest import ROOT, golden_to_torch, load_golden

CASES = [str(c) for c in load_golden("arnorm")["cases"]]


def _case(name, device="cpu"):
    g = load_golden("arnorm")
    world, rows, hidden, has_res, has_norm = (int(v) for v in g[f"{name}.meta"])
    dt = str(g[f"{name}.dtype"])
    t = lambda key: golden_to_torch(g[f"{name}.{key}"], dt, device)  # noqa: E731
    c = {"world": world, "rows": rows, "hidden": hidden, "dtype": dt,
         "xs": [t(f"x{r}") for r in range(world)], "sum": t("sum"), "y": t("y"),
         "weight": t("weight") if has_norm else None,
         "residual_in": t("residual_in") if has_res else None,
         "residual_out": t("residual_out") if has_res else None}
    return c


def _ulps(a, b, dtype_name):
    """|a-b| in units of the last place of max(|a|,|b|) for the given storage dtype."""
    mant = {"bf16": 7, "f16": 10, "f32": 23}[dtype_name]
    a, b = a.double().cpu(), b.double().cpu()
    scale = torch.maximum(a.abs(), b.abs()).clamp_min(1e-30)
    ulp = torch.exp2(torch.floor(torch.log2(scale)) - mant)
    return ((a - b).abs() / ulp)


# ------------------------------------------------------------------------------------------------ CPU
@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_rmsnorm_on_recorded_sum(name):
    """The norm half, exactly: oracle(rms_norm) on the recorded all-reduce result == the reference module."""
    c = _case(name)
    if c["weight"] is None:
        assert torch.equal(O.all_reduce_sum(c["xs"]), c["sum"]) and torch.equal(c["y"], c["sum"])
        return
    y, res = O.rms_norm(c["sum"], c["weight"], 1e-6, c["residual_in"])
    assert torch.equal(y, c["y"])
    if c["residual_in"] is not None:
        assert torch.equal(res, c["residual_out"])


@pytest.mark.parametrize("name", CASES)
def test_oracle_sum_against_recorded_all_reduce(name):
    """The collective half: exact for two ranks; for more ranks the recorded in-dtype accumulation rounds
    world-2 extra times, each by at most half a unit in the last place of a partial sum."""
    c = _case(name)
    total = O.all_reduce_sum(c["xs"])
    if c["world"] == 2:
        assert torch.equal(total, c["sum"])
        y, res = O.allreduce_add_rms_norm(c["xs"], c["weight"], 1e-6, c["residual_in"])
        assert torch.equal(y, c["y"])
        if res is not None and c["residual_out"] is not None:
            assert torch.equal(res, c["residual_out"])
    else:
        mant = {"bf16": 7, "f16": 10, "f32": 23}[c["dtype"]]
        bound = c["world"] * 2.0 ** -(mant + 1) * sum(x.double().abs() for x in c["xs"])  # one half-ulp per rounding
        assert bool(((total.double() - c["sum"].double()).abs() <= bound).all())


def test_header_declares_peer_entry_points():
    text = open(os.path.join(ROOT, "include", "mvb200.h")).read()
    for sym in ("mvb200_peer_alloc", "mvb200_peer_export", "mvb200_peer_open", "mvb200_peer_allreduce_add_rms_norm"):
        assert sym in text


# ------------------------------------------------------------------------------------------------ GPU
def _run_group(eps, streams, xs, weight, eps_val, residuals, outs=None, **kw):
    """Launch every endpoint's kernel on its own stream (they wait for each other on the device)."""
    results = []
    cur = torch.cuda.current_stream()
    for s in streams:
        s.wait_stream(cur)
    for r, (ep, s) in enumerate(zip(eps, streams)):
        with torch.cuda.stream(s):
            results.append(ep.allreduce_add_rms_norm(xs[r], weight, eps_val, residuals[r] if residuals else None,
                                                     outs[r] if outs else None, **kw))
    for s in streams:
        cur.wait_stream(s)
    return results


@pytest.fixture(scope="module")
def comm_mod():
    from minivllm_b200 import comm

    return comm


@pytest.mark.gpu
@pytest.mark.parametrize("algo", [1, 2])
@pytest.mark.parametrize("name", CASES)
def test_kernel_matches_oracle_and_golden(comm_mod, cuda_device, name, algo):
    c = _case(name, cuda_device)
    world = c["world"]
    eps = comm_mod.PeerAllReduce.local_group(world, 64 * 1024, cuda_device)
    streams = [torch.cuda.Stream(cuda_device) for _ in range(world)]
    try:
        for rep in range(3):  # epochs 1..3: both parities, flag reuse
            residuals = [c["residual_in"].clone() for _ in range(world)] if c["residual_in"] is not None else None
            res = _run_group(eps, streams, c["xs"], c["weight"], 1e-6, residuals, algo=algo)
            torch.cuda.synchronize()
            want_y, want_res = O.allreduce_add_rms_norm([x.cpu() for x in c["xs"]], None if c["weight"] is None else c["weight"].cpu(),
                                                        1e-6, None if c["residual_in"] is None else c["residual_in"].cpu())
            for r in range(world):
                y, new_res = res[r]
                assert torch.equal(y, res[0][0]), "ranks must hold bit-identical results"
                if want_res is not None:
                    assert torch.equal(new_res.cpu(), want_res)           # residual: exact
                    assert new_res.data_ptr() == residuals[r].data_ptr()  # in place
                u = _ulps(y, want_y, c["dtype"])
                assert float(u.max()) <= 1.0 and float((u > 0).float().mean()) < 0.02
                if world == 2:  # also the recorded reference output itself
                    ug = _ulps(y, c["y"], c["dtype"])
                    assert float(ug.max()) <= 1.0 and float((ug > 0).float().mean()) < 0.02
                    if c["residual_out"] is not None:
                        assert torch.equal(new_res, c["residual_out"])
    finally:
        for ep in eps:
            ep.close()


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_kernel_varying_rows_many_epochs(comm_mod, cuda_device, world):
    """Row counts that change the CTA count from call to call (8, 1, 100 > 32 CTAs per rank here, 16, ...).
    With all ranks on ONE GPU their CTAs must be co-resident, hence local_group's 32-CTA cap; one rank per
    GPU — the deployment — runs one CTA per row up to 256."""
    hidden = 2560
    row_counts = [8, 1, 100, 16, 8, 128, 3, 64, 65, 8, 8]
    eps = comm_mod.PeerAllReduce.local_group(world, 128 * hidden * 2, cuda_device)
    streams = [torch.cuda.Stream(cuda_device) for _ in range(world)]
    g = torch.Generator().manual_seed(world)
    weight = (1.0 + 0.1 * torch.randn(hidden, generator=g)).to(torch.bfloat16).to(cuda_device)
    try:
        for it, rows in enumerate(row_counts):
            xs = [torch.randn(rows, hidden, generator=g).to(torch.bfloat16).to(cuda_device) for _ in range(world)]
            res0 = torch.randn(rows, hidden, generator=g).to(torch.bfloat16).to(cuda_device)
            residuals = [res0.clone() for _ in range(world)]
            out = _run_group(eps, streams, xs, weight, 1e-6, residuals, algo=1 + (it * 7 // 3) % 2)  # mixes both kernels
            if it % 3 == 2:
                torch.cuda.synchronize()  # otherwise back-to-back launches (programmatic dependent launch overlap)
            want_y, want_res = O.allreduce_add_rms_norm([x.cpu() for x in xs], weight.cpu(), 1e-6, res0.cpu())
            for r in range(world):
                assert torch.equal(out[r][1].cpu(), want_res), (it, rows, r)
                assert float(_ulps(out[r][0], want_y, "bf16").max()) <= 1.0, (it, rows, r)
        # plain all-reduce through the same buffers
        xs = [torch.randn(8, hidden, generator=g).to(torch.bfloat16).to(cuda_device) for _ in range(world)]
        cur = torch.cuda.current_stream()
        for s in streams:
            s.wait_stream(cur)
        ys = []
        for ep, s, x in zip(eps, streams, xs):
            with torch.cuda.stream(s):
                ys.append(ep.all_reduce(x))
        torch.cuda.synchronize()
        for y in ys:
            assert torch.equal(y.cpu(), O.all_reduce_sum([x.cpu() for x in xs]))
    finally:
        for ep in eps:
            ep.close()


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_twoshot_sharded_residual(comm_mod, cuda_device, world):
    """gather_residual=False: y is complete everywhere, the residual is updated only on the rows a rank owns
    (row r -> rank (r % cap) % world, cap = the CTA cap rounded down to a multiple of world)."""
    rows, hidden, cap = 45, 1024, 32
    eps = comm_mod.PeerAllReduce.local_group(world, rows * hidden * 2, cuda_device, max_ctas=cap)
    streams = [torch.cuda.Stream(cuda_device) for _ in range(world)]
    g = torch.Generator().manual_seed(40 + world)
    weight = (1.0 + 0.1 * torch.randn(hidden, generator=g)).to(torch.bfloat16).to(cuda_device)
    try:
        for it in range(3):
            xs = [torch.randn(rows, hidden, generator=g).to(torch.bfloat16).to(cuda_device) for _ in range(world)]
            res0 = torch.randn(rows, hidden, generator=g).to(torch.bfloat16).to(cuda_device)
            residuals = [res0.clone() for _ in range(world)]
            out = _run_group(eps, streams, xs, weight, 1e-6, residuals, algo=2, gather_residual=False)
            torch.cuda.synchronize()
            want_y, want_res = O.allreduce_add_rms_norm([x.cpu() for x in xs], weight.cpu(), 1e-6, res0.cpu())
            owner = (torch.arange(rows) % (cap - cap % world)) % world
            for r in range(world):
                assert torch.equal(out[r][0], out[0][0])
                assert float(_ulps(out[r][0], want_y, "bf16").max()) <= 1.0
                got = residuals[r].cpu()
                assert torch.equal(got[owner == r], want_res[owner == r])
                assert torch.equal(got[owner != r], res0.cpu()[owner != r])
    finally:
        for ep in eps:
            ep.close()


@pytest.mark.gpu
def test_kernel_replays_from_cuda_graphs(comm_mod, cuda_device):
    world, rows, hidden = 2, 8, 1024
    eps = comm_mod.PeerAllReduce.local_group(world, rows * hidden * 2, cuda_device)
    streams = [torch.cuda.Stream(cuda_device) for _ in range(world)]
    g = torch.Generator().manual_seed(5)
    weight = torch.ones(hidden, dtype=torch.bfloat16, device=cuda_device)
    xs = [torch.zeros(rows, hidden, dtype=torch.bfloat16, device=cuda_device) for _ in range(world)]
    res = [torch.zeros(rows, hidden, dtype=torch.bfloat16, device=cuda_device) for _ in range(world)]
    outs = [torch.zeros(rows, hidden, dtype=torch.bfloat16, device=cuda_device) for _ in range(world)]
    try:
        _run_group(eps, streams, xs, weight, 1e-6, res, outs)  # warm-up, epoch 1
        torch.cuda.synchronize()
        graphs = []
        for r in range(world):  # capture does not execute: both ranks record one call (epoch comes from memory)
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                eps[r].allreduce_add_rms_norm(xs[r], weight, 1e-6, res[r], outs[r])
            graphs.append(gr)
        for it in range(4):
            new = [torch.randn(rows, hidden, generator=g).to(torch.bfloat16) for _ in range(world)]
            r0 = torch.randn(rows, hidden, generator=g).to(torch.bfloat16)
            for r in range(world):
                xs[r].copy_(new[r])
                res[r].copy_(r0)
            torch.cuda.synchronize()
            for r in range(world):
                with torch.cuda.stream(streams[r]):
                    graphs[r].replay()
            torch.cuda.synchronize()
            want_y, want_res = O.allreduce_add_rms_norm(new, weight.cpu(), 1e-6, r0)
            for r in range(world):
                assert torch.equal(res[r].cpu(), want_res)
                assert float(_ulps(outs[r], want_y, "bf16").max()) <= 1.0
    finally:
        del graphs
        for ep in eps:
            ep.close()


@pytest.mark.gpu
def test_error_contracts(comm_mod, cuda_device):
    ep = comm_mod.PeerAllReduce.local_group(1, 4096, cuda_device)[0]
    try:
        x = torch.ones(2, 512, dtype=torch.bfloat16, device=cuda_device)
        assert torch.equal(ep.all_reduce(x), x)  # world 1 is the identity
        with pytest.raises(ValueError, match="does not fit"):
            ep.all_reduce(torch.ones(8, 512, dtype=torch.bfloat16, device=cuda_device))
        with pytest.raises(ValueError, match="does not fit"):
            ep.all_reduce(torch.ones(2, 100, dtype=torch.bfloat16, device=cuda_device)[:, :99])
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            ep.all_reduce(x.cpu())
        with pytest.raises(ValueError):
            comm_mod.PeerAllReduce(0, 9, 1024, _defer=True)
    finally:
        ep.close()


@pytest.mark.gpu
def test_two_processes_over_cuda_ipc(cuda_device):
    """Real transport: two processes, CUDA IPC handles exchanged through torch.distributed, NVLink P2P."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29731", os.path.join(ROOT, "tests", "mp_peer_allreduce_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0 and "PEER_ALLREDUCE_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
