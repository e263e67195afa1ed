#!/usr/bin/env python
"""Run under torchrun on N GPUs of one node: parity and latency of the fused peer all-reduce + add + RMSNorm
(minivllm_b200/comm.py) against NCCL all_reduce followed by the harness add_rms_norm kernel.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/mp_peer_allreduce_check.py

Rank 0 prints one JSON line per shape and finally PEER_ALLREDUCE_OK.  The oracle is used here only as the
checker (this tool is test infrastructure)."""

import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist

    from harness import fused
    from minivllm_b200.comm import NvlsAllReduce, PeerAllReduce
    from oracle import allreduce_oracle as O

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    gloo = dist.new_group(backend="gloo")
    shapes = [(8, 2560), (1, 3584), (8, 1024), (32, 2560), (64, 2560), (256, 1024), (256, 8192), (512, 8192)]
    nvls_note = "ok"
    try:  # the NVLS (multimem) variant needs a multicast mapping of a symmetric buffer: NVSwitch fabric + driver support
        comm = NvlsAllReduce(rank, world, max(r * h for r, h in shapes) * 2, group=gloo, device=dev)
    except Exception as e:  # noqa: BLE001
        nvls_note = repr(e)[:200]
        comm = PeerAllReduce(rank, world, max(r * h for r, h in shapes) * 2, group=gloo, device=dev)
    has_nvls = isinstance(comm, NvlsAllReduce)
    if rank == 0:
        print(json.dumps({"nvls": has_nvls, "note": nvls_note}), flush=True)
    ok = True
    for rows, hidden in shapes:
        g = torch.Generator().manual_seed(1000 + rows + hidden)  # same stream on every rank
        xs = [torch.randn(rows, hidden, generator=g).to(torch.bfloat16) for _ in range(world)]
        res0 = torch.randn(rows, hidden, generator=g).to(torch.bfloat16)
        w = (1.0 + 0.1 * torch.randn(hidden, generator=g)).to(torch.bfloat16)
        x, wd = xs[rank].to(dev), w.to(dev)
        res = res0.to(dev)
        want_y, want_res = O.allreduce_add_rms_norm(xs, w, 1e-6, res0)
        exact_res, dy = True, 0.0
        nvls_res_ulps, nvls_dy = None, None
        for algo in (1, 2, 3) if has_nvls else (1, 2):
            res = res0.to(dev)
            y, new_res = comm.allreduce_add_rms_norm(x, wd, 1e-6, res, algo=algo)
            torch.cuda.synchronize()
            if algo == 3:
                # the switch sums in its own order and converts fp32 -> bf16 its own way (as NCCL's NVLS path does):
                # the new residual is held to one bf16 unit in the last place of the LARGEST addend of the element
                a, b = new_res.cpu().float(), want_res.float()
                big = torch.stack([t_.float().abs() for t_ in xs] + [res0.float().abs()]).max(dim=0).values
                big = torch.maximum(big, b.abs())  # ... or of the sum itself, whichever is larger
                nvls_res_ulps = ((a - b).abs() / (big.clamp_min(2.0 ** -120) * 2.0 ** -7)).max().item()
                nvls_dy = (y.cpu().float() - want_y.float()).abs().max().item()
                continue
            exact_res = exact_res and bool(torch.equal(new_res.cpu(), want_res))
            dy = max(dy, (y.cpu().float() - want_y.float()).abs().max().item())
        # NCCL + separate norm kernel on the same inputs
        x2, res2 = xs[rank].to(dev), res0.to(dev)
        dist.all_reduce(x2)
        y2, r2 = fused.add_rms_norm(x2, wd, 1e-6, res2)
        d_nccl = (y.float() - y2.float()).abs().max().item()

        def timed(fn, iters=200):
            for _ in range(20):
                fn()
            dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(iters):
                fn()
            b.record()
            torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b) / iters * 1e3], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t.item()

        out = torch.empty_like(x)

        def run_peer():
            comm.allreduce_add_rms_norm(x, wd, 1e-6, res, out, algo=1)

        def run_two():
            comm.allreduce_add_rms_norm(x, wd, 1e-6, res, out, algo=2)

        def run_two_sharded():
            comm.allreduce_add_rms_norm(x, wd, 1e-6, res, out, algo=2, gather_residual=False)

        def run_nvls():
            comm.allreduce_add_rms_norm(x, wd, 1e-6, res, out, algo=3)

        def run_nvls_sharded():
            comm.allreduce_add_rms_norm(x, wd, 1e-6, res, out, algo=3, gather_residual=False)

        def run_nccl():
            dist.all_reduce(x2)
            fused.add_rms_norm(x2, wd, 1e-6, res2)

        def graphed(fn, n=20):
            fn()
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                for _ in range(n):
                    fn()
            return lambda: gr.replay(), n

        t_peer = timed(run_peer)
        t_nccl = timed(run_nccl)
        gp, n = graphed(run_peer)
        t_peer_g = timed(gp, 50) / n
        gn, n = graphed(run_nccl)
        t_nccl_g = timed(gn, 50) / n
        g2, n = graphed(run_two)
        t_two_g = timed(g2, 50) / n
        g3, n = graphed(run_two_sharded)
        t_two_sh_g = timed(g3, 50) / n
        t_nvls_g = t_nvls_sh_g = None
        if has_nvls:
            g4, n = graphed(run_nvls)
            t_nvls_g = round(timed(g4, 50) / n, 2)
            g5, n = graphed(run_nvls_sharded)
            t_nvls_sh_g = round(timed(g5, 50) / n, 2)
        good = exact_res and dy <= 0.0625 and d_nccl <= 0.125 and (nvls_res_ulps is None or (nvls_res_ulps <= 4.0 and nvls_dy <= 0.0625))
        ok = ok and good
        if rank == 0:
            print(json.dumps({"rows": rows, "hidden": hidden, "world": world, "bytes": rows * hidden * 2,
                              "residual_exact": exact_res, "max_abs_vs_oracle": dy, "max_abs_vs_nccl_path": d_nccl,
                              "us_peer_fused": round(t_peer, 2), "us_nccl_plus_norm": round(t_nccl, 2),
                              "us_peer_fused_graph": round(t_peer_g, 2), "us_nccl_plus_norm_graph": round(t_nccl_g, 2),
                              "us_twoshot_graph": round(t_two_g, 2), "us_twoshot_sharded_residual_graph": round(t_two_sh_g, 2),
                              "us_nvls_graph": t_nvls_g, "us_nvls_sharded_residual_graph": t_nvls_sh_g,
                              "nvls_residual_max_ulps": nvls_res_ulps, "nvls_max_abs_vs_oracle": nvls_dy}), flush=True)
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("PEER_ALLREDUCE_OK" if flag.item() == 1 else "PEER_ALLREDUCE_MISMATCH", flush=True)
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
