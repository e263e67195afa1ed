#!/usr/bin/env python
"""Time the reference's real CUDA attention (flash-attn 2 call sites, attention.py:422-447) beside the
mvb200 kernels on the same logical K/V — the bar SURVEY section 1 names.  Tools only: the product never
dispatches to flash-attn.  flash-attn pages K/V in 256-token blocks, mvb200 in the reference's 64.

    python tools/fa2_compare.py [--out gpurun_out/fa2_compare.json]
"""

from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(fn, iters=20, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    return ts[len(ts) // 2] * 1e3  # median, microseconds


def decode_case(name, b, ctx, h, kvh, d=128):
    import minivllm_b200 as m
    from flash_attn import flash_attn_with_kvcache

    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    out = {"case": name, "B": b, "ctx": ctx, "H": h, "KVH": kvh}
    bytes_ = b * ctx * kvh * d * 2 * 2
    q = torch.randn(b, h, d, device=dev, generator=g).to(torch.bfloat16)
    lens = torch.full((b,), ctx, dtype=torch.int32, device=dev)
    for page, key in ((64, "mvb200"), (256, "fa2")):
        per = (ctx + page - 1) // page
        nb = b * per
        kc = torch.randn(nb, page, kvh, d, device=dev, generator=g).to(torch.bfloat16)
        vc = torch.randn(nb, page, kvh, d, device=dev, generator=g).to(torch.bfloat16)
        bt = torch.randperm(nb, device=dev, generator=g).to(torch.int32).view(b, per)
        if key == "mvb200":
            fn = lambda: m.paged_decode_attention(q, kc, vc, bt, lens, d**-0.5)  # noqa: E731
        else:
            q4 = q.unsqueeze(1)
            fn = lambda: flash_attn_with_kvcache(q4, kc, vc, cache_seqlens=lens, block_table=bt,  # noqa: E731
                                                 softmax_scale=d**-0.5, causal=True)
        try:
            us = timed(fn)
            out[key] = {"us": us, "GBps": bytes_ / us / 1e3}
        except Exception as e:  # noqa: BLE001
            out[key] = {"error": (type(e).__name__ + ": " + str(e))[:300]}
        del kc, vc
    return out


def prefill_case(name, seqs, cached, new, h, kvh, d=128):
    import minivllm_b200 as m
    from flash_attn import flash_attn_varlen_func

    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    lk = cached + new
    flops = 4 * h * d * seqs * (new * cached + new * (new + 1) // 2)
    out = {"case": name, "seqs": seqs, "cached": cached, "new": new, "H": h, "KVH": kvh, "flop": flops}
    q = torch.randn(seqs * new, h, d, device=dev, generator=g).to(torch.bfloat16)
    cu_q = torch.arange(0, seqs + 1, device=dev, dtype=torch.int32) * new
    cu_k = torch.arange(0, seqs + 1, device=dev, dtype=torch.int32) * lk
    for page, key in ((64, "mvb200"), (256, "fa2")):
        per = (lk + page - 1) // page
        nb = seqs * per
        kc = torch.randn(nb, page, kvh, d, device=dev, generator=g).to(torch.bfloat16)
        vc = torch.randn(nb, page, kvh, d, device=dev, generator=g).to(torch.bfloat16)
        bt = torch.randperm(nb, device=dev, generator=g).to(torch.int32).view(seqs, per)
        if key == "mvb200":
            fn = lambda: m.prefill_varlen_attention(q, None, None, cu_q, cu_k, new, lk, d**-0.5,  # noqa: E731
                                                    k_cache=kc, v_cache=vc, block_tables=bt)
        else:
            fn = lambda: flash_attn_varlen_func(q, kc, vc, cu_seqlens_q=cu_q, cu_seqlens_k=cu_k, max_seqlen_q=new,  # noqa: E731
                                                max_seqlen_k=lk, softmax_scale=d**-0.5, causal=True, block_table=bt)
        try:
            us = timed(fn)
            out[key] = {"us": us, "TFLOPs": flops / us / 1e6}
        except Exception as e:  # noqa: BLE001
            out[key] = {"error": (type(e).__name__ + ": " + str(e))[:300]}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "fa2_compare.json"))
    a = ap.parse_args()
    rows = [
        decode_case("cfg2 decode 256x2048 KVH8", 256, 2048, 16, 8),
        decode_case("cfg2 TP8-rank 256x2048 KVH1", 256, 2048, 2, 1),
        decode_case("cfg3 decode 8x2304 KVH8", 8, 2304, 32, 8),
        decode_case("cfg4 decode 1x32768 KVH4", 1, 32768, 28, 4),
        decode_case("cfg5-rank decode 512x1024 KVH1", 512, 1024, 8, 1),
        prefill_case("cfg3p 8x(1024+1024) H32", 8, 1024, 1024, 32, 8),
        prefill_case("cfg3p1 1x(4096+4096) H32", 1, 4096, 4096, 32, 8),
        prefill_case("cfg3p TP8-rank H4", 8, 1024, 1024, 4, 1),
    ]
    for r in rows:
        print(json.dumps(r))
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(rows, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
