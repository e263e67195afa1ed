#!/usr/bin/env python
"""CUDA-event timing of the HBM-bound helper kernels of the path at BASELINE sizes (one GPU):
store_kvcache, qk_norm_rope_store, greedy_argmax, decode_advance.  Prints one JSON line per case with the
algorithmic bytes and the fraction of the measured copy peak (MEASURED_PEAKS.json).

L2 is flushed between timed launches (192 MB write) because several of these working sets fit the 126 MB L2.
"""

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch

    import minivllm_b200 as m

    dev = torch.device("cuda:0")
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        src = "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        peak, src = 6500.0, "fallback 6500 GB/s"
    flush = torch.empty(192 << 20, dtype=torch.uint8, device=dev)

    def timed(fn, iters=20, warm=5):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(iters):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            tot += a.elapsed_time(b)
        return tot / iters * 1e3  # us

    def report(name, nbytes, us, **extra):
        gbs = nbytes / us * 1e-3
        print(json.dumps({"kernel": name, "algorithmic_bytes": nbytes, "us_per_launch": round(us, 2), "GB_per_s": round(gbs, 1),
                          "frac_of_measured_hbm_peak": round(gbs / peak, 3), "peak_GB_per_s": peak, "peak_source": src, **extra}),
              flush=True)

    bf = torch.bfloat16
    # ---- store_kvcache: cfg3 prefill (8192 new tokens, KVH*D = 1024) and cfg2 decode (256 tokens)
    for n, kvh in ((8192, 8), (16384, 1), (256, 8)):
        d, nb = 128, max(2 * n // 64, 64)
        k = torch.randn(n, kvh, d, device=dev).to(bf)
        v = torch.randn(n, kvh, d, device=dev).to(bf)
        kc = torch.zeros(nb, 64, kvh, d, dtype=bf, device=dev)
        vc = torch.zeros_like(kc)
        slots = torch.randperm(nb * 64, device=dev)[:n].to(torch.int32)
        us = timed(lambda: m.store_kvcache(k, v, kc, vc, slots))
        report("store_kvcache", n * kvh * d * 2 * 2 * 2 + 4 * n, us, tokens=n, kv_heads=kvh)
    # ---- qk_norm_rope_store: Qwen3-4B rank shapes, prefill 8192 tokens and decode 8 / 256 tokens
    for n, h, kvh in ((8192, 32, 8), (256, 16, 8), (8, 32, 8)):
        d, nb, max_pos = 128, max(2 * n // 64, 64), 4096
        qkv = torch.randn(n, (h + 2 * kvh) * d, device=dev).to(bf)
        q, k, v = qkv.split([h * d, kvh * d, kvh * d], dim=-1)
        q, k, v = q.view(n, h, d), k.view(n, kvh, d), v.view(n, kvh, d)
        pos = torch.randint(0, max_pos, (n,), device=dev)
        inv = 1.0 / (1e6 ** (torch.arange(0, d, 2, dtype=torch.float, device=dev) / d))
        fr = torch.einsum("i,j->ij", torch.arange(max_pos, dtype=torch.float, device=dev), inv)
        cos, sin = fr.cos().contiguous(), fr.sin().contiguous()
        qw, kw = torch.ones(d, dtype=bf, device=dev), torch.ones(d, dtype=bf, device=dev)
        kc = torch.zeros(nb, 64, kvh, d, dtype=bf, device=dev)
        vc = torch.zeros_like(kc)
        slots = torch.randperm(nb * 64, device=dev)[:n].to(torch.int32)
        us = timed(lambda: m.ops.qk_norm_rope_store(q, k, v, pos, cos, sin, qw, kw, 1e-6, slots, kc, vc))
        nbytes = n * ((h + 2 * kvh) * d * 2 + (h + kvh) * d * 2 + 2 * kvh * d * 2 + 12)
        report("qk_norm_rope_store", nbytes, us, tokens=n, heads=h, kv_heads=kvh)
    # ---- greedy_argmax over the Qwen vocabulary
    for rows in (8, 256, 512):
        vocab = 151936
        logits = torch.randn(rows, vocab, device=dev).to(bf)
        out = torch.empty(rows, dtype=torch.int64, device=dev)
        ws = torch.empty(max(m.ops.greedy_argmax_workspace_numel(rows, vocab), 1), dtype=torch.int64, device=dev)
        us = timed(lambda: m.ops.greedy_argmax(logits, out, workspace=ws))
        report("greedy_argmax", rows * vocab * 2 + rows * 8, us, rows=rows, vocab=vocab)
    # ---- decode_advance (latency-bound: 512 rows x 40 B)
    b, mb = 512, 64
    st = {"next_tokens": torch.zeros(b, dtype=torch.int64, device=dev), "appends": torch.full((b,), -1, dtype=torch.int32, device=dev),
          "input_ids": torch.zeros(b, dtype=torch.int64, device=dev), "positions": torch.zeros(b, dtype=torch.int64, device=dev),
          "slot_mapping": torch.zeros(b, dtype=torch.int32, device=dev), "context_lens": torch.full((b,), 100, dtype=torch.int32, device=dev),
          "block_tables": torch.zeros(b, mb, dtype=torch.int32, device=dev)}

    def adv():
        st["context_lens"].fill_(100)
        m.ops.decode_advance(st["next_tokens"], st["appends"], st["input_ids"], st["positions"], st["slot_mapping"],
                             st["context_lens"], st["block_tables"], 4096)

    us = timed(adv)
    report("decode_advance (+ a fill kernel)", b * 44, us, rows=b)


if __name__ == "__main__":
    main()
