#!/usr/bin/env python
"""Per-tile pipeline timeline of the tcgen05 prefill kernel (debug tool, B200 only).

    python minivllm_b200/build.py --trace        # here: builds libmvb200_trace.so (-DPF_TRACE)
    python tools/prefill_trace.py                  # on the GPU box

One CTA (longest query tile of a middle head/sequence) records clock64 at every hand-off of its TMA
producer, MMA issuer and first softmax warp; this prints the deltas per 64-key tile.
"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["MVB200_LIB"] = os.path.join(ROOT, "minivllm_b200", "libmvb200_trace.so")
os.environ.setdefault("MVB200_PREFILL_KERNEL", "fa2q")  # the timeline instrumentation lives in the 128-key-tile kernel
V1 = "--v1" in sys.argv  # trace the one-query-tile kernel instead of the persistent two-tile one
if V1:
    os.environ["MVB200_PREFILL_V1"] = "1"

import numpy as np
import torch

import minivllm_b200 as m
from minivllm_b200 import _native


def main():
    dev = torch.device("cuda:0")
    S, C, NQ, H, KVH, D, BS = 8, 1024, 1024, 32, 8, 128, 64
    torch.manual_seed(0)
    lk = C + NQ
    nb_seq = lk // BS
    nb = S * nb_seq + 8
    kc = torch.randn(nb, BS, KVH, D, device=dev).to(torch.bfloat16)
    vc = torch.randn(nb, BS, KVH, D, device=dev).to(torch.bfloat16)
    perm = torch.randperm(nb)[: S * nb_seq].view(S, nb_seq).to(torch.int32).to(dev)
    q = torch.randn(S * NQ, H, D, device=dev).to(torch.bfloat16)
    cu_q = torch.arange(0, S + 1, dtype=torch.int32, device=dev) * NQ
    cu_k = torch.arange(0, S + 1, dtype=torch.int32, device=dev) * lk
    for _ in range(3):
        out = m.ops.prefill_varlen_attention(q, None, None, cu_q, cu_k, NQ, lk, D**-0.5, k_cache=kc, v_cache=vc, block_tables=perm)
    torch.cuda.synchronize()
    if not V1:
        return report_v2()
    buf = (ctypes.c_ulonglong * (3 * 64 * 8))()
    fn = _native.lib().mvb200_debug_read_prefill_trace
    fn.argtypes = [ctypes.c_void_p]
    rc = fn(buf)
    assert rc == 0, rc
    t = np.frombuffer(buf, dtype=np.uint64).astype(np.int64).reshape(3, 64, 8)
    t0 = t[1, 0, 0]
    n = 32
    rel = lambda x: int(x - t0)
    print("MMA issuer: j | K_full wait | PVdone(j-2) wait | QK issued@ | [pv j-1] V_full wait | P_full wait | PV issued@")
    for j in range(n):
        r = t[1, j]
        print(f"  {j:2d} | start {rel(r[0]):7d} | Kfull +{r[1]-r[0]:5d} | Sfree +{r[2]-r[1]:5d} | issue +{r[3]-r[2]:4d} (@{rel(r[3])}) "
              f"|| PV({j}): Vfull@{rel(r[4])} Pfull +{r[5]-r[4]:5d} issue +{r[6]-r[5]:4d} (@{rel(r[6])})")
    print("softmax warp 2: j | wait S_full | LDTM | max/rescale | exp | STTM | arrive | period")
    prev = None
    for j in range(n):
        r = t[2, j]
        per = (r[6] - prev) if prev is not None else 0
        prev = r[6]
        print(f"  {j:2d} | @{rel(r[0]):7d} | S_full +{r[1]-r[0]:5d} | ld +{r[2]-r[1]:4d} | max +{r[3]-r[2]:4d} | exp +{r[4]-r[3]:4d} "
              f"| st +{r[5]-r[4]:4d} | arr +{r[6]-r[5]:4d} | busy {r[6]-r[1]:5d} | period {per:5d}")
    cta_report(8 * 32 * 8)
    print("producer: j | K_EMPTY wait | V_EMPTY wait")
    for j in range(n):
        r = t[0, j]
        print(f"  {j:2d} | @{rel(r[0]):7d} | Kempty +{r[1]-r[0]:5d} | Vempty +{r[2]-r[1]:5d}")


def report_v2():
    """Second work item of CTA 0 (steady state: its loads were prefetched during the first)."""
    buf = (ctypes.c_ulonglong * (4 * 64 * 8))()
    fn = _native.lib().mvb200_debug_read_prefill2_trace
    fn.argtypes = [ctypes.c_void_p]
    assert fn(buf) == 0
    t = np.frombuffer(buf, dtype=np.uint64).astype(np.int64).reshape(4, 64, 8)
    t0 = t[0, 62, 0]
    rel = lambda x: int(x - t0)
    n = 17
    for q in (0, 1):
        print(f"MMA issuer, chain of query tile {q}: j | first P chunk issued @ | last chunk issued | QK(j+1) issued | period")
        prev = None
        for j in range(n):
            r = t[q, j]
            if r[1] == 0 or r[1] < t0:
                break
            per = (r[1] - prev) if prev is not None else 0
            prev = r[1]
            qk = f"qk +{r[6]-r[4]:5d}" if r[6] > r[4] else "qk   (none)"
            print(f"  {j:2d} | @{rel(r[1]):7d} | lastP +{r[4]-r[1]:5d} | {qk} | period {per:5d}")
    for wg in (0, 1):
        print(f"softmax warpgroup {wg}: j | wait S_full | LDTM | mask/max/rescale | exp+STTM+arrivals | busy | period")
        prev = None
        for j in range(n):
            r = t[2 + wg, j]
            if r[1] == 0 or r[1] < t[0, 62, 0]:
                break
            per = (r[6] - prev) if prev is not None else 0
            prev = r[6]
            print(f"  {j:2d} | @{rel(r[0]):7d} | S_full +{r[1]-r[0]:5d} | ld +{r[2]-r[1]:4d} | max +{r[3]-r[2]:4d} | exp +{r[4]-r[3]:4d} "
                  f"| busy {r[6]-r[1]:5d} | period {per:5d}")
        print(f"  epilogue done @{rel(t[2 + wg, 63, 0])}")


def cta_report(num_ctas):
    buf = (ctypes.c_ulonglong * (8192 * 4))()
    fn = _native.lib().mvb200_debug_read_prefill_ctas
    fn.argtypes = [ctypes.c_void_p]
    assert fn(buf) == 0
    c = np.frombuffer(buf, dtype=np.uint64).astype(np.int64).reshape(8192, 4)[:num_ctas]
    smid = c[:, 0] & 0xFFFFFFFF
    nkt = c[:, 0] >> 32
    live = nkt > 0
    t0 = c[live, 1].min()
    start, first, end = c[:, 1] - t0, c[:, 2] - t0, c[:, 3] - t0
    dur = (end - start)[live]
    span = end[live].max()
    print(f"CTAs {live.sum()}  kernel span {span/1e3:.1f} us  sum(CTA time) {dur.sum()/1e3:.0f} us  avg concurrency {dur.sum()/span:.1f} (296 slots)")
    A = np.stack([np.ones(live.sum()), nkt[live]], 1).astype(np.float64)
    coef = np.linalg.lstsq(A, dur.astype(np.float64), rcond=None)[0]
    print(f"CTA duration fit: {coef[0]:.0f} ns fixed + {coef[1]:.0f} ns per 64-key tile;  first S after {np.median((first-start)[live]):.0f} ns (median)")
    for q in (18, 24, 32):
        sel = live & (nkt == q)
        if sel.any():
            print(f"  n_kt={q}: {sel.sum()} CTAs, duration median {np.median((end-start)[sel]):.0f} ns  min {(end-start)[sel].min()}  max {(end-start)[sel].max()}")
    # per-SM idle: time with fewer than 2 resident CTAs
    last_start = start[live].max()
    print(f"last CTA starts at {last_start/1e3:.1f} us; tail after that = {(span-last_start)/1e3:.1f} us")
    busy = np.zeros(int(smid.max()) + 1)
    for sm in range(len(busy)):
        sel = live & (smid == sm)
        busy[sm] = (end - start)[sel].sum()
    print(f"per-SM sum of CTA time / (2*span): min {busy.min()/(2*span):.2f} median {np.median(busy)/(2*span):.2f} max {busy.max()/(2*span):.2f}")
    # launch gaps: on each SM, sort by start; gap between an end and the next start
    order = np.argsort(start)
    print("first 6 CTA starts (ns):", start[order][:6], " ends of earliest:", end[order][:3])


if __name__ == "__main__":
    main()
