"""Build libmvb200.so (sm_100a) in-tree with nvcc.  No torch headers, no JIT cache."""

from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libmvb200.so")
SOURCES = [
    "host_util.cu",
    "store_kvcache.cu",
    "qk_rope_store.cu",
    "decode_meta.cu",
    "peer_allreduce.cu",
    "paged_decode.cu",
    "prefill_generic.cu",
    "prefill_fa2q.cu",
    "prefill_dbs.cu",
    "prefill_api.cu",
]
# tuning build only (--trace): the first-generation prefill kernel, kept under probes/ (not product code) for A/B runs
TUNING_SOURCES = [os.path.join(os.path.dirname(HERE), "probes", "prefill_tcgen05.cu")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=hidden",
    "-cudart", "shared",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(os.path.dirname(HERE), "include", "mvb200.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, trace: bool = False) -> str:
    """Compile every .cu for sm_100a and link the shared library.  Returns its path.

    ``trace=True`` builds libmvb200_trace.so from the same sources with -DPF_TRACE (the prefill kernel
    records a per-tile clock64 timeline; used by tools/prefill_trace.py, never by the product path)."""
    if trace:
        return _build(os.path.join(HERE, "libmvb200_trace.so"), os.path.join(HERE, "build_trace"),
                      ["-DPF_TRACE", "-DMVB200_TUNING"], verbose, SOURCES + TUNING_SOURCES)
    if not force and not needs_build():
        return LIB
    return _build(LIB, os.path.join(HERE, "build"), [], verbose, SOURCES)


def _build(LIB: str, objdir: str, extra: list, verbose: bool, sources: list) -> str:
    os.makedirs(objdir, exist_ok=True)
    nvcc = _nvcc()
    procs = []
    objs = []
    for src in sources:
        obj = os.path.join(objdir, os.path.basename(src).replace(".cu", ".o"))
        objs.append(obj)
        cmd = [nvcc, *NVCC_FLAGS, *extra, "-I", CSRC, "-c", os.path.join(CSRC, src), "-o", obj]  # join keeps absolute paths
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            failed = True
            sys.stderr.write(f"--- nvcc failed on {src} ---\n{out}\n")
        elif verbose and out:
            sys.stderr.write(f"--- {src} ---\n{out}\n")
    if failed:
        raise RuntimeError("nvcc compilation failed")
    link = [nvcc, "-shared", "-cudart", "shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB, *objs]
    r = subprocess.run(link, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, trace="--trace" in sys.argv))
