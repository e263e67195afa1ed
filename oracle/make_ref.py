#!/usr/bin/env python
"""Recipe: place the UNMODIFIED reference package under the git-ignored ``oracle/_ref/``.

The reference (jianzhnie/mini-vllm) is pure Python: there is nothing to compile.  The
GPU box has no ``/root/reference``; ``oracle/_ref/`` is listed in ``.gitignore`` (reference
sources never enter the history) but not in ``.gpurunignore``, so this copy travels with the
snapshot and lets the GPU tests and ``bench.py`` run the reference's own engine
(``LLM.generate``), its ``PageAttention`` and its example flow on the box.

Run from ``__graft_entry__.build()`` whenever ``/root/reference`` exists; a no-op elsewhere.
What is taken: ``minivllm/**/*.py``, ``examples/cpu_inference_opt.py`` (BASELINE config 1's
script) and the reference's own hot-path unit tests (``tests/test_page_attention.py``,
``tests/test_npu_backend.py``).  A manifest with per-file SHA-256 is written beside them.
"""

from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
EXTRA = ["examples/cpu_inference_opt.py", "tests/test_page_attention.py", "tests/test_npu_backend.py", "LICENSE"]


def make(source: str | None = None, verbose: bool = False) -> str | None:
    source = source or os.environ.get("MINIVLLM_REFERENCE", "/root/reference")
    if not os.path.isdir(os.path.join(source, "minivllm")):
        return DEST if os.path.isdir(os.path.join(DEST, "minivllm")) else None
    files = []
    for base, dirs, names in os.walk(os.path.join(source, "minivllm")):
        dirs[:] = [d for d in dirs if d != "__pycache__"]
        files += [os.path.relpath(os.path.join(base, n), source) for n in names if n.endswith(".py")]
    files += [f for f in EXTRA if os.path.exists(os.path.join(source, f))]
    if os.path.isdir(DEST):
        shutil.rmtree(DEST)
    manifest = {}
    for rel in sorted(files):
        dst = os.path.join(DEST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(source, rel), dst)
        with open(dst, "rb") as f:
            manifest[rel] = hashlib.sha256(f.read()).hexdigest()
    with open(os.path.join(DEST, "MANIFEST.json"), "w") as f:
        json.dump({"source": source, "files": manifest}, f, indent=1)
    if verbose:
        print(f"oracle/_ref: {len(manifest)} files from {source}")
    return DEST


if __name__ == "__main__":
    make(sys.argv[1] if len(sys.argv) > 1 else None, verbose=True)
