#!/usr/bin/env python
"""Opcode counts per kernel of minivllm_b200/libmvb200.so (cuobjdump -sass): the evidence that the prefill kernels are
tcgen05 / TMEM / TMA code (UTCHMMA, LDTM/STTM, UTMALDG, UTCBAR), that decode is TMA + mma.sync (UTMALDG, HMMA, LDSM)
and that the NVLS all-reduce issues multimem instructions.  Writes profiles/r02_sass_summary.txt.

    python tools/sass_summary.py [path/to/lib.so]
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "minivllm_b200", "libmvb200.so")
WATCH = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMAPF", "UBLKCP", "SYNCS", "HMMA", "LDSM", "MUFU.EX2", "FFMA2", "FADD2",
         "F2FP", "USETMAXREG", "ELECT", "LDGMC", "STGMC", "REDG", "MEMBAR", "ACQBULK", "LDL", "STL", "BAR.SYNC", "ATOMG", "NANOSLEEP"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for ln in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", ln)
        if m and cur:
            op = m.group(1)
            kernels[cur]["_total"] += 1
            for w in WATCH:
                if op.startswith(w):
                    kernels[cur][w] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    lines = [f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)} — opcode counts per kernel (instructions, not executions)", ""]
    for (name, c), pretty in zip(kernels.items(), demangle):
        short = re.sub(r"\(.*", "", pretty)
        watched = "  ".join(f"{w}={c[w]}" for w in WATCH if c[w])
        lines.append(f"{short}\n    instructions={c['_total']}  {watched}")
    text = "\n".join(lines) + "\n"
    dst = os.path.join(ROOT, "profiles", "r02_sass_summary.txt")
    open(dst, "w").write(text)
    print(text[:3000])


if __name__ == "__main__":
    main()
