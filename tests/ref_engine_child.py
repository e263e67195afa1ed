"""Child process of tests/test_reference_engine.py: one reference ``LLM`` with the sm_100a layer swapped in.

Runs call A then call B of a golden case on ONE engine (B hits the prefix cache of A) and prints a JSON
line with the token streams, the top-2 logits of every sampled row and the kernel families launched.
Kept out of the pytest process because importing the reference registers a ``colorama`` stand-in, hides
``flash_attn`` and rebinds ``Attention`` in the reference's modules.
"""

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> None:
    spec = json.loads(sys.argv[1])
    import numpy as np

    from harness import ref_engine

    z = np.load(os.path.join(ROOT, "tests", "golden", spec.get("golden", "engine_named.npz")))
    tag = spec["tag"]
    calls = []
    for call in ("a", "b"):
        n = int(z[f"{tag}.{call}.n"])
        calls.append([z[f"{tag}.{call}.prompt{i}"].tolist() for i in range(n)])
    info = ref_engine.run(
        spec["preset"], device=spec.get("device", "cuda"), dtype=spec["dtype"], dropin=spec.get("dropin", True),
        executor=spec.get("executor", False), enforce_eager=not spec.get("graphs", False), calls=calls,
        max_tokens=32, record_logits=True, config_overrides=spec.get("overrides") or {},
        hide_flash_attn=spec.get("hide_flash_attn"), fuse_qk=spec.get("fuse_qk", False),
        max_num_seqs=8, max_model_len=1024, num_kvcache_blocks=64,
    )
    print("RESULT " + json.dumps(info))


if __name__ == "__main__":
    main()
