"""The drop-in run for real: the reference's own ``LLM.generate`` with the sm_100a layer swapped in.

The reference package is imported from ``/root/reference`` (build container) or from its copy under the
git-ignored ``oracle/_ref/`` (what travels to the GPU box; ``oracle/make_ref.py``).  Everything from
``LLM`` down to the sampler is the reference's code; ``minivllm_b200.install_into_reference()`` replaces the
``Attention`` layer, and with ``executor=True`` the executor's metadata preparation / CUDA-Graph capture.

Golden streams (``tests/golden/engine_named.npz``, ``gen_golden.py::gen_engine_named``) are the reference's
eager CPU engine on the NAMED architectures (Qwen3-0.6B dims, 28 layers, 151 936-token vocabulary; OPT-125M
dims = BASELINE config 1) with seeded random weights, 32 greedy tokens per prompt.  Bars:

* float32: token-for-token equal.  The only tolerated difference is a proven near-tie — the reference's own
  top-2 logits at that step differ by < 2e-4 and the GPU picked the other of those two candidates (GPU and
  CPU GEMMs sum in a different order); the number of such steps is printed and must stay <= 1 per case.
* bfloat16: same rule with the near-tie bound at 4 bf16 ulps of the logit (0.0625); the stream is compared up
  to the first tolerated divergence (free-running generation cannot be compared after it).
* call B's prompts share 2-3 leading blocks with call A's: on the one GPU engine they HIT the prefix cache
  (asserted) and must still produce the streams the reference gives them on a fresh engine.
"""

import json
import os
import subprocess
import sys

import numpy as np
import pytest

from tests.conftest import ROOT, load_golden

CHILD = os.path.join(ROOT, "tests", "ref_engine_child.py")


def _reference_available() -> bool:
    from oracle.reference_engine import locate_reference

    return locate_reference() is not None


def _run_child(spec: dict, timeout: int = 900) -> dict:
    env = dict(os.environ)
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    r = subprocess.run([sys.executable, CHILD, json.dumps(spec)], capture_output=True, text=True, timeout=timeout, env=env)
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT ")]
    assert r.returncode == 0 and lines, f"child failed (rc {r.returncode}):\n{r.stdout[-3000:]}\n{r.stderr[-6000:]}"
    return json.loads(lines[-1][7:])


def _compare(tag: str, info: dict, tie_gap: float) -> dict:
    g = load_golden("engine_named")
    report = {"tag": tag, "tokens": 0, "equal": 0, "near_ties": 0, "compared_after_tie": 0, "max_top1_diff": 0.0}
    for call, res in zip(("a", "b"), info["calls"]):
        n = int(g[f"{tag}.{call}.n"])
        assert len(res["tokens"]) == n
        for i in range(n):
            want = g[f"{tag}.{call}.tokens{i}"].tolist()
            top2 = g[f"{tag}.{call}.top2_{i}"]
            got = res["tokens"][i]
            assert len(got) == len(want) == 32
            for t, (a, b) in enumerate(zip(got, want)):
                report["tokens"] += 1
                if a == b:
                    report["equal"] += 1
                    report["max_top1_diff"] = max(report["max_top1_diff"], abs(res["top2"][i][t][0] - float(top2[t, 0])))
                    continue
                gap = float(top2[t, 0] - top2[t, 1])
                # The reference itself samples through multinomial(softmax(logits / 1e-8))
                # (sampling/functional.py:63-68,434-436): on a tie it may take ANY of the tied candidates.  The golden
                # records its two top ones; a tie can be three-way or wider (verdict r2, 1c: the reference re-run picked
                # an id outside its own recorded top-2), so a GPU token outside the recorded pair is accepted when
                # this run's winning logit sits within the tie gap of the reference's maximum — it belongs to the
                # same tie set.
                in_recorded_pair = a in (int(top2[t, 2]), int(top2[t, 3]))
                in_tie_set = abs(float(res["top2"][i][t][0]) - float(top2[t, 0])) <= tie_gap
                assert gap <= tie_gap and (in_recorded_pair or in_tie_set), (
                    f"{tag} call {call} prompt {i} step {t}: got {a}, reference {b} "
                    f"(reference top-2 gap {gap:.3e}, runner-up {int(top2[t, 3])}, this run's top logit "
                    f"{float(res['top2'][i][t][0]):.4f} vs {float(top2[t, 0]):.4f})\n got  {got}\n want {want}")
                report["near_ties"] += 1
                report["ties_outside_recorded_pair"] = report.get("ties_outside_recorded_pair", 0) + (not in_recorded_pair)
                report["compared_after_tie"] += len(want) - t - 1
                break  # free-running streams legitimately differ after a tie
    return report


# ------------------------------------------------------------------------------------------- CPU


def test_golden_engine_named_is_complete():
    g = load_golden("engine_named")
    assert set(g["cases"].tolist()) >= {"qwen3-0.6b.float32", "qwen3-0.6b-untied.float32",
                                        "qwen3-0.6b-untied.bfloat16", "opt-125m.float32"}
    for tag in g["cases"].tolist():
        for call in ("a", "b"):
            for i in range(int(g[f"{tag}.{call}.n"])):
                assert g[f"{tag}.{call}.tokens{i}"].shape == (32,)
                assert g[f"{tag}.{call}.top2_{i}"].shape == (32, 4)
        # call B really shares whole blocks with call A (what makes it a prefix hit on a warm engine)
        assert g[f"{tag}.b.prompt0"][:128].tolist() == g[f"{tag}.a.prompt0"][:128].tolist()
        assert g[f"{tag}.b.prompt1"][:192].tolist() == g[f"{tag}.a.prompt2"][:192].tolist()


@pytest.mark.skipif(not os.path.isdir("/root/reference/minivllm"), reason="needs /root/reference (build container)")
def test_make_ref_copy_is_the_unmodified_reference(tmp_path):
    """``oracle/make_ref.py`` output == the reference's files, byte for byte, and imports from the copy."""
    import hashlib

    from oracle import make_ref

    dest = make_ref.make()
    manifest = json.load(open(os.path.join(dest, "MANIFEST.json")))
    assert len(manifest["files"]) > 40
    for rel, digest in manifest["files"].items():
        with open(os.path.join("/root/reference", rel), "rb") as f:
            assert hashlib.sha256(f.read()).hexdigest() == digest, rel
    # the copy (not /root/reference) drives a tiny engine on CPU and reproduces round 1's recorded stream
    code = (
        "import os, sys, json; sys.path.insert(0, %r); os.environ['MINIVLLM_REFERENCE'] = %r\n"
        "import numpy as np\n"
        "from harness import ref_engine\n"
        "z = np.load(%r)\n"
        "n = int(z['tiny-qwen3.n']); prompts = [z[f'tiny-qwen3.prompt{i}'].tolist() for i in range(n)]\n"
        "info = ref_engine.run('tiny-qwen3', device='cpu', dtype='float32', calls=[prompts], max_num_seqs=8,\n"
        "                      max_model_len=1024, num_kvcache_blocks=64)\n"
        "assert info['reference_root'] == %r, info['reference_root']\n"
        "assert info['calls'][0]['tokens'] == [z[f'tiny-qwen3.tokens{i}'].tolist() for i in range(n)]\n"
        "print('COPY_OK')\n"
    ) % (ROOT, dest, os.path.join(ROOT, "tests", "golden", "engine.npz"), dest)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert "COPY_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.skipif(not _reference_available(), reason="needs the reference (or oracle/_ref)")
def test_executor_patch_keeps_reference_streams_on_cpu():
    """``reference_patch.install_executor()`` alone (metadata preparation rebinding, reference attention on
    CPU): the reference engine must produce exactly its recorded OPT-125M streams, prefix-free call A."""
    info = _run_child({"tag": "opt-125m.float32", "preset": "opt-125m", "dtype": "float32", "device": "cpu",
                       "dropin": False, "executor": True, "hide_flash_attn": True})
    g = load_golden("engine_named")
    n = int(g["opt-125m.float32.a.n"])
    assert info["calls"][0]["tokens"] == [g[f"opt-125m.float32.a.tokens{i}"].tolist() for i in range(n)]
    assert info["attention_classes"] == ["minivllm.models.layers.attention.Attention"]


# ------------------------------------------------------------------------------------------- GPU

CASES = [
    # tag, preset, dtype, overrides, near-tie bound on the reference's top-2 logit gap
    ("qwen3-0.6b.float32", "qwen3-0.6b", "float32", {}, 2e-4),
    ("qwen3-0.6b-untied.float32", "qwen3-0.6b", "float32", {"tie_word_embeddings": False}, 2e-4),
    ("qwen3-0.6b-untied.bfloat16", "qwen3-0.6b", "bfloat16", {"tie_word_embeddings": False}, 0.0625),
    ("opt-125m.float32", "opt-125m", "float32", {}, 2e-4),
]


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["layer_eager", "executor_graphs", "executor_graphs_fused_qk"])
@pytest.mark.parametrize("tag,preset,dtype,overrides,tie_gap", CASES, ids=[c[0] for c in CASES])
def test_reference_llm_generate_with_b200_layer(cuda_device, tag, preset, dtype, overrides, tie_gap, mode):
    if not _reference_available():
        pytest.fail("oracle/_ref is missing: __graft_entry__.build() must run oracle/make_ref.py before the snapshot")
    graphs = mode != "layer_eager"
    fuse_qk = mode.endswith("fused_qk")  # QwenAttention.forward rebound to the one-kernel q/k norm + RoPE + store
    if fuse_qk and not preset.startswith("qwen"):
        pytest.skip("the fused q/k path rebinds QwenAttention only")
    info = _run_child({"tag": tag, "preset": preset, "dtype": dtype, "overrides": overrides,
                       "executor": graphs, "graphs": graphs, "fuse_qk": fuse_qk})
    # the reference's engine really ran our layer and our kernels
    assert info["attention_classes"] == ["minivllm_b200.attention.Attention"], info["attention_classes"]
    seen = set(info["kernels_seen"])
    assert any(k.startswith("decode") for k in seen) and any(k.startswith("prefill") for k in seen) \
        and any(k.startswith("store") or k.startswith("qk_norm_rope") for k in seen), seen
    if fuse_qk:
        assert any(k.startswith("qk_norm_rope") for k in seen), seen
    if graphs:
        assert info["graphs_captured"] == [1, 2, 4, 8], info["graphs_captured"]
    else:
        assert info["graphs_captured"] == []
    assert info["calls"][0]["cache_hits"] == 0
    assert info["calls"][1]["cache_hits"] >= 5, "call B must hit the prefix cache of call A"  # 2 + 3 blocks
    rep = _compare(tag, info, tie_gap)
    print(f"\n[reference-engine parity] {mode}: {json.dumps(rep)} kernels={sorted(seen)}")
    # float32: at most one proven tie per case.  bfloat16: logits of magnitude ~3 have an ulp of 0.0156, the
    # reference's own CPU stream contains exact top-2 ties; every divergence has been proven to be one above
    # (both candidates are the reference's), at most one per prompt since a free-running stream ends there.
    assert rep["near_ties"] <= (1 if dtype == "float32" else 7), rep
    if dtype == "float32":
        assert rep["max_top1_diff"] < 2e-3, rep


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["qwen3-0.6b-untied.bfloat16", "qwen3-0.6b-untied.float32"])
def test_reference_engine_tp2_streams_on_two_gpus(cuda_device, tag, tmp_path):
    """TP = 2 on two GPUs (verdict r2, missing 1): the reference engine + drop-in with KV heads sharded per rank, graphs
    captured on every rank and the decode-size all-reduces on the peer kernel must reproduce the streams the reference
    recorded at TP = 1 (``tools/tp_parity.py``; weights via the TP-aware ``seeded_load``).  Skipped on a one-GPU box."""
    import socket

    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    if not _reference_available():
        pytest.fail("oracle/_ref is missing: __graft_entry__.build() must run oracle/make_ref.py before the snapshot")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = tmp_path / "tp_parity.json"
    env = dict(os.environ)
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tools", "tp_parity.py"), "--tag", tag, "--out", str(out)],
                       capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0 and out.exists(), f"torchrun rc {r.returncode}:\n{r.stdout[-3000:]}\n{r.stderr[-6000:]}"
    rec = json.loads(out.read_text())
    print(f"\n[reference-engine parity, TP 2] {json.dumps(rec)}")
    assert rec["ok"], rec.get("error")
    assert rec["attention_classes"] == ["minivllm_b200.attention.Attention"]
    assert rec["kv_heads_per_rank"] == 4 and rec["graphs_captured"] == [1, 2, 4, 8]
    assert rec["allreduce"]["collective"] == "mvb200_peer_one_shot" and rec["allreduce"]["peer_calls"] > 0, rec["allreduce"]
    assert rec["cache_hits"][1] >= 5
    assert rec["parity"]["near_ties"] <= (1 if tag.endswith("float32") else 7), rec["parity"]
