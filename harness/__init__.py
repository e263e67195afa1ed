"""Test/bench harness that hosts the hot path end to end on a box without the reference.

NOT product code: plain-torch Qwen2/Qwen3 decoder blocks (cuBLAS GEMMs via F.linear, eager
RMSNorm/RoPE/SiLU, NCCL all-reduce) around ``minivllm_b200.Attention`` + a minimal prefill-then-decode
engine with the reference's block/slot layout and prefix hashing.  The scheduler, sampler and model
definitions of the reference are out of scope; this exists so that end-to-end tokens/s
and greedy token-stream parity can be measured on the GPU box.
"""
