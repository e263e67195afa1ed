"""CPU oracle for the device-side decode bookkeeping (TEST INFRASTRUCTURE — never imported by the
product path; only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use it).

Restates, in numpy, what the reference computes on the host every decode step:

* ``prepare_decode``      InferenceExecutor._prepare_decode_input      (engine/inference_executor.py:551-617)
                          + _prepare_block_tables_decode               (:651-688)
* ``may_append``          BlockManager.may_append, the allocation half (engine/block_manager.py, len % BS == 1)
* ``greedy``              Sampler temperature == 0 / GreedySampler      (sampling/sampler.py:104-106, :118-130)

Pinned: ``prepare_decode`` reproduces the tensors recorded from the unmodified reference executor in
tests/golden/metadata.npz (steps d0..d3, including a block-boundary crossing) — tests/test_decode_meta.py.
``greedy`` is checked against ``torch.argmax`` on CPU, the call the reference makes.
"""

from __future__ import annotations

import numpy as np


def prepare_decode(seqs, block_size: int = 64, width: int | None = None) -> dict:
    """One row per sequence: last token, position len-1, context_len len,
    slot = block_table[-1]*BS + last_block_num_tokens - 1 (0 without a table, :563-574); block table padded
    with -1 to the longest table (:672-679) or to ``width`` columns."""
    b = len(seqs)
    ids = np.zeros(b, np.int64)
    pos = np.zeros(b, np.int64)
    ctx = np.zeros(b, np.int32)
    slots = np.zeros(b, np.int32)
    mb = max((len(s.block_table) for s in seqs), default=0)
    if width is not None:
        mb = width
    table = np.full((b, mb), -1, np.int32)
    for i, s in enumerate(seqs):
        n = len(s.token_ids)
        ids[i] = s.token_ids[-1]
        pos[i] = n - 1
        ctx[i] = n
        if s.block_table:
            last_block_num_tokens = n - (len(s.block_table) - 1) * block_size  # sequence.py last_block_num_tokens
            slots[i] = s.block_table[-1] * block_size + last_block_num_tokens - 1
            table[i, : len(s.block_table)] = s.block_table
    return {"input_ids": ids, "positions": pos, "context_lens": ctx, "slot_mapping": slots, "block_tables": table}


class HostSeq:
    """The three fields of the reference's Sequence the decode prep reads (engine/sequence.py:31)."""

    def __init__(self, token_ids, block_table):
        self.token_ids = [int(t) for t in token_ids]
        self.block_table = [int(b) for b in block_table]

    def __len__(self):
        return len(self.token_ids)


def may_append(seq: HostSeq, free_blocks: list, block_size: int = 64):
    """Allocation half of BlockManager.may_append: a sequence whose newest token is the first of a block
    (len % BS == 1) takes the next free block.  Returns the block id or None."""
    if len(seq) % block_size == 1:
        bid = free_blocks.pop(0)
        seq.block_table.append(bid)
        return bid
    return None


def greedy(logits: np.ndarray) -> np.ndarray:
    """argmax over the last axis, first occurrence on ties, NaN as the maximum (torch.argmax)."""
    x = np.asarray(logits, dtype=np.float32)
    out = np.empty(x.shape[0], np.int64)
    for r in range(x.shape[0]):
        row = x[r]
        nan = np.flatnonzero(np.isnan(row))
        out[r] = nan[0] if nan.size else int(np.argmax(row))
    return out
