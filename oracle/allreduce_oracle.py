"""CPU oracle for the all-reduce + residual add + RMSNorm sequence (TEST INFRASTRUCTURE — imported only by
tests/, __graft_entry__.smoke() and bench.py's CPU legs; never by the product path).

Restates, with torch CPU operators in the reference's own order and rounding points:

* ``RowParallelLinear.forward``: ``dist.all_reduce(y)``            (models/layers/linear.py:314-319)
  -> SUM over ranks, result in the model dtype.  Here: fp32 accumulation in rank order, one rounding.
  For world 2 this is exactly what any implementation returns (T(a+b)); for world > 2 NCCL/gloo round
  the partial sums to the model dtype along the way, so they can differ from this by a few units in the
  last place of the *sum* (and from each other) — tests state that tolerance.
* ``RMSNorm.forward(x, residual)``                                  (models/layers/layernorm.py:81-131)
  ``x32 = x.float() + residual.float(); residual = x32.to(T);
    y = (x32 * rsqrt(mean(x32^2) + eps) * weight).to(T)``;  without residual ``x32 = x.float()``.

Pinned against tests/golden/arnorm.npz, recorded from the reference's own ``RMSNorm`` module applied to an
in-dtype accumulation over ranks (tests/golden/gen_golden.py arnorm) — tests/test_peer_allreduce.py.
"""

from __future__ import annotations

import torch


def all_reduce_sum(xs: list[torch.Tensor]) -> torch.Tensor:
    acc = torch.zeros_like(xs[0], dtype=torch.float32)
    for x in xs:
        acc = acc + x.float()
    return acc.to(xs[0].dtype)


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: torch.Tensor | None = None):
    """(y, new_residual) as RMSNorm.forward; new_residual is None without a residual."""
    orig = x.dtype
    x32 = x.float()
    new_residual = None
    if residual is not None:
        x32 = x32 + residual.float()
        new_residual = x32.to(orig)
    var = x32.pow(2).mean(dim=-1, keepdim=True)
    y = (x32 * torch.rsqrt(var + eps) * weight).to(orig)
    return y, new_residual


def allreduce_add_rms_norm(xs: list[torch.Tensor], weight: torch.Tensor | None, eps: float,
                           residual: torch.Tensor | None = None):
    """What every rank holds after ``dist.all_reduce`` + ``RMSNorm.forward`` (``weight=None``: the plain
    all-reduce, plus the residual when given)."""
    total = all_reduce_sum(xs)
    if weight is None:
        if residual is None:
            return total, None
        r = (total.float() + residual.float()).to(total.dtype)
        return r, r
    return rms_norm(total, weight, eps, residual)
