"""One-shot all-reduce over NVLink peer memory, fused with the residual add + RMSNorm that follow it.

Caller-side fusion of SURVEY.md 8(f) rank 3: replaces ``dist.all_reduce`` after the row-parallel
``o_proj`` / ``down_proj`` (reference models/layers/linear.py:314-319) plus the next
``RMSNorm.forward(x, residual)`` (models/layers/layernorm.py:81-105) for decode-size messages.  The kernel
and its contract are in ``csrc/peer_allreduce.cu`` / ``include/mvb200.h``; this module only owns the
buffers and the one-time handle exchange (``torch.distributed`` is the plumbing, CUDA IPC the transport).
Larger messages (prefill) stay on NCCL: check ``fits()``.
"""

from __future__ import annotations

import ctypes

import torch

from . import _native
from .ops import _code, _require_cuda, _stream


class PeerSetupError(RuntimeError):
    """Raised on EVERY rank of the group when the peer mapping could not be set up on at least one of them."""


class PeerAllReduce:
    """Per-rank endpoint.  Build one per process with ``PeerAllReduce(rank, world, max_bytes, group=...)``
    (handles travel through ``all_gather_object``) or ``PeerAllReduce.local_group`` for ``world`` endpoints
    inside one process (tests on a single GPU: the kernel only needs peer *pointers*)."""

    def __init__(self, rank: int, world: int, max_bytes: int, *, group=None, device=None, max_ctas: int = 0,
                 _defer: bool = False):
        lib = _native.lib()
        if not 1 <= world <= 8:
            raise ValueError(f"PeerAllReduce supports 1..8 ranks, got {world}")
        self.rank, self.world = rank, world
        self.max_ctas = int(max_ctas)  # 0: library default; identical on every rank
        self.device = torch.device(device if device is not None else ("cuda", torch.cuda.current_device()))
        self.slot_bytes = (int(max_bytes) + 255) // 256 * 256
        self._lib = lib
        self._opened: list[int] = []
        self._data = self._flags = self._epoch = None
        self.data_ptrs = (ctypes.c_void_p * 8)()
        self.flag_ptrs = (ctypes.c_void_p * 8)()
        if _defer or world == 1:
            self._alloc_own()
            return
        # Several processes: success must be agreed on.  A rank that fails alone (no P2P to one peer, IPC limits,
        # out of memory) must still take part in every collective below, or its peers wait for it forever; and the
        # peers must learn of it, or they would spin in the kernel's flag wait while it has gone back to NCCL.
        # So: nothing that can fail sits between the collectives unguarded, and the outcome is all-gathered.
        import torch.distributed as dist

        error: Exception | None = None
        mine = None
        try:
            self._alloc_own()
            mine = (self._export(self._data), self._export(self._flags))
        except Exception as e:  # noqa: BLE001  reported collectively below
            error = e
        everyone: list = [None] * world
        dist.all_gather_object(everyone, mine, group=group)
        if error is None and all(h is not None for h in everyone):
            try:
                with torch.cuda.device(self.device):
                    for r, (hd, hf) in enumerate(everyone):
                        if r != rank:
                            self.data_ptrs[r], self.flag_ptrs[r] = self._open(hd), self._open(hf)
            except Exception as e:  # noqa: BLE001
                error = e
        # doubles as the barrier: nobody launches before every rank has mapped every peer
        outcome: list = [None] * world
        dist.all_gather_object(outcome, None if error is None else f"{type(error).__name__}: {error}", group=group)
        bad = {r: msg for r, msg in enumerate(outcome) if msg is not None}
        if bad:
            self.close()
            raise PeerSetupError(f"peer all-reduce unavailable, set-up failed on rank(s) {sorted(bad)}: "
                                 + "; ".join(f"[{r}] {m}" for r, m in sorted(bad.items())))

    def _alloc_own(self) -> None:
        lib = self._lib
        with torch.cuda.device(self.device):
            self._data = self._alloc(2 * (self.world + 2) * self.slot_bytes)
            self._flags = self._alloc(lib.mvb200_peer_flag_bytes())
            self._epoch = self._alloc(lib.mvb200_peer_epoch_bytes())
        self.data_ptrs[self.rank], self.flag_ptrs[self.rank] = self._data, self._flags

    # ---- construction helpers -------------------------------------------------------------------
    @classmethod
    def local_group(cls, world: int, max_bytes: int, device=None, max_ctas: int = 32) -> list["PeerAllReduce"]:
        """``world`` endpoints on one device.  Their kernels wait for each other, so all of their CTAs must
        be co-resident on that one GPU: ``max_ctas`` caps the grid (32 x 8 ranks fits a B200 several times)."""
        eps = [cls(r, world, max_bytes, device=device, max_ctas=max_ctas, _defer=True) for r in range(world)]
        for a in eps:
            for b in eps:
                a.data_ptrs[b.rank], a.flag_ptrs[b.rank] = b._data, b._flags
        return eps

    def _alloc(self, nbytes: int) -> int:
        p = ctypes.c_void_p()
        _native.check(self._lib.mvb200_peer_alloc(nbytes, ctypes.byref(p)), "mvb200_peer_alloc")
        return p.value

    def _export(self, ptr: int) -> bytes:
        buf = ctypes.create_string_buffer(64)
        _native.check(self._lib.mvb200_peer_export(ptr, buf), "mvb200_peer_export")
        return buf.raw

    def _open(self, handle: bytes) -> int:
        p = ctypes.c_void_p()
        _native.check(self._lib.mvb200_peer_open(handle, ctypes.byref(p)), "mvb200_peer_open")
        self._opened.append(p.value)
        return p.value

    def close(self) -> None:
        for p in self._opened:
            self._lib.mvb200_peer_close(p)
        self._opened = []
        for name in ("_data", "_flags", "_epoch"):
            p = getattr(self, name, None)
            if p:
                self._lib.mvb200_peer_free(p)
                setattr(self, name, None)

    # ---- the op ---------------------------------------------------------------------------------
    def fits(self, x: torch.Tensor) -> bool:
        row = x.size(-1) * x.element_size()
        return x.dim() == 2 and x.numel() * x.element_size() <= self.slot_bytes and row % 16 == 0 and row <= 16384

    #: messages up to this many bytes use the one-shot kernel, larger ones the two-shot kernel
    #: (tests/mp_peer_allreduce_check.py prints both; 8 x B200: equal at 512 KiB (19.4 vs 19.6 us), two-shot ahead
    #: from there on: 4 MiB 62 vs 35 us, 8 MiB 115 vs 52 us; NCCL all_reduce + norm kernel: 33 / 60 / 79 us)
    ONE_SHOT_MAX_BYTES = 1024 * 1024

    def allreduce_add_rms_norm(self, x: torch.Tensor, weight: torch.Tensor | None, eps: float,
                               residual: torch.Tensor | None = None, out: torch.Tensor | None = None, *,
                               algo: int | None = None, gather_residual: bool = True):
        """``y, residual`` of ``RMSNorm(all_reduce(x), residual)``; ``residual`` is updated in place.
        ``weight=None``: no normalisation (``y = all_reduce(x) [+ residual]``).
        ``algo``: 1 one-shot, 2 two-shot, ``None`` picks by size.  ``gather_residual=False`` (two-shot only)
        leaves the residual row-sharded: this rank then updates only the rows it owns, which is all a
        decoder needs when every consumer of the residual is one of these calls.  Every rank must make the
        same choice."""
        _require_cuda(x, weight, residual, out)
        if x.dim() != 2 or x.stride(1) != 1:
            raise ValueError(f"peer all-reduce expects a [rows, hidden] tensor with dense rows, got {tuple(x.shape)}")
        rows, hidden = x.shape
        if not self.fits(x):
            raise ValueError(f"message of {x.numel() * x.element_size()} bytes does not fit the {self.slot_bytes}-byte slots "
                             "(or the row is not a multiple of 16 bytes / exceeds 16 KiB): use NCCL for this size")
        for name, t in (("weight", weight), ("residual", residual), ("out", out)):
            if t is not None and t.dtype != x.dtype:
                raise ValueError(f"{name} dtype {t.dtype} != x dtype {x.dtype}")
        if weight is not None and (weight.numel() != hidden or not weight.is_contiguous()):
            raise ValueError("weight must be a contiguous [hidden] vector")
        if residual is not None and (residual.shape != x.shape or residual.stride(1) != 1):
            raise ValueError("residual must match x and have dense rows")
        if out is None:
            out = torch.empty((rows, hidden), dtype=x.dtype, device=x.device)
        elif out.shape != x.shape or out.stride(1) != 1:
            raise ValueError("out must match x and have dense rows")
        if algo is None:
            algo = 1 if x.numel() * x.element_size() <= self.ONE_SHOT_MAX_BYTES or self.world == 1 else 2
        rs = lambda t: t.stride(0) if rows > 1 else hidden  # noqa: E731
        rc = self._lib.mvb200_peer_allreduce_add_rms_norm(
            self.data_ptrs, self.flag_ptrs, self._epoch, self.slot_bytes, self.rank, self.world,
            x.data_ptr(), rs(x), residual.data_ptr() if residual is not None else None,
            rs(residual) if residual is not None else 0, weight.data_ptr() if weight is not None else None,
            out.data_ptr(), rs(out), rows, hidden, float(eps), _code(x.dtype), self.max_ctas,
            int(algo), 1 if gather_residual else 0, _stream(),
        )
        _native.check(rc, "mvb200_peer_allreduce_add_rms_norm")
        return out, residual

    def all_reduce(self, x: torch.Tensor, out: torch.Tensor | None = None, algo: int | None = None) -> torch.Tensor:
        return self.allreduce_add_rms_norm(x, None, 0.0, None, out, algo=algo)[0]


class NvlsAllReduce(PeerAllReduce):
    """``PeerAllReduce`` whose large messages go through the NVSwitch (NVLS / multimem): the owner of a row reads
    it with ``multimem.ld_reduce`` — the switch sums the ranks' copies — and broadcasts the finished row with one
    ``multimem.st`` (``csrc/peer_allreduce.cu::nvls_allreduce_kernel``).  The symmetric buffer and its multicast
    mapping come from ``torch.distributed._symmetric_memory`` (plumbing); flags and epochs are the parent's.
    Needs the default NCCL process group (one process per GPU) and a fabric with multicast support."""

    def __init__(self, rank: int, world: int, max_bytes: int, *, group=None, device=None, max_ctas: int = 0):
        super().__init__(rank, world, max_bytes, group=group, device=device, max_ctas=max_ctas)
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm

        if world < 2:
            raise ValueError("NvlsAllReduce needs at least two ranks")
        self._sym = symm.empty(6 * self.slot_bytes, dtype=torch.uint8, device=self.device)
        self._sym.zero_()
        hdl = symm.rendezvous(self._sym, dist.group.WORLD.group_name)
        self._sym_handle = hdl
        self.multicast_ptr = int(getattr(hdl, "multicast_ptr", 0) or 0)
        if not self.multicast_ptr:
            raise RuntimeError("this fabric / driver gives no multicast mapping (NVLS unavailable): use PeerAllReduce")
        torch.cuda.synchronize(self.device)
        dist.barrier()

    #: ``algo=None`` keeps the parent's choice (one-shot / two-shot peer stores): measured on 8 x B200
    #: (profiles/r02_peer_allreduce_tp8_run26.txt) the NVLS kernel does not beat the two-shot one yet — 8 MiB:
    #: 69 / 56 us (gathered / sharded residual) against 64 / 52 us; the operator is bound by its four flag
    #: round trips and the copy in / out of the symmetric buffer, not by NVLink bytes.  Ask for it with ``algo=3``.
    NVLS_MIN_BYTES = None

    def allreduce_add_rms_norm(self, x, weight, eps, residual=None, out=None, *, algo=None, gather_residual=True):
        """As ``PeerAllReduce.allreduce_add_rms_norm``; ``algo=3`` (or ``None`` for large messages) = NVLS."""
        nbytes = x.numel() * x.element_size()
        if algo is None and self.NVLS_MIN_BYTES is not None and nbytes > self.NVLS_MIN_BYTES:
            algo = 3
        if algo != 3:
            return super().allreduce_add_rms_norm(x, weight, eps, residual, out, algo=algo, gather_residual=gather_residual)
        _require_cuda(x, weight, residual, out)
        if x.dim() != 2 or x.stride(1) != 1 or not self.fits(x):
            raise ValueError(f"NVLS all-reduce expects a [rows, hidden] tensor with dense rows that fits the {self.slot_bytes}-byte slots")
        rows, hidden = x.shape
        for name, t in (("weight", weight), ("residual", residual), ("out", out)):
            if t is not None and t.dtype != x.dtype:
                raise ValueError(f"{name} dtype {t.dtype} != x dtype {x.dtype}")
        if out is None:
            out = torch.empty((rows, hidden), dtype=x.dtype, device=x.device)
        rs = lambda t: t.stride(0) if rows > 1 else hidden  # noqa: E731
        rc = self._lib.mvb200_nvls_allreduce_add_rms_norm(
            self.multicast_ptr, self._sym.data_ptr(), self.slot_bytes, self.flag_ptrs, self._epoch, self.rank, self.world,
            x.data_ptr(), rs(x), residual.data_ptr() if residual is not None else None,
            rs(residual) if residual is not None else 0, weight.data_ptr() if weight is not None else None,
            out.data_ptr(), rs(out), rows, hidden, float(eps), _code(x.dtype), self.max_ctas,
            1 if gather_residual else 0, _stream(),
        )
        _native.check(rc, "mvb200_nvls_allreduce_add_rms_norm")
        return out, residual
