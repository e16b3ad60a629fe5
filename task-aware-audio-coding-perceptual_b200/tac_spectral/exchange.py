"""Peer-mapped exchange buffers for the one-shot NVLink sum of a step (``include/tac_spectral.h``, "multi-GPU exchange").

SURVEY.md 8e prescribes one packed all-reduce of ``[loss sums | per view n, sum z, sum z z^T]`` per step for the
batch-sharded form of ``train_pkl_file_task_aware.py:597-629`` / ``pkl_file_model_updated.py:489-511``.  Here that
exchange is peer stores + flags issued by the kernels that produce the payload (``tac_dpca_stats_push``,
``tac_loss_finalize``) and spun on by the kernels that consume it (``tac_dpca_eig_pop``): no communication-library
kernel sits between them.  This module only does the plumbing: one zero-filled buffer per rank that every process of
the node has mapped, through ``torch.distributed._symmetric_memory`` (default) or CUDA IPC (``tac_ipc_alloc``).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Sequence

import torch

from . import _lib as L


class PeerBuffer:
    """``nbytes`` of zero-filled device memory per rank; ``ptrs[r]`` = rank r's buffer as mapped in this process."""

    def __init__(self, nbytes: int, group, device, backend: str | None = None):
        import torch.distributed as dist
        self.group, self.device, self.nbytes = group, torch.device(device), int(nbytes)
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > L.XCHG_MAX_WORLD:
            raise ValueError(f"the peer exchange supports up to {L.XCHG_MAX_WORLD} ranks of one node, got {self.world}")
        backend = (backend or os.environ.get("TAC_XCHG_BACKEND", "auto")).lower()
        if backend not in ("auto", "symm", "ipc"):
            raise ValueError(f"unknown exchange backend {backend!r} (auto | symm | ipc)")
        self.ptrs: List[int] = []
        self._keep = None
        self._ipc_local = None
        self._ipc_opened: List[int] = []
        self.backend = None
        if backend in ("auto", "symm"):
            err = None
            try:
                self._init_symm()
            except Exception as e:                                   # every rank must take the same decision: vote below
                err = e
            ok = torch.tensor([0 if err is not None else 1], dtype=torch.int32, device=self.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok.item()) == 1:
                self.backend = "symm"
            elif backend == "symm":
                raise RuntimeError(f"symmetric-memory rendezvous failed on at least one rank ({err!r})")
            else:
                self._keep, self.ptrs = None, []
        if self.backend is None:
            self._init_ipc()
            self.backend = "ipc"
        torch.cuda.synchronize(self.device)
        dist.barrier(group=group, device_ids=[self.device.index] if dist.get_backend(group) == "nccl" else None)

    def _init_symm(self):
        import torch.distributed._symmetric_memory as symm_mem
        with torch.cuda.device(self.device):
            t = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=self.device)
            name = getattr(self.group, "group_name", None)
            try:
                hdl = symm_mem.rendezvous(t, self.group)
            except Exception:
                if name is None:
                    raise
                symm_mem.enable_symm_mem_for_group(name)             # older releases want the group registered first
                hdl = symm_mem.rendezvous(t, name)
            t.zero_()
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        if len(ptrs) != self.world or ptrs[self.rank] != t.data_ptr():
            raise RuntimeError("unexpected symmetric-memory handle layout")
        self._keep, self.ptrs = (t, hdl), ptrs

    def _init_ipc(self):
        """CUDA IPC mapping.  Allocation and the opens can fail on ONE rank only (no peer access to some GPU, a rank on
        another node): every rank still takes part in the collectives, then all vote, and all raise together."""
        import torch.distributed as dist
        lib = L.lib()
        err = None
        with torch.cuda.device(self.device):
            p = C.c_void_p()
            handle = (C.c_ubyte * 64)()
            try:
                L.check(lib.tac_ipc_alloc(self.nbytes, C.byref(p), handle))
                self._ipc_local = p.value
                mine = bytes(handle)
            except RuntimeError as e:
                err, mine = e, b""
            handles: List[bytes] = [b""] * self.world
            dist.all_gather_object(handles, mine, group=self.group)
            ptrs = []
            for r, h in enumerate(handles):
                if r == self.rank:
                    ptrs.append(self._ipc_local)
                    continue
                if err is not None or len(h) != 64:
                    err = err or RuntimeError(f"rank {r} could not allocate its exchange buffer")
                    continue
                q = C.c_void_p()
                try:
                    L.check(lib.tac_ipc_open((C.c_ubyte * 64).from_buffer_copy(h), C.byref(q)))
                except RuntimeError as e:
                    err = e
                    continue
                self._ipc_opened.append(q.value)
                ptrs.append(q.value)
            ok = torch.tensor([0 if err is not None else 1], dtype=torch.int32, device=self.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
            if int(ok.item()) != 1:
                self.close()
                raise RuntimeError(f"CUDA IPC mapping of the exchange buffers failed on at least one rank ({err!r})")
            self.ptrs = ptrs

    def close(self):
        """Unmap / free (IPC backend); the symmetric-memory tensor goes with its last reference."""
        lib = L.lib()
        with torch.cuda.device(self.device):
            for q in self._ipc_opened:
                lib.tac_ipc_close(q)
            self._ipc_opened = []
            if self._ipc_local is not None:
                lib.tac_ipc_free(self._ipc_local)
                self._ipc_local = None
        self._keep = None


class Exchange:
    """Named channels (``{"stats": payload, "sums": payload}``, payload in float64 values) carved out of one
    :class:`PeerBuffer`; ``desc(name)`` is the ``tac_xchg`` descriptor the C ABI takes (host struct, by pointer)."""

    def __init__(self, channels: Dict[str, int], group, device, backend: str | None = None):
        import torch.distributed as dist
        world = dist.get_world_size(group)
        lib = L.lib()
        self.offsets: Dict[str, int] = {}
        off = 0
        for name, payload in channels.items():
            n = int(lib.tac_xchg_bytes(world, int(payload)))
            if n == 0:
                raise ValueError(f"bad exchange channel {name!r}: world {world}, payload {payload}")
            self.offsets[name] = off
            off += (n + 511) // 512 * 512
        self.buf = PeerBuffer(off, group, device, backend)
        self.world, self.rank, self.backend = self.buf.world, self.buf.rank, self.buf.backend
        self._desc = {}
        for name, payload in channels.items():
            d = L.XchgStruct()
            d.world, d.rank, d.payload = self.world, self.rank, int(payload)
            for r in range(self.world):
                d.peer[r] = self.buf.ptrs[r] + self.offsets[name]
            self._desc[name] = d

    def desc(self, name: str):
        return C.byref(self._desc[name])

    def check(self) -> None:
        """Raise if a pop of any channel gave up waiting for a peer (``tac_xchg_status``).  Synchronises the device.

        The kernels never trap on a missing peer (that would kill the CUDA context of this rank, and of the others in
        turn); they record the failure and the host reports it here, the way NCCL's watchdog does for an all-reduce.
        The bound is ``TAC_XCHG_TIMEOUT_S`` (default 600 s, 0 = unbounded) or :func:`set_timeout`."""
        torch.cuda.synchronize(self.buf.device)
        lib = L.lib()
        with torch.cuda.device(self.buf.device):
            for name, d in self._desc.items():
                w = C.c_uint(0)
                L.check(lib.tac_xchg_status(C.byref(d), C.byref(w)))
                if w.value:
                    raise RuntimeError(f"peer exchange channel {name!r}: pop #{w.value - 1} on rank {self.rank} timed out waiting "
                                       f"for a peer's push; results since then are undefined (raise TAC_XCHG_TIMEOUT_S or "
                                       f"barrier the ranks after long rank-local pauses)")

    def close(self):
        self.buf.close()


def set_timeout(seconds: float) -> None:
    """Wait bound of the pops enqueued from now on (``tac_xchg_set_timeout``): 0 = unbounded, negative = default."""
    L.check(L.lib().tac_xchg_set_timeout(float(seconds)))


def push(x, src: torch.Tensor) -> None:
    """Generic push of a float64 device tensor (``tac_xchg_push``) on the current stream."""
    with L.device_of(src):
        L.check(L.lib().tac_xchg_push(x, src.data_ptr(), L.stream_of(src)))


def pop(x, dst: torch.Tensor) -> None:
    """Generic pop (wait for all ranks' pushes, rank-ordered float64 sum) into ``dst`` on the current stream."""
    with L.device_of(dst):
        L.check(L.lib().tac_xchg_pop(x, dst.data_ptr(), L.stream_of(dst)))
