"""Peer-visible blocks for the fused gradient reduce + Adam + parameter all-gather of the sharded mapping step
(include/monogs_b200.h, mgs_peer_*).

Every rank owns one block  [parameter arena | gradient arena | loss_out f32[4] | flags]  allocated by the library
(cudaMalloc, so that a CUDA IPC handle covers exactly the block) and maps the blocks of the other ranks of the node
through CUDA IPC; the 64-byte handles travel once through torch.distributed (all_gather_object), nothing else does.
`PeerBlock.local_group` builds the same structure for several "ranks" inside one process on one GPU — the protocol
(flags, slices, rank-ordered sums) is identical, which is how the single-GPU tests exercise it.
"""
import ctypes as C
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

from . import _lib as L


class _Raw:
    """A device range exposed through __cuda_array_interface__ so torch can wrap it without owning it."""

    def __init__(self, ptr: int, n: int, typestr: str, owner):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}
        self._owner = owner


class PeerBlock:
    """backend="ipc": cudaMalloc + CUDA IPC (unicast peer mappings; the TMA kernel).  backend="symm": the block is a
    torch.distributed._symmetric_memory allocation (CUDA VMM), whose rendezvous returns the peer mappings AND an NVSwitch
    multicast mapping of the same memory: mgs_peer_reduce_adam then reduces inside the switch (multimem.ld_reduce) and
    broadcasts the parameters with multimem.st."""

    def __init__(self, numel: int, dev: torch.device, backend: str = "ipc", side_floats: int = 4):
        assert numel > 0 and numel % 4 == 0, "arena length must be a positive multiple of 4 floats"
        assert backend in ("ipc", "symm")
        self.lib = L.lib()
        self.numel, self.dev, self.backend = numel, dev, backend
        self.flag_bytes = int(self.lib.mgs_peer_flags_bytes())
        assert side_floats >= 4 and side_floats % 4 == 0
        self.side = side_floats        # floats of loss_out: the sums of the arena's trailing sum-only segment
        self.nbytes = 2 * numel * 4 + 4 * side_floats + self.flag_bytes
        self._opened: List[int] = []
        self.mc_base = 0
        if backend == "symm":
            import torch.distributed._symmetric_memory as symm_mem
            self._symm = symm_mem.empty(self.nbytes // 4, dtype=torch.float32, device=dev)
            self._symm.zero_()
            torch.cuda.synchronize(dev)
            self.ptr = self._symm.data_ptr()
            self.params, self.grads = self._symm[:numel], self._symm[numel:2 * numel]
            self.loss_out = self._symm[2 * numel:2 * numel + side_floats]
            self.flags = self._symm[2 * numel + side_floats:].view(torch.int32)
        else:
            p = C.c_void_p()
            with torch.cuda.device(dev):
                L.check(self.lib.mgs_peer_alloc(self.nbytes, C.byref(p)), "mgs_peer_alloc")
            self.ptr = p.value
            self.params = torch.as_tensor(_Raw(self.ptr, numel, "<f4", self), device=dev)
            self.grads = torch.as_tensor(_Raw(self.ptr + 4 * numel, numel, "<f4", self), device=dev)
            self.loss_out = torch.as_tensor(_Raw(self.ptr + 8 * numel, side_floats, "<f4", self), device=dev)
            self.flags = torch.as_tensor(_Raw(self.ptr + 8 * numel + 4 * side_floats, self.flag_bytes // 4, "<i4", self), device=dev)
        self.world, self.rank = 1, 0
        self.bases = [self.ptr]
        self._group, self._connected = None, False

    # ------------------------------------------------------------------------------------------------
    def handle(self) -> bytes:
        buf = C.create_string_buffer(L.PEER_HANDLE_BYTES)
        L.check(self.lib.mgs_peer_export(C.c_void_p(self.ptr), buf), "mgs_peer_export")
        return buf.raw

    def connect(self, group=None):
        """Exchange the handles over `group` and map every other rank's block (collective)."""
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        self._group, self._connected = group, True
        assert world <= L.PEER_MAX, f"at most {L.PEER_MAX} ranks share peer blocks (one NVLink domain)"
        if self.backend == "symm":
            import torch.distributed._symmetric_memory as symm_mem
            h = symm_mem.rendezvous(self._symm, group=(group if group is not None else dist.group.WORLD).group_name)
            if not h.multicast_ptr:
                raise RuntimeError("symmetric memory gave no multicast mapping (no NVSwitch multicast on this box): use "
                                   "exchange=\"peer\"")
            self._handle = h
            self.world, self.rank, self.bases, self.mc_base = world, rank, [int(p) for p in h.buffer_ptrs], int(h.multicast_ptr)
            assert self.bases[rank] == self.ptr
            dist.barrier(group)
            return self
        handles: List[Optional[bytes]] = [None] * world
        dist.all_gather_object(handles, (self.handle(), self.numel), group=group)
        bases = []
        for r, (h, n) in enumerate(handles):
            assert n == self.numel, "every rank must hold the same arena layout"
            if r == rank:
                bases.append(self.ptr)
                continue
            p = C.c_void_p()
            with torch.cuda.device(self.dev):
                L.check(self.lib.mgs_peer_open(C.create_string_buffer(h, L.PEER_HANDLE_BYTES), C.byref(p)), "mgs_peer_open")
            self._opened.append(p.value)
            bases.append(p.value)
        self.world, self.rank, self.bases = world, rank, bases
        dist.barrier(group)      # nobody signals a block before every rank has mapped it
        return self

    @staticmethod
    def local_group(numel: int, world: int, dev: torch.device, side_floats: int = 4) -> List["PeerBlock"]:
        """`world` blocks in THIS process on one device, wired to each other by their plain pointers."""
        blocks = [PeerBlock(numel, dev, side_floats=side_floats) for _ in range(world)]
        for r, b in enumerate(blocks):
            b.world, b.rank, b.bases = world, r, [x.ptr for x in blocks]
        return blocks

    # ------------------------------------------------------------------------------------------------
    def exchange(self, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor, segment_ends: Sequence[int],
                 segment_lrs: Sequence[float], timeout_ms: int = 0) -> "L.MgsPeerExchange":
        x = L.MgsPeerExchange()
        x.timeout_ms = timeout_ms
        x.world, x.rank, x.numel = self.world, self.rank, self.numel
        for r, b in enumerate(self.bases):
            x.param_arena[r], x.grad_arena[r] = b, b + 4 * self.numel
            x.loss_out[r], x.flags[r] = b + 8 * self.numel, b + 8 * self.numel + 4 * self.side
        assert exp_avg.numel() == self.numel and exp_avg_sq.numel() == self.numel
        x.exp_avg, x.exp_avg_sq = exp_avg.data_ptr(), exp_avg_sq.data_ptr()
        assert len(segment_ends) == len(segment_lrs) <= L.PEER_MAX_SEGMENTS
        if self.mc_base:
            x.param_multicast, x.grad_multicast = self.mc_base, self.mc_base + 4 * self.numel
            x.loss_multicast = self.mc_base + 8 * self.numel
        x.n_segments = len(segment_ends)
        for i, (e, lr) in enumerate(zip(segment_ends, segment_lrs)):
            x.segment_end[i], x.segment_lr[i] = int(e), float(lr)
        return x

    def slice_of(self, r: int):
        """[begin, end) in floats of rank r's slice (the same arithmetic as mgs_peer_reduce_adam)."""
        n4 = self.numel // 4
        per = (n4 + self.world - 1) // self.world
        b = min(per * r, n4)
        return 4 * b, 4 * min(b + per, n4)

    def error_epoch(self, stream_ptr) -> int:
        e = C.c_uint32(0)
        L.check(self.lib.mgs_peer_error(C.c_void_p(self.flags.data_ptr()), C.byref(e), stream_ptr), "mgs_peer_error")
        return e.value

    def close(self):
        """Drain this device, unmap the other ranks' blocks, then free this one.  Collective for a connected block: an
        exporter must not free its allocation while another process still has it mapped, so the ranks meet between the two
        steps.  The synchronize comes FIRST: a peer_reduce_adam / peer_wait kernel still queued on this rank reads and
        writes the peers' mappings."""
        if self.backend == "symm":
            if getattr(self, "ptr", None):
                torch.cuda.synchronize(self.dev)
            self.ptr = None       # the symmetric allocation is released with its tensors
            return
        if getattr(self, "ptr", None):
            with torch.cuda.device(self.dev):
                torch.cuda.synchronize(self.dev)
                for p in self._opened:
                    self.lib.mgs_peer_close(C.c_void_p(p))
                self._opened = []
                if self._connected:
                    dist.barrier(self._group)
                self.lib.mgs_peer_free(C.c_void_p(self.ptr))
            self.ptr = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False

    def __del__(self):
        # Last resort for a block that was never closed: release the local allocation and mappings WITHOUT the collective
        # barrier (a finalizer can not rendezvous).  A connected block should be closed explicitly, on every rank.
        try:
            if getattr(self, "backend", None) == "ipc" and getattr(self, "ptr", None):
                with torch.cuda.device(self.dev):
                    torch.cuda.synchronize(self.dev)
                    for p in self._opened:
                        self.lib.mgs_peer_close(C.c_void_p(p))
                    self.lib.mgs_peer_free(C.c_void_p(self.ptr))
                self.ptr = None
        except Exception:   # interpreter shutdown: the driver reclaims everything
            pass
