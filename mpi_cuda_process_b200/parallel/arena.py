"""Symmetric device arena: the same allocation on every rank, mapped into every peer process.

This is the B200 replacement for the reference's communication layer.  The reference moves the
halo through host memory: full-grid D2H copy, ``w-2`` blocking one-element ``MPI_Send`` calls, the
mirror-image ``MPI_Recv`` loop, and a full-grid H2D copy (MDF_kernel.cu:166-183).  Here every rank
``cudaMalloc``s one arena, exports it with CUDA IPC and maps its peers' arenas, so a kernel holds
plain device pointers into the neighbour GPUs and stores halo rows over NVLink/NVSwitch itself.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence

import torch

from ..ops import _lib

_TORCH_TYPESTR = {
    torch.float32: "<f4", torch.float64: "<f8", torch.int32: "<i4", torch.int64: "<i8",
    torch.uint8: "|u1", torch.int8: "|i1",
}


class _CudaArrayHolder:
    def __init__(self, ptr: int, shape, typestr: str, owner=None):
        self.__cuda_array_interface__ = {
            "shape": tuple(int(s) for s in shape), "typestr": typestr, "data": (int(ptr), False),
            "version": 3, "strides": None,
        }
        self._owner = owner


def tensor_from_ptr(ptr: int, shape: Sequence[int], dtype: torch.dtype, device: torch.device, owner=None) -> torch.Tensor:
    """Zero-copy torch view of raw device memory (kept alive by ``owner``)."""
    holder = _CudaArrayHolder(ptr, shape, _TORCH_TYPESTR[dtype], owner)
    t = torch.as_tensor(holder, device=device)
    t._mcp_owner = holder  # keep the arena alive as long as the view
    return t


class SymmetricArena:
    """One zero-filled device allocation per rank + peer mappings of the same allocation.

    ``peer_ptr(r)`` is the address of rank ``r``'s arena in *this* process (own base for ``r == rank``).
    """

    def __init__(self, ctx, nbytes: int, peers: Optional[Sequence[int]] = None):
        self.ctx = ctx
        self.nbytes = int(nbytes)
        self.lib = _lib.cuda_lib()
        base = C.c_void_p()
        _lib.check(self.lib.mcp_arena_alloc(self.nbytes, C.byref(base)), "mcp_arena_alloc")
        self.base = int(base.value)
        self._peer_ptrs: Dict[int, int] = {ctx.rank: self.base}
        self._opened: List[int] = []
        self._closed = False
        if ctx.distributed:
            hsz = self.lib.mcp_ipc_handle_size()
            hbuf = (C.c_ubyte * hsz)()
            _lib.check(self.lib.mcp_ipc_get_handle(base, hbuf), "mcp_ipc_get_handle")
            handles = ctx.all_gather_object(bytes(hbuf))
            want = set(range(ctx.world)) if peers is None else set(peers)
            want.discard(ctx.rank)
            for r in sorted(want):
                hb = (C.c_ubyte * hsz).from_buffer_copy(handles[r])
                p = C.c_void_p()
                _lib.check(self.lib.mcp_ipc_open_handle(hb, C.byref(p)), f"mcp_ipc_open_handle(rank {r})")
                self._peer_ptrs[r] = int(p.value)
                self._opened.append(int(p.value))
            ctx.barrier()

    def peer_ptr(self, rank: int, offset: int = 0) -> int:
        return self._peer_ptrs[rank] + offset

    def has_peer(self, rank: int) -> bool:
        return rank in self._peer_ptrs

    def view(self, offset: int, shape, dtype: torch.dtype) -> torch.Tensor:
        return tensor_from_ptr(self.base + offset, shape, dtype, self.ctx.device, owner=self)

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        if self.ctx.distributed:
            try:
                self.ctx.barrier()  # nobody may still be storing into our arena
            except Exception:
                pass
        for p in self._opened:
            self.lib.mcp_ipc_close_handle(C.c_void_p(p))
        self._opened.clear()
        self.lib.mcp_arena_free(C.c_void_p(self.base))

    def __del__(self):
        try:
            if not self._closed and not self.ctx.distributed:
                self.close()
        except Exception:
            pass


class SharedBuffer:
    """A device buffer owned by ONE rank and mapped into every other rank (CUDA IPC): the destination of the fused
    gather + pack, where every rank's pack kernel stores its block straight into rank 0's output over NVLink.
    The reference has no gather at all (every rank keeps a replica of the whole grid, MDF_kernel.cu:137-144)."""

    def __init__(self, ctx, nbytes: int, owner: int = 0):
        self.ctx, self.nbytes, self.owner = ctx, int(nbytes), owner
        self.lib = _lib.cuda_lib()
        self.ptr = 0
        self._mapped = False
        hsz = self.lib.mcp_ipc_handle_size()
        handle = None
        if ctx.rank == owner:
            base = C.c_void_p()
            _lib.check(self.lib.mcp_arena_alloc(self.nbytes, C.byref(base)), "mcp_arena_alloc")
            self.ptr = int(base.value)
            if ctx.distributed:
                hbuf = (C.c_ubyte * hsz)()
                _lib.check(self.lib.mcp_ipc_get_handle(base, hbuf), "mcp_ipc_get_handle")
                handle = bytes(hbuf)
        if ctx.distributed:
            handle = ctx.broadcast_object(handle, owner)
            if ctx.rank != owner:
                hb = (C.c_ubyte * hsz).from_buffer_copy(handle)
                p = C.c_void_p()
                _lib.check(self.lib.mcp_ipc_open_handle(hb, C.byref(p)), "mcp_ipc_open_handle")
                self.ptr = int(p.value)
                self._mapped = True
            ctx.barrier()

    def view(self, shape, dtype: torch.dtype) -> torch.Tensor:
        assert self.ctx.rank == self.owner
        return tensor_from_ptr(self.ptr, shape, dtype, self.ctx.device, owner=self)

    def close(self) -> None:
        if not self.ptr:
            return
        if self.ctx.distributed:
            self.ctx.barrier()
        if self._mapped:
            self.lib.mcp_ipc_close_handle(C.c_void_p(self.ptr))
        elif self.ctx.rank == self.owner:
            self.lib.mcp_arena_free(C.c_void_p(self.ptr))
        self.ptr = 0
