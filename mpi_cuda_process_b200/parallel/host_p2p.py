"""Host-side halo exchange over gloo (CPU tensors).

Two message shapes:
  * ``row``     one message per ghost row block (what a sane MPI code does);
  * ``element`` the reference's pattern: columns 1..w-2 of the border row as ``w-2`` separate blocking
    one-element messages, tag 1 (MDF_kernel.cu:167-169,181-183).
Used by the CPU plumbing path (BASELINE.json config 1) and by the host-staged GPU baseline.  Unlike the
reference (rank 1 sends to and receives from itself, MDF_kernel.cu:201,215 -> deadlock for g >= 2) the
neighbour table comes from the decomposition and both directions are mirrored.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist


def exchange_rows(u: torch.Tensor, ghost: int, own: int, up: Optional[int], down: Optional[int], group=None,
                  mode: str = "row") -> int:
    """Fill the ghost rows of ``u`` (layout ``[ghost | own | ghost]`` along axis 0) from the neighbours.

    Returns the number of messages this rank sent.  ``u`` must be a CPU tensor.
    """
    assert u.device.type == "cpu"
    sent = 0
    if mode == "row":
        ops, bufs = [], []
        if up is not None:
            sb = u[ghost:2 * ghost].contiguous()
            rb = torch.empty_like(sb)
            ops += [dist.P2POp(dist.isend, sb, up, group), dist.P2POp(dist.irecv, rb, up, group)]
            bufs.append((rb, slice(0, ghost)))
            sent += 1
        if down is not None:
            sb = u[own:own + ghost].contiguous()
            rb = torch.empty_like(sb)
            ops += [dist.P2POp(dist.isend, sb, down, group), dist.P2POp(dist.irecv, rb, down, group)]
            bufs.append((rb, slice(ghost + own, ghost + own + ghost)))
            sent += 1
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
            for rb, sl in bufs:
                u[sl].copy_(rb)
        return sent
    if mode != "element":
        raise ValueError(mode)
    if u.dim() != 2:
        raise ValueError("element mode is the reference's 2-D pattern")
    w = u.shape[1]
    # Deadlock-free ordering of blocking one-element messages along the open chain: even ranks send first.
    rank = dist.get_rank(group)

    def send_rows(dst, rows):
        nonlocal sent
        for r in rows:
            row = u[r]
            for j in range(1, w - 1):
                dist.send(row[j:j + 1].clone(), dst, group=group, tag=1)
                sent += 1

    def recv_rows(src, rows):
        for r in rows:
            row = u[r]
            for j in range(1, w - 1):
                t = torch.empty(1, dtype=u.dtype)
                dist.recv(t, src, group=group, tag=1)
                row[j] = t[0]

    up_send = range(ghost, 2 * ghost)
    up_recv = range(0, ghost)
    dn_send = range(own, own + ghost)
    dn_recv = range(ghost + own, own + 2 * ghost)
    if rank % 2 == 0:
        if down is not None:
            send_rows(down, dn_send); recv_rows(down, dn_recv)
        if up is not None:
            send_rows(up, up_send); recv_rows(up, up_recv)
    else:
        if up is not None:
            recv_rows(up, up_recv); send_rows(up, up_send)
        if down is not None:
            recv_rows(down, dn_recv); send_rows(down, dn_send)
    return sent


def exchange_axis(u: torch.Tensor, axis: int, ghost: int, own: int, lo: Optional[int], hi: Optional[int], group=None) -> int:
    """Generic ghost exchange along ``axis`` (used by the 3-D pencil CPU path)."""
    if lo is None and hi is None:
        return 0
    v = u.movedim(axis, 0)
    ops, bufs, sent = [], [], 0
    if lo is not None:
        sb = v[ghost:2 * ghost].contiguous(); rb = torch.empty_like(sb)
        ops += [dist.P2POp(dist.isend, sb, lo, group), dist.P2POp(dist.irecv, rb, lo, group)]
        bufs.append((rb, slice(0, ghost))); sent += 1
    if hi is not None:
        sb = v[own:own + ghost].contiguous(); rb = torch.empty_like(sb)
        ops += [dist.P2POp(dist.isend, sb, hi, group), dist.P2POp(dist.irecv, rb, hi, group)]
        bufs.append((rb, slice(ghost + own, ghost + own + ghost))); sent += 1
    for w in dist.batch_isend_irecv(ops):
        w.wait()
    for rb, sl in bufs:
        v[sl].copy_(rb)
    return sent
