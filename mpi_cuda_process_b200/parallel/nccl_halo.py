"""NCCL halo exchange (baseline): grouped ncclSend/ncclRecv of the ghost rows once per generation.

This is the "NCCL-halo variant" BASELINE.json asks to be beaten: partitioned storage, a plain
(unfused) stencil kernel, and the exchange as a separate library call between kernels.  It keeps the
reference's structure (compute, then exchange the border, MDF_kernel.cu:162-183) with the host staging
and per-element messages removed.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist


def exchange_axis(u: torch.Tensor, axis: int, ghost: int, own: int, lo: Optional[int], hi: Optional[int], group=None) -> int:
    """Fill ghost layers of ``u`` along ``axis`` (layout ``[ghost | own | ghost]``).  CUDA tensors, NCCL group."""
    if lo is None and hi is None:
        return 0
    v = u.movedim(axis, 0)
    ops, post, sent = [], [], 0
    if lo is not None:
        sb = v[ghost:2 * ghost]
        rb = v[0:ghost]
        if not sb.is_contiguous():
            sb = sb.contiguous(); tmp = torch.empty_like(sb); post.append((rb, tmp)); rb = tmp
        ops += [dist.P2POp(dist.isend, sb, lo, group), dist.P2POp(dist.irecv, rb, lo, group)]
        sent += 1
    if hi is not None:
        sb = v[own:own + ghost]
        rb = v[ghost + own:ghost + own + ghost]
        if not sb.is_contiguous():
            sb = sb.contiguous(); tmp = torch.empty_like(sb); post.append((rb, tmp)); rb = tmp
        ops += [dist.P2POp(dist.isend, sb, hi, group), dist.P2POp(dist.irecv, rb, hi, group)]
        sent += 1
    for w in dist.batch_isend_irecv(ops):
        w.wait()
    for dst, src in post:
        dst.copy_(src)
    return sent
