"""Process/device bring-up and teardown.

Reference: ``MPI_Init / MPI_Comm_size / MPI_Comm_rank / MPI_Get_processor_name`` (MDF_kernel.cu:128-131),
four ``cudaStreamCreate`` calls issued *before* MPI_Init with no ``cudaSetDevice`` anywhere, so every
rank lands on GPU 0 (MDF_kernel.cu:114-121, defect D2), and ``MPI_Finalize`` (MDF_kernel.cu:234).
Here: one process per GPU, ``LOCAL_RANK -> cuda:N`` bound first, ``torch.distributed`` for the plumbing
(NCCL on GPUs for bootstrap/baselines, gloo for host-side control and the CPU path).
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist


@dataclass
class DistContext:
    rank: int = 0
    world: int = 1
    local_rank: int = 0
    device: torch.device = torch.device("cpu")
    owns_process_group: bool = False
    _cpu_group: Optional[object] = None

    @property
    def is_cuda(self) -> bool:
        return self.device.type == "cuda"

    @property
    def distributed(self) -> bool:
        return self.world > 1

    # -- host-side control plane -----------------------------------------------------------
    def cpu_group(self):
        """A gloo group for host objects / CPU tensors (lazily created next to an NCCL default group)."""
        if not self.distributed:
            return None
        if self._cpu_group is None:
            if dist.get_backend() == "gloo":
                self._cpu_group = dist.group.WORLD
            else:
                self._cpu_group = dist.new_group(backend="gloo")
        return self._cpu_group

    def barrier(self) -> None:
        if self.distributed:
            if self.is_cuda:
                torch.cuda.synchronize(self.device)
            dist.barrier(group=self.cpu_group())

    def all_gather_object(self, obj):
        if not self.distributed:
            return [obj]
        out = [None] * self.world
        dist.all_gather_object(out, obj, group=self.cpu_group())
        return out

    def broadcast_object(self, obj, src: int = 0):
        if not self.distributed:
            return obj
        box = [obj]
        dist.broadcast_object_list(box, src=src, group=self.cpu_group())
        return box[0]

    def all_reduce_max(self, value: float) -> float:
        if not self.distributed:
            return value
        t = torch.tensor([value], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.cpu_group())
        return float(t[0])

    def close(self) -> None:
        if self.owns_process_group and dist.is_initialized():
            try:
                dist.destroy_process_group()
            except Exception:
                pass
            self.owns_process_group = False


def init_context(device: Optional[str] = None, backend: Optional[str] = None) -> DistContext:
    """Create the context from the torchrun environment (or single-process defaults).

    ``device``: "cuda", "cpu" or None (cuda if available).  Binds the CUDA device *before* anything
    else touches the GPU.
    """
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    use_cuda = (device or ("cuda" if torch.cuda.is_available() else "cpu")).startswith("cuda")
    if use_cuda and not torch.cuda.is_available():
        raise RuntimeError("CUDA requested but no GPU is visible")
    if use_cuda:
        ndev = torch.cuda.device_count()
        if local_rank >= ndev and os.environ.get("MCP_ALLOW_GPU_SHARING") != "1":
            # Two ranks time-slicing one GPU cannot run the fused kernels: they are cooperative launches that fill
            # every SM and spin on each other's flags (they would stall until the flag timeout).  One rank per GPU.
            raise RuntimeError(f"LOCAL_RANK {local_rank} but only {ndev} GPU(s) visible: one rank per GPU is required "
                               "(set MCP_ALLOW_GPU_SHARING=1 for the unfused baselines)")
        dev = torch.device("cuda", local_rank % max(ndev, 1))
        torch.cuda.set_device(dev)
    else:
        dev = torch.device("cpu")
    owns = False
    if dist.is_initialized():
        world, rank = dist.get_world_size(), dist.get_rank()
    elif world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        be = backend or ("nccl" if use_cuda else "gloo")
        kwargs = {}
        if be == "nccl":
            kwargs["device_id"] = dev
        dist.init_process_group(backend=be, rank=rank, world_size=world,
                                timeout=datetime.timedelta(seconds=600), **kwargs)
        owns = True
    return DistContext(rank=rank, world=world, local_rank=local_rank, device=dev, owns_process_group=owns)
