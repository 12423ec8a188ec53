"""Decomposition, process bring-up and the communication layers (fused P2P arena, NCCL and host baselines)."""
from .decomp import Block, Decomposition, choose_pencil_grid, split_axis  # noqa: F401
from .context import DistContext, init_context  # noqa: F401
