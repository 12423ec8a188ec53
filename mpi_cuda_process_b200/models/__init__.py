"""Workload families ("models") of the framework.

The reference ships two programs that differ only in cell type and update rule (``diff kernel.cu
MDF_kernel.cu``): the 2-D 5-point MDF heat/Jacobi update on floats and Conway's Game of Life on ints.
Each workload here is described by a :class:`Workload` record (rule, dtypes, traffic model, oracle, operator code
and time-block limits of the fused kernel, boundary rule); ``Solver`` and the CLI dispatch through the registry.  The
3-D 7-point heat equation is added as the extension BASELINE.json names.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, Tuple

from .. import oracle


@dataclass(frozen=True)
class Workload:
    name: str
    ndim: int
    dtypes: Tuple[str, ...]              # storage types ("bits" = bit-packed life cells)
    stencil_points: int
    flops_per_update: int
    reference: str                      # file:line of the rule in the reference
    init: Callable
    step: Callable
    description: str
    op_code: int = -1                   # MCP_OP_* of the fused 2-D kernel (-1: not a 2-D operator)
    max_time_block: Tuple[Tuple[str, int], ...] = ()   # (dtype, largest register-resident time block) pairs

    def time_block_limit(self, dtype: str) -> int:
        """Largest number of generations the fused kernel advances per pass over the field for this dtype."""
        return dict(self.max_time_block).get(dtype, 1)

    def evolving_box(self, cfg) -> Tuple[int, int, int, int]:
        """Inclusive global box (ylo, yhi, xlo, xhi) of the cells that evolve; cells outside keep their value."""
        h, w = cfg.height, cfg.width
        if self.name == "life" and cfg.bc == "dead-outside":
            return 0, h - 1, 0, w - 1
        return 1, h - 2, 1, w - 2

    def bytes_per_update(self, dtype: str) -> int:
        """Compulsory HBM traffic of one update: one read + one write of the cell."""
        return 2 * (8 if dtype == "fp64" else 4)


from .mdf2d import MDF2D  # noqa: E402
from .life2d import LIFE2D  # noqa: E402
from .heat3d import HEAT3D  # noqa: E402

REGISTRY: Dict[str, Workload] = {w.name: w for w in (MDF2D, LIFE2D, HEAT3D)}


def get(name: str) -> Workload:
    try:
        return REGISTRY[name]
    except KeyError:
        raise ValueError(f"unknown workload {name!r}; have {sorted(REGISTRY)}") from None
