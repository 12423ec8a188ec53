"""Conway's Game of Life (B3/S23) on int32 cells.

Reference: ``game_of_life`` kernel.cu:10-68 (neighbour sum :61, rule :66), ``create_universe``
kernel.cu:131-146 (ring dead, interior alive with p = 0.15 from the unseeded glibc ``rand()``).
Two boundary readings exist in the reference text: "dead-outside" (the kernel's edge
branches, which are dead code because of unsigned compares) and "fixed-ring" (init + halo exchange).
"""
from .. import oracle
from . import Workload

LIFE2D = Workload(
    name="life", ndim=2, dtypes=("int32", "bits"), stencil_points=9, flops_per_update=10,
    reference="kernel.cu:61,66",
    init=oracle.life_init_reference, step=oracle.life_step,
    description="Game of Life B3/S23, 8-neighbourhood",
    # bit-packed cells: measured 24.4 TLUPS at TB=2 vs 16.3 at TB=4 (36 window registers per level)
    op_code=1, max_time_block=(("int32", 4), ("bits", 2)),
)
