"""3-D 7-point heat equation (extension; BASELINE.json config 3).

No counterpart in the reference, which is strictly 2-D (MDF_kernel.cu:13-18 index math).  Same structure as
the MDF rule lifted to three axes: ``u' = u + c*(sum of 6 neighbours - 6u)`` with ``c = 1/6`` (3-D stability
limit), boundary shell held fixed, every operation individually rounded so the oracle is exact.
"""
from .. import oracle
from . import Workload

HEAT3D = Workload(
    name="heat3d", ndim=3, dtypes=("fp32", "fp64"), stencil_points=7, flops_per_update=9,
    reference="extension (no 3-D code in the reference)",
    init=oracle.heat3d_init, step=oracle.heat3d_step,
    description="3-D 7-point heat, Dirichlet shell",
    max_time_block=(("fp32", 1), ("fp64", 1)),
)
