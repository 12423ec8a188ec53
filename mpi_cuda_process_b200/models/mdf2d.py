"""2-D 5-point finite-difference heat/Jacobi update ("MDF", Metodo das Diferencas Finitas).

Reference: ``run_mdf`` MDF_kernel.cu:10-22 (rule at :20), ``create_universe`` MDF_kernel.cu:88-99.
``u' = u + 0.25*(uE + uW + uN + uS - 4u)`` = explicit FTCS step at the 2-D stability limit = one Jacobi
sweep for Laplace's equation; Dirichlet ring held at 100, steady state 100 everywhere.
"""
from .. import oracle
from . import Workload

MDF2D = Workload(
    name="mdf", ndim=2, dtypes=("fp32", "fp64"), stencil_points=5, flops_per_update=7,
    reference="MDF_kernel.cu:20",
    init=oracle.mdf_init, step=oracle.mdf_step,
    description="2-D 5-point heat/Jacobi, Dirichlet ring",
    op_code=0, max_time_block=(("fp32", 8), ("fp64", 8)),
)
