"""mpi_cuda_process_b200 -- a B200-native multi-GPU stencil framework.

Provides the capabilities of the CUDA+MPI reference ``Rodrigovicente/MPI-CUDA-Process`` (2-D 5-point
heat/Jacobi "MDF" solver, Game of Life; row-slab decomposition with halo exchange; stdin run interface),
re-designed for one 8xB200 box: fields are torch tensors in a CUDA-IPC symmetric arena, the hot path is
one persistent sm_100a kernel that fuses the halo exchange (NVLink peer stores + flags), temporal
blocking and the residual reduction into the stencil sweep.
"""
from .config import RunConfig  # noqa: F401

__version__ = "0.1.0"
__all__ = ["RunConfig", "Solver", "RunStats", "oracle", "io"]


def __getattr__(name):
    if name in ("Solver", "RunStats"):
        from . import solver
        return getattr(solver, name)
    if name in ("oracle", "io"):
        import importlib
        return importlib.import_module(f".{name}", __name__)
    raise AttributeError(name)
