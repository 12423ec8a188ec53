#!/usr/bin/env python
"""A hot-edged plate relaxing towards its steady state: the reference's MDF problem (ring held at 100, interior 0,
MDF_kernel.cu:20,88-99), run until the largest change per generation drops below a tolerance -- something the
reference cannot do (fixed generation count, MDF_kernel.cu:153).

    python examples/heat_plate.py                                   # one GPU, or the CPU backend without one
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/heat_plate.py --size 16384

On GPUs the whole solve is ONE kernel launch per rank: the kernel all-reduces the residual over NVLink after every
time block and stops itself.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_cuda_process_b200 import RunConfig, Solver  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--size", type=int, default=512)
ap.add_argument("--tol", type=float, default=1e-2)
ap.add_argument("--max-generations", type=int, default=200000)
ap.add_argument("--dtype", default="fp32", choices=["fp32", "fp64"])
ap.add_argument("--out", default=None, help="write the final field as .npy (rank 0)")
a = ap.parse_args()

cfg = RunConfig(workload="mdf", height=a.size, width=a.size, dtype=a.dtype, tol=a.tol, residual="linf",
                max_generations=a.max_generations)
with Solver(cfg) as s:
    s.initialize()
    st = s.run_to_tolerance()
    s.check_status()
    field = s.gather()                       # rank 0: the global field; None elsewhere
    if s.ctx.rank == 0:
        mid = field[a.size // 2, a.size // 2]
        print(f"{a.size} x {a.size} {a.dtype} on {s.ctx.world} rank(s) [{s.backend}]: {st.generations} generations, "
              f"{st.launches} launch(es), stop: {st.stop_reason}, last max|du| = "
              f"{st.residual_history[-1][1]:.3e}, centre = {mid:.4f}, {st.glups:.1f} GLUPS")
        if a.out:
            import numpy as np
            np.save(a.out, field)
