#!/usr/bin/env python
"""A glider walking across a Game of Life universe, printed in the reference's print_array format (kernel.cu:115-129:
'0' for a live cell, ' ' for a dead one) every few generations -- the per-generation output the reference carries
commented out in its loop (kernel.cu:232,265).  With several ranks the glider crosses the slab cuts.

    python examples/life_glider.py
    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 examples/life_glider.py
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_cuda_process_b200 import RunConfig, Solver  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--height", type=int, default=24)
ap.add_argument("--width", type=int, default=32)
ap.add_argument("--generations", type=int, default=32)
ap.add_argument("--every", type=int, default=8)
a = ap.parse_args()

universe = np.zeros((a.height, a.width), dtype=np.int32)
universe[1:4, 1:4] = [[0, 1, 0], [0, 0, 1], [1, 1, 1]]          # a glider heading down and to the right

cfg = RunConfig(workload="life", height=a.height, width=a.width)
with Solver(cfg) as s:
    s.initialize(universe)                   # every rank passes the global array and keeps its slab
    done = 0
    while done < a.generations:
        n = min(a.every, a.generations - done)
        s.run(n)
        done += n
        frame = s.gather_ascii()             # bytes of print_array, assembled on rank 0
        if s.ctx.rank == 0:
            sys.stdout.write(f"generation {done}, {int(frame.count(b'0'))} live cells")
            sys.stdout.flush()
            sys.stdout.buffer.write(frame)
            sys.stdout.buffer.flush()
