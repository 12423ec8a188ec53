"""One short launch of one kernel family for an ncu capture (run plain first, then under ncu; see profiles/README.md).

    python scripts/ncu_case.py mdf2d --height 16384 --width 16384 --time-block 8 --generations 16 [--chunk-rows R]
    python scripts/ncu_case.py heat3d --depth 256 --height 512 --width 1024 --generations 2
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_cuda_process_b200 import RunConfig, Solver  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("what", choices=["mdf2d", "life", "heat3d"])
ap.add_argument("--depth", type=int, default=256)
ap.add_argument("--height", type=int, default=16384)
ap.add_argument("--width", type=int, default=16384)
ap.add_argument("--time-block", type=int, default=8)
ap.add_argument("--generations", type=int, default=16)
ap.add_argument("--chunk-rows", type=int, default=0)
ap.add_argument("--dtype", default="fp32")
ap.add_argument("--residual", action="store_true")
ap.add_argument("--cluster3d", action="store_true", help="heat3d through the plain backend's cluster + DSMEM kernel")
a = ap.parse_args()
kw = dict(backend="nccl" if a.cluster3d else "fused", time_block=a.time_block,
          extra={"chunk_rows": a.chunk_rows, "cluster3d": a.cluster3d})
if a.what == "heat3d":
    cfg = RunConfig(workload="heat3d", depth=a.depth, height=a.height, width=a.width, dtype=a.dtype, **kw)
elif a.what == "life":
    cfg = RunConfig(workload="life", height=a.height, width=a.width, dtype=a.dtype, seed=1, **kw)
else:
    cfg = RunConfig(workload="mdf", height=a.height, width=a.width, dtype=a.dtype, residual="linf", **kw)
with Solver(cfg) as s:
    s.initialize()
    s.run(a.generations, residual=a.residual)      # warm-up launch (skipped with ncu -s)
    st = s.run(a.generations, residual=a.residual)
    s.check_status()
    print(f"{a.what} {cfg.shape} {a.dtype} TB={s.time_block}: {st.glups:.1f} GLUPS, {st.launches} launch(es)")
