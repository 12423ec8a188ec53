import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from mpi_cuda_process_b200 import RunConfig, Solver, oracle

def mdf(h, w, tb, g, res):
    cfg = RunConfig(workload="mdf", height=h, width=w, backend="fused", time_block=tb, residual="linf")
    with Solver(cfg) as s:
        s.initialize()
        s.run(g, residual=res)
        got = s.gather()
    print("mdf", h, w, tb, g, res, "ok" if got.tobytes() == oracle.mdf_run(oracle.mdf_init(h, w), g).tobytes() else "BAD")

def life(h, w, tb, g, dens=None, backend="fused", bc="dead-outside"):
    f = oracle.life_init_reference(h, w) if dens is None else (np.random.default_rng(9).random((h, w)) < dens).astype(np.int32)
    cfg = RunConfig(workload="life", height=h, width=w, backend=backend, time_block=tb, bc=bc)
    with Solver(cfg) as s:
        s.initialize(f)
        s.run(g)
        got = s.gather()
        full = s.bufs[s.cur].cpu().numpy()
    ref = oracle.life_run(f, g, bc)
    bad = np.argwhere(got != ref)
    print(f"life {backend} h={h} w={w} tb={tb} g={g}: mismatches={len(bad)}",
          (f"rows {sorted(set(bad[:,0].tolist()))[:20]} cols {sorted(set(bad[:,1].tolist()))[:20]}" if len(bad) else ""),
          "pad nonzero:", int((full[:, w:] != 0).sum()), "ghost nonzero:", int((full[:s.ghost] != 0).sum() + (full[-s.ghost:] != 0).sum()))

which = sys.argv[1]
if which == "a":
    mdf(300, 500, 4, 10, True); life(130, 200, 2, 6)
elif which == "b":
    mdf(300, 500, 4, 10, False); life(130, 200, 2, 6)
elif which == "c":
    life(150, 275, 4, 23, backend="nccl", bc="fixed-ring"); life(40, 70, 4, 12, 0.5)
elif which == "d":
    life(150, 275, 4, 23, backend="fused", bc="fixed-ring"); life(40, 70, 4, 12, 0.5)
