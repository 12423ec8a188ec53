import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from mpi_cuda_process_b200 import RunConfig, Solver, oracle

def run(h, w, tb, g, dens=None, chunk=0, bc="dead-outside"):
    if dens is None:
        f = oracle.life_init_reference(h, w)
    else:
        f = (np.random.default_rng(9).random((h, w)) < dens).astype(np.int32)
    cfg = RunConfig(workload="life", height=h, width=w, backend="fused", time_block=tb, bc=bc, extra={"chunk_rows": chunk})
    with Solver(cfg) as s:
        s.initialize(f)
        s.run(g)
        got = s.gather()
    ref = oracle.life_run(f, g, bc)
    bad = np.argwhere(got != ref)
    print(f"h={h} w={w} tb={tb} g={g} dens={dens} chunk={chunk} bc={bc}: mismatches={len(bad)}",
          (f"rows {sorted(set(bad[:,0].tolist()))[:12]} cols {sorted(set(bad[:,1].tolist()))[:12]}" if len(bad) else ""))

run(40, 70, 4, 12, 0.5)
run(40, 70, 4, 4, 0.5)
run(40, 70, 4, 4, 0.5, chunk=40)
run(40, 70, 1, 1, 0.5, chunk=40)
run(40, 70, 1, 1, 0.5)
run(40, 70, 2, 2, 0.5, chunk=40)
run(130, 200, 2, 6)
run(130, 200, 2, 2)
run(130, 200, 2, 2, chunk=130)
run(130, 200, 1, 1)
run(130, 200, 4, 4)
run(150, 275, 2, 6)
run(150, 275, 4, 12, 0.5)
run(40, 70, 4, 4, 0.5, chunk=40, bc="fixed-ring")
