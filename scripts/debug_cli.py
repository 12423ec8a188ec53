import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from mpi_cuda_process_b200 import RunConfig, Solver, oracle
for (h, w, tb, g, chunk) in [(256, 256, 4, 100, 0), (256, 256, 4, 100, 256), (256, 256, 4, 40, 0), (256, 256, 4, 48, 0), (256, 256, 4, 52, 0), (256, 256, 4, 64, 0), (256, 256, 1, 100, 0), (256, 256, 2, 100, 0), (300, 500, 4, 100, 0)]:
    cfg = RunConfig(workload="mdf", height=h, width=w, backend="fused", time_block=tb, extra={"chunk_rows": chunk})
    with Solver(cfg) as s:
        s.initialize()
        s.run(g)
        s.check_status()
        got = s.gather()
    ref = oracle.mdf_run(oracle.mdf_init(h, w), g)
    bad = np.argwhere(got != ref)
    print(h, w, tb, g, chunk, "mismatches", len(bad), (bad[:, 0].min(), bad[:, 0].max(), bad[:, 1].min(), bad[:, 1].max()) if len(bad) else "",
          "maxdiff", float(np.abs(got - ref).max()), flush=True)
