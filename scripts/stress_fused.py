import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from mpi_cuda_process_b200 import RunConfig, Solver, oracle

def trial(workload, h, w, tb, g, seed, chunk=0):
    rng = np.random.default_rng(seed)
    if workload == "life":
        f = (rng.random((h, w)) < 0.4).astype(np.int32)
        cfg = RunConfig(workload="life", height=h, width=w, backend="fused", time_block=tb, extra={"chunk_rows": chunk})
        ref = oracle.life_run(f, g)
    else:
        f = oracle.mdf_init(h, w)
        f[1:-1, 1:-1] = rng.uniform(0, 100, (h - 2, w - 2)).astype(np.float32)
        cfg = RunConfig(workload="mdf", height=h, width=w, backend="fused", time_block=tb, extra={"chunk_rows": chunk})
        ref = oracle.mdf_run(f, g)
    with Solver(cfg) as s:
        s.initialize(f)
        s.run(g)
        got = s.gather()
    bad = np.argwhere(got != ref)
    return len(bad), (bad[:, 0].min(), bad[:, 0].max(), bad[:, 1].min(), bad[:, 1].max()) if len(bad) else None

n = int(sys.argv[1]) if len(sys.argv) > 1 else 30
fails = {}
t0 = time.time()
for wl in ("life", "mdf"):
    for (h, w, tb, g) in [(40, 70, 4, 12), (130, 200, 2, 6), (64, 300, 4, 8), (200, 130, 1, 5)]:
        nf = 0
        first = None
        for i in range(n):
            nb, box = trial(wl, h, w, tb, g, i)
            if nb:
                nf += 1
                first = first or (i, nb, box)
        print(f"{wl} {h}x{w} tb={tb} g={g}: {nf}/{n} failed", first, flush=True)
print("flags", os.environ.get("MCP_DEBUG_FLAGS"), "time", round(time.time() - t0, 1))
