"""Flag-protocol stress of the fused 2-D kernel: random grids, rank counts, time blocks, tile heights and launch
sequences, every halo push and publish preceded by a pseudo-random 5-200 us stall (debug bit 8), all ranks emulated on
ONE GPU inside one cooperative launch (tests/_emu.py).  Every case is compared bit for bit with the oracle.
On a multi-GPU box the same injected delays run across real NVLink in tests/mgpu_worker.py ("fused+delays").

    python scripts/stress_fused.py [--cases 20] [--seed 1]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from mpi_cuda_process_b200 import oracle  # noqa: E402
from _emu import EmulatedRanks  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--cases", type=int, default=20)
ap.add_argument("--seed", type=int, default=1)
ap.add_argument("--only", type=int, default=-1, help="run only this case index")
ap.add_argument("--prefetch", type=int, default=1)
ap.add_argument("--delays", type=int, default=1)
a = ap.parse_args()
rng = np.random.default_rng(a.seed)
for case in range(a.cases):
    tb = int(rng.choice([4, 8]))
    n = int(rng.integers(2, 9))
    h = int(rng.integers(n * tb * 3, n * tb * 3 + 600))
    w = int(rng.integers(40, 1500))
    chunk = int(rng.choice([0, tb, 3 * tb, 40]))
    sync = int(rng.choice([1, 2]))
    f0 = oracle.mdf_init(h, w)
    f0[1:-1, 1:-1] = rng.uniform(0, 100, (h - 2, w - 2)).astype(np.float32)
    sflags = bool(rng.integers(0, 2))
    plan = [(int(rng.integers(1, 6)), int(rng.choice([0, 0, 2]))) for _ in range(int(rng.integers(1, 4)))]
    if a.only >= 0 and case != a.only:
        continue
    # MCP_OPT_DEBUG_FLAGS = 8 (injected delays), MCP_OPT_SYNC_MODE, MCP_OPT_PREFETCH
    emu = EmulatedRanks(f0, n, tb, strip_flags=sflags, chunk_rows=chunk,
                        options={2: 8 if a.delays else 0, 3: sync, 4: a.prefetch})
    g = 0
    for sweeps, res_mode in plan:
        emu.run(sweeps * tb, residual_mode=res_mode)
        g += sweeps * tb
    ok = emu.status() == [0] * n and emu.field().tobytes() == oracle.mdf_run(f0, g).tobytes()
    emu.close()
    print(f"case {case}: {h}x{w} ranks={n} tb={tb} chunk_rows={chunk} sync={sync} strip_flags={sflags} plan={plan} "
          f"prefetch={a.prefetch} delays={a.delays} generations={g}: "
          f"{'ok' if ok else 'MISMATCH'}", flush=True)
    if not ok and a.only < 0:
        sys.exit(1)
print("stress done")
