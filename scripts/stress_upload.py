"""R7 of docs/PROTOCOL.md over real NVLink: the host overwrites each rank's slab between launches with NO barrier between
the ranks -- safe only because a launch drains its incoming halo pushes before it ends.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/stress_upload.py [--delays]

Every rank keeps a pinned copy of its slab (owned rows + ghost rows) after G0 generations of a hashed field, then
repeats `load_local(copy); run(G1)` REPS times without any host synchronisation in between (no timing events, no
collectives: `run(timed=False)`), optionally with the kernel's injected delays switched on.  Each repetition restarts
from the same state, so the final field must be generation G0 + G1 of the oracle run: compared by device checksum
(`Solver.checksum`) against one straight run of G0 + G1 generations on the NCCL-halo backend.  A late
halo push of repetition i landing in a slab that repetition i + 1 has already uploaded would change the checksum.
The same path is what `bench.py`'s end-to-end loop exercises (`commit_input` -> `run_async`).  Needs >= 2 GPUs; not
part of pytest (the round's GPU budget was spent when it was written: run it before relying on it).
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_cuda_process_b200 import RunConfig, Solver  # noqa: E402
from mpi_cuda_process_b200.parallel.context import init_context  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--height", type=int, default=4096)
ap.add_argument("--width", type=int, default=4096)
ap.add_argument("--g0", type=int, default=16)
ap.add_argument("--g1", type=int, default=40)
ap.add_argument("--reps", type=int, default=12)
ap.add_argument("--delays", action="store_true", help="pseudo-random stalls in front of halo pushes and publishes")
a = ap.parse_args()

ctx = init_context(device="cuda")


def restarted(extra: dict) -> int:
    cfg = RunConfig(workload="mdf", height=a.height, width=a.width, backend="fused", init="hash", seed=5, time_block=8,
                    extra=extra)
    with Solver(cfg, ctx=ctx) as s:
        s.initialize()
        s.run(a.g0)
        torch.cuda.synchronize()
        ctx.barrier()                                   # every neighbour's last push of this run has landed
        host = torch.empty(s.store_shape, dtype=s.tdtype).pin_memory()
        host.copy_(s.bufs[s.cur])                       # owned rows + the ghost rows the neighbours pushed
        torch.cuda.synchronize()
        for _ in range(a.reps):
            s.load_local(host)                          # asynchronous H2D on the compute stream, no barrier
            s.run(a.g1, timed=False)                    # no events, no collectives
        s.check_status()
        return s.checksum()


def straight(backend: str) -> int:
    cfg = RunConfig(workload="mdf", height=a.height, width=a.width, backend=backend, init="hash", seed=5, time_block=8)
    with Solver(cfg, ctx=ctx) as s:
        s.initialize()
        s.run(a.g0 + a.g1)
        return s.checksum()


got = restarted({"debug_flags": 8} if a.delays else {})
want = straight("nccl")
if ctx.rank == 0:
    print(json.dumps({"case": "upload between launches without a host barrier", "world": ctx.world,
                      "grid": [a.height, a.width], "reps": a.reps, "delays": a.delays,
                      "restarted_fused": hex(got), "straight_nccl": hex(want), "ok": got == want}), flush=True)
ctx.close()
sys.exit(0 if got == want else 1)
