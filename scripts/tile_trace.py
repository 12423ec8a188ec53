"""Per-tile timing of the fused 2-D kernel (in-kernel globaltimer trace): which tiles are the critical path, and how
long do tiles wait for their dependencies (previous-sweep tiles, neighbour GPUs' halo flags)?

    python scripts/tile_trace.py [--height H --width W --time-block TB --chunk-rows R]
    python -m torch.distributed.run --nproc-per-node 8 ... scripts/tile_trace.py      (one line per rank)
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_cuda_process_b200 import RunConfig, Solver  # noqa: E402
from mpi_cuda_process_b200.parallel.context import init_context  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--height", type=int, default=16384)
ap.add_argument("--width", type=int, default=16384)
ap.add_argument("--time-block", type=int, default=8)
ap.add_argument("--chunk-rows", type=int, default=0)
ap.add_argument("--generations", type=int, default=800)
ap.add_argument("--prefetch", type=int, default=None)
a = ap.parse_args()
ctx = init_context(device="cuda")
cfg = RunConfig(workload="mdf", height=a.height, width=a.width, backend="fused", time_block=a.time_block,
                extra={"chunk_rows": a.chunk_rows, "debug_flags": 4, "prefetch": a.prefetch})
with Solver(cfg, ctx=ctx) as s:
    s.initialize()
    s.run(a.generations)
    st = s.run(a.generations)
    s.check_status()
    cap = 1 << 20
    ch, sp = (C.c_int * cap)(), (C.c_int * cap)()
    buf = (C.c_uint32 * cap)()
    n = s._lib.mcp_fused2d_tile_trace(s._plan, s.time_block, buf, ch, sp, cap)
    chunk, strip = np.array(ch[:n]), np.array(sp[:n])
    q = {}
    for which, name in enumerate(("rows", "wait", "publish")):
        s._lib.mcp_fused2d_tile_trace2(s._plan, s.time_block, which, buf, cap)
        q[name] = np.array(buf[:n], dtype=np.float64) / 1e3 / (a.generations // s.time_block)   # us per sweep
    nchunk, nstr = chunk.max() + 1, strip.max() + 1
    edge = (strip == 0) | (strip == nstr - 1)
    tb_ = (chunk == 0) | (chunk == nchunk - 1)
    out = {"rank": ctx.rank, "world": ctx.world, "glups": round(st.glups, 1), "tiles": int(n), "chunks": int(nchunk),
           "strips": int(nstr)}
    for name, m in (("interior", ~edge & ~tb_), ("edge_strips", edge & ~tb_), ("cut_chunks", tb_ & ~edge)):
        if m.any():
            out[name] = {k: [round(float(v[m].mean()), 2), round(float(v[m].max()), 2)] for k, v in q.items()}
    tot = {k: float(v.sum()) for k, v in q.items()}
    out["share_of_warp_time"] = {k: round(v / sum(tot.values()), 4) for k, v in tot.items()}
    print(json.dumps(out), flush=True)
ctx.close()
