"""Per-tile timing of the fused 2-D kernel (in-kernel globaltimer trace): which tiles are the critical path?

    MCP_DEBUG_FLAGS=4 python scripts/tile_trace.py [--height H --width W --time-block TB --chunk-rows R]
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np

os.environ.setdefault("MCP_DEBUG_FLAGS", "4")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_cuda_process_b200 import RunConfig, Solver  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--height", type=int, default=2048)
ap.add_argument("--width", type=int, default=16384)
ap.add_argument("--time-block", type=int, default=8)
ap.add_argument("--chunk-rows", type=int, default=0)
ap.add_argument("--generations", type=int, default=160)
a = ap.parse_args()
cfg = RunConfig(workload="mdf", height=a.height, width=a.width, backend="fused", time_block=a.time_block,
                extra={"chunk_rows": a.chunk_rows})
with Solver(cfg) as s:
    s.initialize()
    s.run(a.generations)
    st = s.run(a.generations)
    cap = 1 << 20
    ns = (C.c_uint32 * cap)()
    ch = (C.c_int * cap)()
    sp = (C.c_int * cap)()
    n = s._lib.mcp_fused2d_tile_trace(s._plan, s.time_block, ns, ch, sp, cap)
    t = np.array(ns[:n], dtype=np.float64) / 1e3
    chunk = np.array(ch[:n])
    strip = np.array(sp[:n])
    rows_buf = (C.c_int * 4096)()
    nchunk = s._lib.mcp_fused2d_chunks(s._plan, s.time_block, rows_buf, 4096)
    rows = np.array(rows_buf[:nchunk])[chunk]
    nstr = strip.max() + 1
    edge = (strip == 0) | (strip == nstr - 1)
    tb = (chunk == 0) | (chunk == nchunk - 1)
    print(f"{a.height}x{a.width} TB={s.time_block}: {st.glups:.0f} GLUPS, {n} tiles, {nchunk} chunks x {nstr} strips")
    for name, m in (("interior", ~edge & ~tb), ("edge strips", edge & ~tb), ("top/bottom chunks", tb & ~edge)):
        if m.any():
            print(f"  {name:18s}: {m.sum():6d} tiles, {np.mean(t[m]):8.1f} us mean, {np.max(t[m]):8.1f} us max, "
                  f"{np.mean(t[m] / (rows[m] + 2 * s.time_block)) * 1e3:7.1f} ns per input row")
