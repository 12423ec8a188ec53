"""Checkpoint / resume (absent in the reference).

Each rank writes its owned block; a manifest records the config, generation and decomposition.  Resume
requires the same grid but not the same world size: blocks are re-cut from the saved pieces.
"""
from __future__ import annotations

import json
import os
from pathlib import Path

import numpy as np


def save(solver, directory: str) -> str:
    d = Path(directory)
    d.mkdir(parents=True, exist_ok=True)
    ctx = solver.ctx
    loc = solver.local_owned()
    if ctx.is_cuda:
        import torch
        torch.cuda.synchronize(ctx.device)
    arr = loc.detach().cpu().contiguous().numpy()
    tmp = d / f"block_{ctx.rank:04d}.npy.tmp"
    with open(tmp, "wb") as f:
        np.save(f, arr)
    os.replace(tmp, d / f"block_{ctx.rank:04d}.npy")
    meta = ctx.all_gather_object({"rank": ctx.rank, "start": list(solver.block.start), "shape": list(arr.shape)})
    if ctx.rank == 0:
        man = {"config": json.loads(solver.cfg.to_json()), "generation": solver.generation, "world": ctx.world,
               "blocks": meta}
        tmpm = d / "manifest.json.tmp"
        tmpm.write_text(json.dumps(man, indent=1))
        os.replace(tmpm, d / "manifest.json")
    ctx.barrier()
    return str(d)


def load(solver, directory: str) -> int:
    d = Path(directory)
    man = json.loads((d / "manifest.json").read_text())
    c = man["config"]
    shape = (c["depth"], c["height"], c["width"]) if c["workload"] == "heat3d" else (c["height"], c["width"])
    if tuple(shape) != tuple(solver.cfg.shape):
        raise ValueError(f"checkpoint grid {shape} != configured {solver.cfg.shape}")
    # Assemble only what this rank needs (its block + ghosts) from the saved pieces.
    G = solver.ghost
    lo = [solver.block.start[a] - G for a in range(solver.ndim - 1)] + [0]
    hi = [solver.block.stop[a] + G for a in range(solver.ndim - 1)] + [solver.cfg.width]
    # bit-packed life stores 32 cells per word: assemble CELLS (what save() wrote), pack them afterwards
    bits = getattr(solver, "bits", False)
    cells_shape = tuple(solver.store_shape[:-1]) + (solver.cfg.width,) if bits else tuple(solver.store_shape)
    loc = np.zeros(cells_shape, dtype=solver.npdtype)
    for b in man["blocks"]:
        bs, bshape = b["start"], b["shape"]
        be = [s + n for s, n in zip(bs, bshape)]
        ov_lo = [max(l, s) for l, s in zip(lo, bs)]
        ov_hi = [min(h, e) for h, e in zip(hi, be)]
        if any(a >= z for a, z in zip(ov_lo, ov_hi)):
            continue
        piece = np.load(d / f"block_{b['rank']:04d}.npy", mmap_mode="r")
        src = tuple(slice(a - s, z - s) for a, z, s in zip(ov_lo, ov_hi, bs))
        dst = tuple(slice(a - l, z - l) for a, z, l in zip(ov_lo, ov_hi, lo))
        loc[dst] = piece[src]
    import torch
    if bits:
        loc = solver._pack_bits(loc)
    t = torch.from_numpy(loc)
    for buf in solver.bufs:
        buf.copy_(t)
    solver.cur = 0
    solver._ghosts_valid = True
    solver._reset_plan_parity()
    solver.generation = int(man["generation"])
    if solver.ctx.is_cuda:
        torch.cuda.synchronize(solver.ctx.device)
    solver.ctx.barrier()
    return solver.generation
