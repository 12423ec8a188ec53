#!/usr/bin/env python
"""Headline benchmark: grid-point updates per second (GLUPS) of the 2-D 5-point MDF solve.

Config (BASELINE.json #2): 16384 x 16384 fp32, 1-D slab decomposition over N GPUs of one box.  One step = one
complete solve of the configured 1000 iterations (``--gens-per-step 1000``, ONE kernel launch per rank); ``--steps K``
repeats that solve K times, so a run executes K x 1000 iterations and GLUPS is their rate.  Strong scaling: the global
grid is fixed.

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 10 --warmup 3

Rank 0 prints ONE JSON line.  ``value`` is device-timed (CUDA events around the K steps, max over ranks);
``e2e`` repeats the measurement through the public API with a pinned-host -> device upload of the rank's slab
and a device -> host read of the step's result (kernel status + global residual of the last iteration) inside every
step; ``e2e_with_gather`` additionally assembles the solution field on rank 0 (fused gather + pack over NVLink) and
copies it to pinned host memory every step.  Each rank binds itself to the CPUs next to its GPU first (so that pinned
buffers come from the GPU's memory node; ``--no-numa-bind`` switches that off; recorded under ``config``).
``--impl reference`` is the reference arm (the reference cannot be built or run here);
``--impl ref-fixed`` runs the patched reference program of ``baseline/`` (a baseline, not the reference arm).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "grid-point updates/sec (GLUPS, device-timed, max over ranks)"


def parse_args(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="fused", choices=["fused", "ours", "reference", "ref-fixed", "nccl", "staged"])
    p.add_argument("--workload", default="mdf", choices=["mdf", "life", "heat3d"],
                   help="mdf = the headline config; heat3d/life for the other BASELINE.json configs")
    p.add_argument("--depth", type=int, default=0)
    p.add_argument("--decomposition", default="slab", choices=["slab", "pencil"])
    p.add_argument("--height", type=int, default=16384)
    p.add_argument("--width", type=int, default=16384)
    p.add_argument("--gens-per-step", type=int, default=1000)
    p.add_argument("--time-block", type=int, default=8)
    p.add_argument("--dtype", default="fp32", choices=["fp32", "fp64", "bits"])
    p.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                   help="weak: --height is per GPU (BASELINE.json config 4 uses 8192 x 8192 per GPU)")
    p.add_argument("--chunk-rows", type=int, default=0, help="rows per tile of the schedule (0 = default)")
    p.add_argument("--sync-mode", type=int, default=None, help="1 = grid barrier between sweeps, 2 = dataflow flags")
    p.add_argument("--prefetch", type=int, default=None, help="0 = no next-tile TMA prefetch (A/B)")
    p.add_argument("--cluster3d", action="store_true", help="--impl nccl --workload heat3d: the cluster/DSMEM kernel")
    p.add_argument("--e2e-depth", type=int, default=1, help="steps enqueued ahead of the result being read (e2e)")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-numa-bind", action="store_true",
                   help="do not bind each rank to the CPUs (and memory node) next to its GPU")
    p.add_argument("--halo-message", default="row", choices=["row", "element"])
    p.add_argument("--coarse-halo-flags", action="store_true", help="one arrival counter per neighbour instead of per-strip flags")
    return p.parse_args(argv)


def reference_arm(args) -> int:
    """The unmodified reference cannot run here; say why in one line (details in DESIGN.md / BASELINE.md)."""
    why = ("reference is two bare .cu programs with no build system or Python package (pip install fails: no "
           "setup.py/pyproject), needs MPI (mpi.h/libmpi/mpirun absent, no network), and as written deadlocks for "
           "g>=2 (MDF_kernel.cu:168 vs :215) and reads ~16 GiB out of bounds on row 0 (MDF_kernel.cu:13-20), "
           "which would fault the GPU")
    rank = int(os.environ.get("RANK", "0"))
    if rank == 0:
        print(json.dumps({"impl": "reference", "unavailable": why, "metric": METRIC, "n_gpus": args.gpus}))
    return 0


def ref_fixed_arm(args) -> int:
    """``--impl ref-fixed``: the reference's own MDF program with baseline/ref_fixed.patch applied, started as N ranks on
    the MPI stand-in (NOT the unmodified reference, hence not ``--impl reference``).  It moves the whole grid over PCIe
    three times per generation, so a handful of generations is all that is timed; GLUPS is still a rate."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0                      # the stand-in launcher starts the ranks itself
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import run_ref_fixed
    gens = max(2, min(args.gens_per_step, 3))
    h = args.height * (args.gpus if args.scaling == "weak" else 1)
    try:
        rec = run_ref_fixed.run(args.gpus, gens, h, args.width)
    except Exception as e:            # no GPU, no reference tree to patch, ...: say so in one line
        print(json.dumps({"impl": "ref-fixed", "unavailable": str(e).strip().splitlines()[0][:300], "metric": METRIC,
                          "n_gpus": args.gpus}))
        return 0
    print(json.dumps({"metric": METRIC, "value": round(rec["glups"] , 4), "unit": "GLUPS", "n_gpus": args.gpus,
                      "impl": "ref-fixed", "higher_is_better": True, "scaling": args.scaling, "dtype": "fp32",
                      "data": "synthetic", "steps": 1, "warmup": 0, "ms_per_step": round(rec["seconds"] * 1e3, 3),
                      "config": {"model": "2-D 5-point MDF (heat/Jacobi), Dirichlet ring=100", "grid": rec["grid"],
                                 "generations": gens, "pattern": rec["pattern"]},
                      "note": "wall clock between barriers inside the program (it has no device timing); "
                              "baseline only, see BASELINE.md section 3"}))
    return 0


def maybe_respawn(args) -> None:
    """``python bench.py --gpus N`` without torchrun: start the N ranks ourselves."""
    if args.gpus > 1 and "RANK" not in os.environ:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29517"),
               os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))


def main(argv=None) -> int:
    args = parse_args(argv)
    if args.impl == "reference":
        return reference_arm(args)
    if args.impl == "ref-fixed":
        return ref_fixed_arm(args)
    maybe_respawn(args)

    import torch
    from mpi_cuda_process_b200 import RunConfig, Solver
    from mpi_cuda_process_b200.parallel.context import init_context
    from mpi_cuda_process_b200.utils.clocks import make_sampler
    from mpi_cuda_process_b200.utils import metrics

    ctx = init_context(device="cuda")
    # host side next to the GPU before the first pinned allocation (torchrun does not bind its children)
    from mpi_cuda_process_b200.utils import numa
    numa_rec = {"bound": False, "why": "disabled"} if args.no_numa_bind else numa.bind_to_gpu(ctx.device.index or 0)
    if ctx.world != args.gpus and ctx.rank == 0:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE={ctx.world}", file=sys.stderr)
    backend = {"ours": "fused"}.get(args.impl, args.impl)
    h = args.height * (ctx.world if args.scaling == "weak" else 1)
    w = args.width
    gps = args.gens_per_step
    depth = args.depth
    if args.workload == "heat3d":
        depth = depth or 512
        if args.scaling == "weak" and args.decomposition == "slab":
            depth, h = depth * ctx.world, args.height
    cfg = RunConfig(workload=args.workload, height=h, width=w, depth=depth if args.workload == "heat3d" else 0,
                    dtype=("bits" if args.dtype == "bits" else "int32") if args.workload == "life" else args.dtype,
                    backend=backend,
                    decomposition=args.decomposition, seed=0 if args.workload == "life" else None,
                    time_block=args.time_block, residual="linf", generations=gps, halo_message=args.halo_message,
                    extra={"chunk_rows": args.chunk_rows, "sync_mode": args.sync_mode, "cluster3d": args.cluster3d, "prefetch": args.prefetch,
                           "coarse_halo_flags": args.coarse_halo_flags})
    K, W = args.steps, max(args.warmup, 0)

    with Solver(cfg, ctx=ctx) as s:
        s.initialize()
        # ---------------- device-timed headline: K steps of `gps` generations ----------------
        for _ in range(W):
            s.run(gps, timed=False)
        s.check_status()
        ctx.barrier()
        torch.cuda.synchronize()
        launches0 = s.total_stats.launches
        sampler = make_sampler(ctx.device.index or 0).start() if ctx.rank == 0 else None
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(K):
            s.run(gps, timed=False)
        ev1.record()
        torch.cuda.synchronize()
        ctx.barrier()
        clocks = sampler.stop() if sampler is not None else None
        ms_total = ctx.all_reduce_max(ev0.elapsed_time(ev1))
        launches = s.total_stats.launches - launches0
        s.check_status()
        cells = cfg.cells
        value = cells * gps * K / (ms_total * 1e-3) / 1e9
        geom = None
        if backend == "fused" and cfg.ndim == 2:
            import ctypes as C
            ns, nc, grid, smem = C.c_int(), C.c_int(), C.c_int(), C.c_int()
            s._lib.mcp_fused2d_geometry(s._plan, s.time_block, C.byref(ns), C.byref(nc), C.byref(grid), C.byref(smem))
            rows_buf = (C.c_int * 4096)()
            nchunk = s._lib.mcp_fused2d_chunks(s._plan, s.time_block, rows_buf, 4096)
            crow = [rows_buf[i] for i in range(max(nchunk, 0))]
            geom = {"strips": ns.value, "chunks": nc.value, "ctas": grid.value, "smem_bytes": smem.value,
                    "chunk_rows_first_max_last": [crow[0], max(crow), crow[-1]] if crow else None}

        # ---------------- end to end through the public API ----------------
        e2e = e2e_gather = None
        if not args.no_e2e:
            host = torch.empty(s.store_shape, dtype=s.tdtype).pin_memory()
            host.copy_(s.bufs[s.cur])                     # a valid slab (incl. ghost rows) to upload every step
            torch.cuda.synchronize()
            h2d = host.numel() * host.element_size()
            tb = s.time_block
            d2h = 32                                       # result block: status, sweeps done, (sum of squares, max)

            # Every step: H2D of the step's input slab from pinned memory, one solve (one launch), D2H of the step's
            # result (status + global residual).  The host reads step i's result AFTER it has enqueued step i+1, so the
            # ranks' launches are not separated by a host round trip per step (the copy itself happens every step).
            pending = []

            marks = []                                    # (event before the step's device work, event after it)

            def one_step():
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                s.commit_input()                          # this step's input (uploaded while the previous step ran)
                s.stage_input(host)                       # pinned host -> device for the next step, overlapped
                pending.append(s.run_async(gps))          # solve + async 32-byte result copy
                e1.record()
                marks.append((e0, e1))
                return pending.pop(0).wait() if len(pending) > args.e2e_depth else None

            def drain():
                r = None
                while pending:
                    r = pending.pop(0).wait()
                return r

            s.stage_input(host)
            for _ in range(min(W, 2)):
                one_step()
            drain()
            torch.cuda.synchronize()
            ctx.barrier()
            t0 = time.perf_counter()
            last = None
            del marks[:]
            l0 = s.total_stats.launches
            for _ in range(K):
                r = one_step()
                last = r if r is not None else last
            last = drain()                                # the last step's result
            torch.cuda.synchronize()                      # includes the upload issued in the last step
            t1 = time.perf_counter()
            e2e_ms = ctx.all_reduce_max((t1 - t0) * 1e3)
            # device timeline of the steps: time inside a step's device work and idle gaps between consecutive steps
            busy = [a.elapsed_time(b) for a, b in marks]
            gaps = [marks[i][1].elapsed_time(marks[i + 1][0]) for i in range(len(marks) - 1)]
            step_busy_ms = ctx.all_reduce_max(sum(busy) / max(len(busy), 1))
            step_gap_ms = ctx.all_reduce_max(sum(gaps) / max(len(gaps), 1)) if gaps else 0.0
            # what bounds a step: the solve, or the host -> device upload of its input?  (device-timed, max over ranks;
            # "overlapped" = while the solve of the previous step runs, "alone" = all ranks uploading, GPUs idle)
            h2d_overlapped_ms = ctx.all_reduce_max(s.last_upload_ms() or 0.0)
            ctx.barrier()
            s.stage_input(host)
            h2d_alone_ms = ctx.all_reduce_max(s.last_upload_ms() or 0.0)
            e2e = {"value": round(cells * gps * K / (e2e_ms * 1e-3) / 1e9, 3), "unit": "GLUPS",
                   "ms_per_step": round(e2e_ms / K, 4), "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "result": "kernel status + global max|u'-u| of the last iteration (32-byte result block)",
                   "last_residual": last, "overlap": "H2D of step i+1 overlaps compute of step i (side stream)",
                   "launches_per_step": (s.total_stats.launches - l0) / max(K, 1),
                   "h2d_ms_overlapped_max": round(h2d_overlapped_ms, 3), "h2d_ms_alone_max": round(h2d_alone_ms, 3),
                   "h2d_gbs_per_gpu_overlapped": round(h2d / (h2d_overlapped_ms * 1e-3) / 1e9, 2) if h2d_overlapped_ms > 0 else None,
                   "bound": "upload" if h2d_overlapped_ms > 0.9 * e2e_ms / K else "solve",
                   "device_ms_per_step_busy_max": round(step_busy_ms, 3),
                   "device_ms_idle_between_steps_max": round(step_gap_ms, 3), "pipeline_depth": args.e2e_depth}

            # ---- the same, with the solution field gathered to rank 0 and copied to the host every step ----
            Kg = max(1, min(K, 4))

            def one_step_gather():
                one_step()
                r = drain()                               # the field is wanted now: wait for this step
                f = s.gather(copy=False)                  # fused gather+pack over NVLink, D2H into pinned memory
                return r, (None if f is None else float(f[0, 0]))

            one_step_gather()
            torch.cuda.synchronize()
            ctx.barrier()
            t0 = time.perf_counter()
            for _ in range(Kg):
                one_step_gather()
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            eg_ms = ctx.all_reduce_max((t1 - t0) * 1e3)
            e2e_gather = {"value": round(cells * gps * Kg / (eg_ms * 1e-3) / 1e9, 3), "unit": "GLUPS", "steps": Kg,
                          "ms_per_step": round(eg_ms / Kg, 4), "h2d_bytes_per_step": int(h2d),
                          "d2h_bytes_per_step": int(d2h + (cells * host.element_size() if ctx.rank == 0 else 0)),
                          "gather": s.last_gather_stats}

        if ctx.rank == 0:
            esize = 8 if args.dtype == "fp64" else (0.125 if args.dtype == "bits" else 4)
            cells_gpu = cells // ctx.world
            halo = 0 if ctx.world == 1 else int(2 * (cells // cfg.shape[0]) * esize)
            roof = metrics.roofline(cells_gpu, gps * K, ms_total, 2 * esize, halo, time_block=s.time_block)
            out = {
                "metric": METRIC, "value": round(value, 3), "unit": "GLUPS", "n_gpus": ctx.world, "steps": K,
                "warmup": W, "ms_per_step": round(ms_total / K, 4), "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": cfg.dtype, "data": "synthetic",
                "impl": args.impl if args.impl != "ours" else "fused",
                "config": {"model": {"mdf": "2-D 5-point MDF (heat/Jacobi), Dirichlet ring=100",
                                     "life": "Game of Life B3/S23 int32", "heat3d": "3-D 7-point heat, Dirichlet shell"}[args.workload],
                           "grid": list(cfg.shape),
                           "global_batch": cells, "seq_len": gps, "generations_per_step": gps,
                           "iterations": gps * K, "parallelism": f"{args.decomposition}{ctx.world}",
                           "time_block": s.time_block, "backend": s.backend,
                           "l2": "inputs larger than L2 (two %d MiB buffers per GPU stream through HBM every pass)"
                                 % (int(cells_gpu * esize) >> 20), "geometry": geom,
                           "host_affinity_rank0": numa_rec},
                "gpu_launches": int(launches),
                "clocks": clocks,
                "roofline": {k: (round(v, 4) if isinstance(v, float) else v) for k, v in roof.items()},
            }
            if e2e is not None:
                out["e2e"] = e2e
                out["e2e_with_gather"] = e2e_gather
            print(json.dumps(out), flush=True)
    ctx.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
