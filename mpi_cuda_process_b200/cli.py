"""Run interface.

Compatible with the reference's: three prompts on stdout and three integers on stdin -- generations,
height, width (MDF_kernel.cu:103-112, kernel.cu:150-159):

    echo "100 256 256" | python -m mpi_cuda_process_b200 --workload mdf

The reference reads stdin on *every* rank before MPI_Init (defect D1: under most launchers only rank 0
has a stdin); here rank 0 prompts and reads, then broadcasts.  Everything else is a flag.  Under
``torchrun`` the same command runs one rank per GPU.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from typing import List, Optional

PROMPTS = (
    "Enter desired number of generations:\n",
    "Enter desired height of universe:\n",
    "Enter desired width of universe:\n",
)


def read_ghw(stream=None, out=None) -> List[int]:
    """Print the reference's three prompts and parse three whitespace-separated ints (scanf("%d") semantics)."""
    stream = stream if stream is not None else sys.stdin
    out = out if out is not None else sys.stdout
    vals: List[int] = []
    pending: List[str] = []
    for prompt in PROMPTS:
        out.write(prompt)
        out.flush()
        while not pending:
            line = stream.readline()
            if line == "":
                raise EOFError("expected three integers on stdin: generations height width")
            pending = line.split()
        vals.append(int(pending.pop(0)))
    return vals


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="mpi_cuda_process_b200", description=__doc__.split("\n\n")[0])
    from . import models
    p.add_argument("--workload", choices=sorted(models.REGISTRY), default="mdf",
                   help="; ".join(f"{w.name}: {w.description}" for w in models.REGISTRY.values()))
    p.add_argument("-g", "--generations", type=int, default=None, help="generations to run (else read from stdin)")
    p.add_argument("--height", type=int, default=None, help="rows of the grid (else read from stdin)")
    p.add_argument("--width", type=int, default=None, help="columns of the grid (else read from stdin)")
    p.add_argument("--depth", type=int, default=0, help="planes of the grid (heat3d)")
    p.add_argument("--dtype", choices=["fp32", "fp64", "bits"], default="fp32",
                   help="bits = bit-packed life cells (fused GPU backend)")
    p.add_argument("--bc", choices=["fixed-ring", "dead-outside"], default=None,
                   help="life: hold the outer ring, or treat cells outside the grid as dead (default); mdf and heat3d "
                        "always hold their boundary")
    p.add_argument("--backend", choices=["auto", "fused", "nccl", "staged", "cpu"], default="auto",
                   help="fused = the persistent sm_100a kernel (default on GPUs); nccl / staged = baselines; cpu = "
                        "torch CPU + gloo")
    p.add_argument("--decomposition", choices=["slab", "pencil"], default="slab",
                   help="slab = cut the first axis; pencil = cut z and y of a 3-D grid")
    p.add_argument("--pencil", type=int, nargs=2, default=(0, 0), metavar=("PZ", "PY"),
                   help="ranks along z and y for --decomposition pencil (0 0 = choose)")
    p.add_argument("--time-block", type=int, default=8,
                   help="generations per pass over the field in the fused 2-D kernel (1, 2, 4, 8)")
    p.add_argument("--tol", type=float, default=None, help="run to tolerance on the global residual")
    p.add_argument("--max-generations", type=int, default=0, help="upper bound for --tol runs")
    p.add_argument("--residual", choices=["none", "l2", "linf"], default="none",
                   help="norm of u' - u to report (and to compare with --tol)")
    p.add_argument("--seed", type=int, default=None,
                   help="life: counter-based random universe (default: the reference's unseeded glibc rand() stream); "
                        "mdf --init hash: seed of the random interior")
    p.add_argument("--prob", type=float, default=0.15, help="life: probability that a cell starts alive (kernel.cu:193)")
    p.add_argument("--init", choices=["ring", "hash", "box"], default="ring", help="mdf initial field")
    p.add_argument("--ring", type=float, default=100.0, help="boundary value (MDF_kernel.cu:93)")
    p.add_argument("--interior", type=float, default=0.0, help="initial interior value")
    p.add_argument("--halo-message", choices=["row", "element"], default="row",
                   help="staged baseline: one message per border row, or the reference's one per element")
    p.add_argument("--strict-fp64-accum", action="store_true",
                   help="unfused backends: evaluate the last step of the update in fp64 like the reference's SASS "
                        "(bit-identical to the default)")
    p.add_argument("--input", default=None,
                   help="initial field from a file (any --output-format; every rank reads it and keeps its block); the "
                        "reference can only start from its built-in field (MDF_kernel.cu:146)")
    p.add_argument("--input-format", choices=["ascii", "npy", "raw", "text"], default=None)
    p.add_argument("--output", default=None, help="write the final field here (rank 0)")
    p.add_argument("--output-format", choices=["ascii", "npy", "raw", "text"], default=None)
    p.add_argument("--print", dest="print_field", action="store_true",
                   help="print the final life grid in the reference's print_array format")
    p.add_argument("--print-every", type=int, default=0, metavar="N",
                   help="print the grid every N generations -- the per-generation print_array calls the reference "
                        "carries commented out inside its loop (kernel.cu:232,265)")
    p.add_argument("--output-every", type=int, default=0, metavar="N",
                   help="also write a frame every N generations: <output stem>.gNNNNNN<ext>")
    p.add_argument("--check", action="store_true", help="compare with the CPU oracle (small grids)")
    p.add_argument("--json", action="store_true", help="print run statistics as one JSON line")
    p.add_argument("--log-jsonl", default=None, help="append run statistics as JSON lines to this file (rank 0)")
    p.add_argument("--no-prompt", action="store_true", help="do not read g/h/w from stdin (flags only)")
    p.add_argument("--checkpoint-dir", default=None, help="write checkpoints (one file per rank + manifest) here")
    p.add_argument("--checkpoint-every", type=int, default=0, help="generations between checkpoints")
    p.add_argument("--resume", default=None, help="checkpoint directory to resume from")
    return p


def main(argv: Optional[List[str]] = None) -> int:
    args = build_parser().parse_args(argv)
    from .config import RunConfig
    from .parallel.context import init_context
    from .solver import Solver
    from . import io as mio, oracle
    from .utils import checkpoint as ckpt

    ctx = init_context(device="cpu" if args.backend == "cpu" else None)
    need_stdin = not args.no_prompt and (args.generations is None or args.height is None or args.width is None)
    ghw = None
    if need_stdin:
        if ctx.rank == 0:
            ghw = read_ghw()
        ghw = ctx.broadcast_object(ghw, 0)
    g = args.generations if args.generations is not None else (ghw[0] if ghw else 1)
    h = args.height if args.height is not None else (ghw[1] if ghw else 256)
    w = args.width if args.width is not None else (ghw[2] if ghw else 256)

    cfg = RunConfig(
        workload=args.workload, generations=g, height=h, width=w, depth=args.depth,
        dtype=("bits" if args.dtype == "bits" else "int32") if args.workload == "life" else args.dtype,
        bc=args.bc,
        ring=args.ring, interior=args.interior, prob=args.prob, init=args.init, seed=args.seed, backend=args.backend,
        decomposition=args.decomposition, pencil=tuple(args.pencil), time_block=args.time_block, tol=args.tol,
        max_generations=args.max_generations, residual=args.residual, halo_message=args.halo_message,
        output=args.output, output_format=args.output_format or ("ascii" if args.workload == "life" else "npy"),
        strict_fp64_accum=args.strict_fp64_accum, checkpoint_every=args.checkpoint_every,
        checkpoint_dir=args.checkpoint_dir)

    rc = 0
    start = None
    with Solver(cfg, ctx=ctx) as solver:
        if args.resume:
            ckpt.load(solver, args.resume)
        elif args.input:
            fmt = args.input_format or ("npy" if args.input.endswith(".npy") else cfg.output_format)
            start = mio.load_field(args.input, fmt, shape=cfg.shape, dtype=solver.npdtype, w=cfg.width)
            if tuple(start.shape) != tuple(cfg.shape):
                raise SystemExit(f"--input holds a field of shape {tuple(start.shape)}, the run is {tuple(cfg.shape)}")
            solver.initialize(start.astype(solver.npdtype, copy=False))
        else:
            solver.initialize()

        def emit(print_it: bool, path: Optional[str]) -> Optional["np.ndarray"]:
            """Collective: gather on every rank, print / write on rank 0.  Returns the gathered field (rank 0)."""
            life = cfg.workload == "life"
            need_field = bool(path) or (print_it and not life)
            fld = solver.gather() if need_field else None
            text = solver.gather_ascii() if (print_it and life) else None
            if ctx.rank == 0:
                if text is not None:
                    sys.stdout.flush()
                    sys.stdout.buffer.write(text)
                    sys.stdout.buffer.flush()
                elif print_it and fld is not None:
                    mio.print_array(fld)
                if path and fld is not None:
                    mio.dump_field(fld, path, cfg.output_format)
            return fld

        def frame_path(generation: int) -> Optional[str]:
            if not args.output:
                return None
            stem, ext = os.path.splitext(args.output)
            return f"{stem}.g{generation:06d}{ext}"

        ckpt_every = cfg.checkpoint_every if cfg.checkpoint_dir else 0
        periods = [n for n in (ckpt_every, args.print_every, args.output_every if args.output else 0) if n > 0]
        if cfg.tol is not None:
            stats = solver.run_to_tolerance()
        elif periods:
            # segments end at every multiple of a period (generations counted from this run's start)
            from .solver import RunStats
            stats = RunStats(cells=cfg.cells)
            done = 0
            while done < g:
                n = min(min((done // q + 1) * q for q in periods), g) - done
                s = solver.run(n, residual=cfg.residual != "none")
                stats.generations += n
                stats.device_ms += s.device_ms
                stats.launches += s.launches
                stats.residual_history += s.residual_history
                done += n
                if ckpt_every and (done % ckpt_every == 0 or done == g):
                    ckpt.save(solver, cfg.checkpoint_dir)
                want_print = args.print_every > 0 and done % args.print_every == 0 and done < g
                want_frame = args.output_every > 0 and done % args.output_every == 0
                if want_print or (want_frame and args.output):
                    emit(want_print, frame_path(done) if want_frame else None)
        else:
            stats = solver.run(g, residual=cfg.residual != "none")
        solver.check_status()
        final_print = args.print_field or args.print_every > 0
        field = emit(final_print, args.output)
        gathered = bool(args.output) or (final_print and cfg.workload != "life")
        if args.check and not gathered:        # decided from the flags: the gather is collective, `field` is rank 0's
            field = solver.gather()
        if ctx.rank == 0:
            if args.check and field is not None:
                ref = oracle.run_field(cfg, start if start is not None else oracle.init_field(cfg), solver.generation)
                same = bool((ref == field).all())
                print(f"oracle check: {'bitwise equal' if same else 'MISMATCH'} after {solver.generation} generations")
                rc = 0 if same else 1
            summary = {
                "workload": cfg.workload, "backend": solver.backend, "world": ctx.world, "shape": list(cfg.shape),
                "generations": stats.generations, "device_ms": round(stats.device_ms, 4),
                "glups": round(stats.glups, 3), "launches": stats.launches,
                "residual": stats.residual_history[-1][1] if stats.residual_history else None,
                "stop_reason": stats.stop_reason or None,
            }
            if args.log_jsonl:
                from .utils.log import JsonlLogger
                lg = JsonlLogger(args.log_jsonl, rank=ctx.rank)
                lg.log(**summary, residual_history=stats.residual_history[-64:])
                lg.close()
            if args.json:
                print(json.dumps(summary))
            else:
                print(f"{cfg.workload} {cfg.shape} x {stats.generations} generations on {ctx.world} rank(s) "
                      f"[{solver.backend}]: {stats.device_ms:.3f} ms, {stats.glups:.2f} GLUPS")
        rc = ctx.broadcast_object(rc, 0)
    ctx.close()
    return rc


if __name__ == "__main__":
    sys.exit(main())
