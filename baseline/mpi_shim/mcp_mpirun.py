#!/usr/bin/env python
"""mpirun stand-in for baseline/mpi_shim: ``mcp_mpirun.py -np N prog [args...]``.

Creates one named pipe per ordered rank pair, starts N copies of ``prog`` with MCP_MPI_RANK / MCP_MPI_SIZE /
MCP_MPI_DIR set, and copies its own stdin to EVERY rank (the reference reads g, h, w with scanf on every rank
before MPI_Init, MDF_kernel.cu:105-112 -- under a real mpirun only rank 0 would see them)."""
import argparse
import os
import subprocess
import sys
import tempfile


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("-np", type=int, default=2)
    ap.add_argument("--timeout", type=float, default=None)
    ap.add_argument("prog", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    if not a.prog:
        ap.error("no program given")
    data = sys.stdin.buffer.read() if not sys.stdin.isatty() else b""
    with tempfile.TemporaryDirectory(prefix="mcp_mpi_") as d:
        for s in range(a.np):
            for t in range(a.np):
                os.mkfifo(os.path.join(d, f"p_{s}_{t}"))
        procs = []
        for r in range(a.np):
            env = dict(os.environ, MCP_MPI_RANK=str(r), MCP_MPI_SIZE=str(a.np), MCP_MPI_DIR=d)
            procs.append(subprocess.Popen(a.prog, env=env, stdin=subprocess.PIPE))
        for p in procs:
            try:
                p.stdin.write(data)
                p.stdin.close()
            except BrokenPipeError:
                pass
        rc = 0
        for p in procs:
            try:
                rc = max(rc, abs(p.wait(timeout=a.timeout)))
            except subprocess.TimeoutExpired:
                p.kill()          # the exact process we started
                rc = max(rc, 124)
        return rc


if __name__ == "__main__":
    sys.exit(main())
