#!/usr/bin/env python
"""Run the ``ref-fixed`` baseline (see build_ref_fixed.py) on N ranks through the MPI stand-in and print ONE JSON line:
GLUPS over the slowest rank's barrier-to-barrier time.

    python baseline/run_ref_fixed.py -np 8 --generations 4 --height 16384 --width 16384 [--dump /tmp/field]
"""
import argparse
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def run(np_: int, g: int, h: int, w: int, dump: str = None, timeout: float = 900.0) -> dict:
    exe = os.path.join(HERE, "_ref", "ref_fixed")
    if not os.path.exists(exe):
        sys.path.insert(0, HERE)
        import build_ref_fixed
        exe = build_ref_fixed.build()
    env = dict(os.environ)
    if dump:
        env["REF_FIXED_DUMP"] = dump
    r = subprocess.run([sys.executable, os.path.join(HERE, "mpi_shim", "mcp_mpirun.py"), "-np", str(np_),
                        "--timeout", str(timeout), exe], input=f"{g}\n{h}\n{w}\n".encode(), capture_output=True,
                       env=env, timeout=timeout + 60)
    recs = []
    for line in r.stdout.decode(errors="replace").splitlines():
        if line.startswith("{"):
            recs.append(json.loads(line))
    if r.returncode != 0 or len(recs) != np_:
        raise RuntimeError(f"ref-fixed failed (rc {r.returncode}, {len(recs)}/{np_} ranks reported)\n"
                           + r.stdout.decode(errors="replace")[-2000:] + r.stderr.decode(errors="replace")[-2000:])
    bad = [x for x in recs if x["cuda"] != "no error"]
    if bad:
        raise RuntimeError(f"CUDA error on rank {bad[0]['rank']}: {bad[0]['cuda']}")
    secs = max(x["seconds"] for x in recs)
    return {"impl": "ref-fixed", "n_gpus": np_, "grid": [h, w], "generations": g, "seconds": round(secs, 6),
            "glups": round(h * w * g / secs / 1e9, 6), "field_sum": sum(x["slab_sum"] for x in recs),
            "pattern": "per generation: 2 full-grid H2D + 1 full-grid D2H per rank, border rows as w-2 one-element "
                       "blocking messages (MDF_kernel.cu:161-183), wall clock between barriers"}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("-np", type=int, default=2)
    ap.add_argument("--generations", type=int, default=4)
    ap.add_argument("--height", type=int, default=16384)
    ap.add_argument("--width", type=int, default=16384)
    ap.add_argument("--dump", default=None)
    a = ap.parse_args()
    print(json.dumps(run(a.np, a.generations, a.height, a.width, a.dump)))
