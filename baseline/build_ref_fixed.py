#!/usr/bin/env python
"""Build ``ref-fixed``: the reference's MDF program with the minimal repairs of ``ref_fixed.patch`` applied, linked
against the MPI stand-in of ``mpi_shim/``.

    python baseline/build_ref_fixed.py            ->  baseline/_ref/ref_fixed  (git-ignored; travels with gpurun)

The patch is kept as a zero-context unified diff against /root/reference/MDF_kernel.cu (CRLF preserved; the
unmodified file is the only thing it can apply to, so no context lines are carried).  What it changes, and nothing
else: the ``create_universe`` call (MDF_kernel.cu:146 vs :88), one loop for every rank with mirrored peers and offsets
(:155-222; N row slabs instead of the hard-wired two), the new field is what comes back from the GPU (:164,177), the
Dirichlet ring is held and the out-of-bounds neighbour reads are gone (:13-20,30-31), cells that are not a rank's own
are no longer overwritten with 100 (:34-36), ``cudaSetDevice`` per rank (:114-121), ceil instead of floor for the
grid (:149), and a timer + JSON line.  The communication pattern is the reference's: the WHOLE grid goes host->device
twice and device->host once per generation, and every border row travels as w-2 blocking one-element messages.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("MCP_REFERENCE_DIR", "/root/reference")
OUT = os.path.join(HERE, "_ref")


def build(verbose: bool = False) -> str:
    src = os.path.join(REF, "MDF_kernel.cu")
    if not os.path.exists(src):
        raise FileNotFoundError(f"{src}: the reference tree is needed to build ref-fixed")
    os.makedirs(OUT, exist_ok=True)
    dst = os.path.join(OUT, "MDF_kernel_fixed.cu")
    shutil.copyfile(src, dst)
    r = subprocess.run(["patch", "--binary", "-s", dst, os.path.join(HERE, "ref_fixed.patch")],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("ref_fixed.patch does not apply:\n" + r.stdout + r.stderr)
    exe = os.path.join(OUT, "ref_fixed")
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-w", "-I", os.path.join(HERE, "mpi_shim"),
           "-o", exe, dst, os.path.join(HERE, "mpi_shim", "mpi_shim.c")]
    if verbose:
        print("+", " ".join(cmd))
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("ref-fixed does not compile:\n" + r.stdout + r.stderr)
    return exe


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
