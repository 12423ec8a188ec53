"""In-tree build of the native layer.

``libmcp_b200.so``  all sm_100a kernels + the C-ABI runtime (nvcc, ``-gencode arch=compute_100a,code=sm_100a``)
``libmcp_host.so``  the CPU reference subset (g++ only; usable on a box without CUDA)

The reference ships no build system at all (no Makefile, no nvcc line, no ``-arch``), so
its kernels would be JIT-compiled from sm_52 PTX on a B200.  Here the arch is pinned and the
shared objects are built next to the sources so that they travel with the tree.
"""
from __future__ import annotations

import contextlib
import fcntl
import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
LIBDIR = HERE / "_lib"
OBJDIR = HERE / "_build"
INCLUDE = HERE.parent.parent / "include"

CUDA_SOURCES = [
    "runtime.cu",
    "init_bc.cu",
    "stencil_plain.cu",
    "stencil2d_fused.cu",
    "heat3d_fused.cu",
    "heat3d_cluster.cu",
    "residual.cu",
    "residual_noalloc.cu",
    "gather_pack.cu",
    "checksum.cu",
    "host_ref.cpp",
]
HOST_SOURCES = ["host_ref.cpp"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off",
    "--fmad=false",   # contraction is always explicit (__fmaf_rn) in the kernels
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _sources_hash(names) -> str:
    h = hashlib.sha256()
    deps = sorted(CSRC.glob("*.cuh")) + sorted(INCLUDE.glob("*.h")) + [CSRC / n for n in names] + [Path(__file__)]
    for p in deps:
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def lib_path() -> Path:
    return LIBDIR / "libmcp_b200.so"


def host_lib_path() -> Path:
    return LIBDIR / "libmcp_host.so"


@contextlib.contextmanager
def _build_lock():
    """One builder at a time per tree: under torchrun every rank imports the package at once, and a rank must
    never load a half-written shared object (outputs are linked under a temporary name and renamed)."""
    LIBDIR.mkdir(exist_ok=True)
    with open(LIBDIR / ".build.lock", "w") as f:
        fcntl.flock(f, fcntl.LOCK_EX)
        try:
            yield
        finally:
            fcntl.flock(f, fcntl.LOCK_UN)


def _run(cmd, verbose):
    if verbose:
        print("+", " ".join(str(c) for c in cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"command failed ({r.returncode}): {' '.join(str(c) for c in cmd)}\n{r.stdout}\n{r.stderr}")
    if verbose and (r.stdout.strip() or r.stderr.strip()):
        print(r.stdout, r.stderr, flush=True)
    return r


def _obj_stale(src: Path, obj: Path, stamp: str) -> bool:
    tag = obj.with_suffix(obj.suffix + ".hash")
    return not (obj.exists() and tag.exists() and tag.read_text() == stamp)


def build_cuda(force: bool = False, verbose: bool = False, ptxas_info: bool = False) -> Path:
    """Build (or confirm up to date) libmcp_b200.so.  Cheap when nothing changed: one hash over the sources."""
    out = lib_path()
    stamp_all = _sources_hash(CUDA_SOURCES)
    stamp_file = LIBDIR / "libmcp_b200.hash"
    if not force and out.exists() and stamp_file.exists() and stamp_file.read_text() == stamp_all:
        return out
    with _build_lock():
        if not force and out.exists() and stamp_file.exists() and stamp_file.read_text() == stamp_all:
            return out          # another process built it while we waited for the lock
        return _build_cuda_locked(out, stamp_all, stamp_file, force, verbose, ptxas_info)


def _build_cuda_locked(out: Path, stamp_all: str, stamp_file: Path, force: bool, verbose: bool,
                       ptxas_info: bool) -> Path:
    OBJDIR.mkdir(exist_ok=True)
    nvcc = _nvcc()
    headers_stamp = hashlib.sha256(
        b"".join(p.read_bytes() for p in sorted(CSRC.glob("*.cuh")) + sorted(INCLUDE.glob("*.h")))
        + " ".join(NVCC_FLAGS).encode()).hexdigest()

    def compile_one(name: str) -> Path:
        src = CSRC / name
        obj = OBJDIR / (name + ".o")
        stamp = hashlib.sha256(src.read_bytes()).hexdigest() + headers_stamp
        if force or _obj_stale(src, obj, stamp):
            cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-c", str(src), "-o", str(obj)]
            if src.suffix == ".cpp":
                cmd[1:1] = ["-x", "cu"]
            if ptxas_info:
                cmd[1:1] = ["-Xptxas", "-v"]
            r = _run(cmd, verbose)
            if ptxas_info:
                (OBJDIR / (name + ".ptxas.txt")).write_text(r.stderr)
            obj.with_suffix(obj.suffix + ".hash").write_text(stamp)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(CUDA_SOURCES))) as ex:
        objs = list(ex.map(compile_one, CUDA_SOURCES))
    tmp = out.with_suffix(f".so.tmp{os.getpid()}")
    _run([nvcc, "-shared", "-o", str(tmp), *[str(o) for o in objs], "-gencode", "arch=compute_100a,code=sm_100a",
          "-Xcompiler", "-fPIC"], verbose)
    os.replace(tmp, out)
    stamp_file.write_text(stamp_all)
    return out


def build_host(force: bool = False, verbose: bool = False) -> Path:
    out = host_lib_path()
    stamp = _sources_hash(HOST_SOURCES)
    stamp_file = LIBDIR / "libmcp_host.hash"
    if not force and out.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return out
    with _build_lock():
        if not force and out.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
            return out
        return _build_host_locked(out, stamp, stamp_file, verbose)


def _build_host_locked(out: Path, stamp: str, stamp_file: Path, verbose: bool) -> Path:
    final = out
    out = final.with_suffix(f".so.tmp{os.getpid()}")
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    base = ["-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", *[str(CSRC / n) for n in HOST_SOURCES],
            "-o", str(out)]
    # OpenMP where the toolchain has it (some g++ installs ship without libgomp): rows are independent, results are
    # bit-identical either way
    attempts = [[cxx, "-fopenmp", *base], [cxx, *base]]
    if shutil.which("g++") and shutil.which("g++") != cxx:
        attempts.insert(1, [shutil.which("g++"), "-fopenmp", *base])
    err = None
    for cmd in attempts:
        try:
            _run(cmd, verbose)
            err = None
            break
        except RuntimeError as e:
            err = e
    if err is not None:
        raise err
    os.replace(out, final)
    stamp_file.write_text(stamp)
    return final


def build_all(force: bool = False, verbose: bool = False, ptxas_info: bool = False):
    return build_cuda(force, verbose, ptxas_info), build_host(force, verbose)


if __name__ == "__main__":
    force = "--force" in sys.argv
    libs = build_all(force=force, verbose="-v" in sys.argv, ptxas_info="--ptxas" in sys.argv)
    for p in libs:
        print(p)
