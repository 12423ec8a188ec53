"""ref-fixed: the reference's MDF program + baseline/ref_fixed.patch on the MPI stand-in.

CPU: the patch applies to the unmodified reference source and the result compiles for sm_100a.
GPU: the patched program, started as 1, 2 and 3 ranks, reproduces the oracle's field bit for bit -- for g >= 2, which is
exactly where the unpatched reference deadlocks (MDF_kernel.cu:168 vs :215)."""
import os
import shutil
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline"))

HAVE_REF = os.path.isdir("/root/reference") and shutil.which("nvcc") is not None and shutil.which("patch") is not None
EXE = os.path.join(ROOT, "baseline", "_ref", "ref_fixed")


@pytest.mark.skipif(not HAVE_REF, reason="needs the reference tree, nvcc and patch")
def test_patch_applies_and_compiles():
    import build_ref_fixed
    exe = build_ref_fixed.build()
    assert os.path.exists(exe)
    src = open(os.path.join(ROOT, "baseline", "_ref", "MDF_kernel_fixed.cu"), "rb").read()
    assert b"create_universe(h_univ, w, h)" in src and b"\r\n" in src       # patched, CRLF preserved


@pytest.mark.gpu
@pytest.mark.skipif(not (os.path.exists(EXE) or HAVE_REF), reason="ref_fixed binary not built and no reference tree")
@pytest.mark.parametrize("ranks", [1, 2, 3])
def test_ref_fixed_matches_oracle(tmp_path, ranks):
    import run_ref_fixed
    from mpi_cuda_process_b200 import oracle
    g, h, w = 7, 45, 70
    dump = str(tmp_path / "field")
    rec = run_ref_fixed.run(ranks, g, h, w, dump=dump, timeout=120)
    assert rec["n_gpus"] == ranks and rec["glups"] > 0
    parts = [np.fromfile(f"{dump}.{r}", dtype=np.float32) for r in range(ranks)]
    got = np.concatenate(parts).reshape(h, w)
    ref = oracle.mdf_run(oracle.mdf_init(h, w), g)
    assert got.tobytes() == ref.tobytes()
