"""The native layer used from plain C (examples/solve_mdf.c): no Python, no CUDA headers, the reference's stdin interface."""
import os
import re
import shutil
import subprocess

import numpy as np
import pytest

from mpi_cuda_process_b200 import oracle
from mpi_cuda_process_b200.ops import build

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs gcc")
def test_c_example_matches_oracle(tmp_path):
    libdir = str(build.lib_path().parent)
    exe = str(tmp_path / "solve_mdf")
    subprocess.run(["gcc", "-O2", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "examples", "solve_mdf.c"),
                    "-L", libdir, "-lmcp_b200", f"-Wl,-rpath,{libdir}", "-o", exe], check=True)
    g, h, w = 100, 200, 300
    r = subprocess.run([exe], input=f"{g} {h} {w}\n", capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("Enter desired number of generations:\nEnter desired height of universe:\n"
                               "Enter desired width of universe:\n")
    m = re.search(r"u\[\d+\]\[1\] = (\S+), u\[\d+\]\[\d+\] = (\S+), last max\|du\| = (\S+) \((\d+) launches\)", r.stdout)
    assert m, r.stdout
    u = oracle.mdf_init(h, w)
    prev = u
    for _ in range(g):
        prev, u = u, oracle.mdf_step(u)
    assert np.float32(float(m.group(1).rstrip(","))) == u[h // 2, 1]
    assert np.float32(float(m.group(2).rstrip(","))) == u[h // 2, w // 2]
    assert np.float32(float(m.group(3))) == np.float32(oracle.residual(u, prev, "linf"))   # printed with 9 digits
