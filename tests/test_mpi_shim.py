"""baseline/mpi_shim: the MPI stand-in works, and the reference compiles unmodified against it (compile only)."""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "baseline", "mpi_shim")

RING = r"""
#include <stdio.h>
#include <mpi.h>
int main(int argc, char** argv) {
    int g = 0; if (scanf("%d", &g) != 1) return 3;          /* every rank reads stdin, like the reference */
    int rank, size; MPI_Status st;
    MPI_Init(&argc, &argv); MPI_Comm_size(MPI_COMM_WORLD, &size); MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    float row[8], got[8];
    for (int i = 0; i < 8; ++i) row[i] = rank * 100.0f + i + g;
    int next = (rank + 1) % size, prev = (rank + size - 1) % size;
    for (int rep = 0; rep < 50; ++rep) {
        if (rank % 2 == 0) { for (int i = 0; i < 8; ++i) MPI_Send(&row[i], 1, MPI_FLOAT, next, 1, MPI_COMM_WORLD);
                             for (int i = 0; i < 8; ++i) MPI_Recv(&got[i], 1, MPI_FLOAT, prev, 1, MPI_COMM_WORLD, &st); }
        else               { for (int i = 0; i < 8; ++i) MPI_Recv(&got[i], 1, MPI_FLOAT, prev, 1, MPI_COMM_WORLD, &st);
                             for (int i = 0; i < 8; ++i) MPI_Send(&row[i], 1, MPI_FLOAT, next, 1, MPI_COMM_WORLD); }
        for (int i = 0; i < 8; ++i) if (got[i] != prev * 100.0f + i + g) return 4;
    }
    MPI_Barrier(MPI_COMM_WORLD);
    printf("rank %d of %d ok\n", rank, size);
    MPI_Finalize();
    return 0;
}
"""


def test_shim_ring_of_one_element_messages(tmp_path):
    src = tmp_path / "ring.c"
    src.write_text(RING)
    exe = tmp_path / "ring"
    subprocess.run(["gcc", "-O1", "-I", SHIM, str(src), os.path.join(SHIM, "mpi_shim.c"), "-o", str(exe)], check=True)
    r = subprocess.run([sys.executable, os.path.join(SHIM, "mcp_mpirun.py"), "-np", "4", "--timeout", "60", str(exe)],
                       input=b"7\n", capture_output=True, timeout=120)
    assert r.returncode == 0, r.stdout.decode() + r.stderr.decode()
    out = r.stdout.decode()
    assert all(f"rank {k} of 4 ok" in out for k in range(4))


@pytest.mark.skipif(shutil.which("nvcc") is None or not os.path.isdir("/root/reference"),
                    reason="needs nvcc and the reference tree")
@pytest.mark.parametrize("name", ["MDF_kernel.cu", "kernel.cu"])
def test_reference_compiles_unmodified_against_the_shim(tmp_path, name):
    obj = tmp_path / (name + ".o")
    r = subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-I", SHIM, "-c",
                        os.path.join("/root/reference", name), "-o", str(obj)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert obj.exists()
