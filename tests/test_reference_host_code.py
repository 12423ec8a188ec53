"""The reference's OWN code as the oracle ("the reference's own CPU create_universe text compiles as
plain C++ -- usable directly at test time").

Nothing of the reference lives in this repository: at test time the text of ``create_universe`` and ``print_array``
(MDF_kernel.cu:72-99, kernel.cu:115-146) and of the two ``__device__`` update functions ``run_mdf`` and
``game_of_life`` (MDF_kernel.cu:10-22, kernel.cu:10-68; plain C inside) is cut out of the files under /root/reference,
compiled with g++ next to a small generated ``main`` and run; its output is compared with this framework's initial
conditions, writer and update rules -- bit for bit.  CPU only; skipped where the reference tree is absent (GPU boxes)."""
import os
import re
import shutil
import subprocess

import numpy as np
import pytest

from mpi_cuda_process_b200 import io as mio
from mpi_cuda_process_b200 import oracle

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF) or shutil.which("g++") is None,
                                reason="needs the reference tree and g++")


def _function_text(path: str, signature: str) -> str:
    """The text of one top-level function, from its signature to the matching closing brace."""
    src = open(path, "rb").read().decode("latin-1").replace("\r\n", "\n")
    start = src.index(signature)
    depth, i = 0, src.index("{", start)
    while True:
        depth += {"{": 1, "}": -1}.get(src[i], 0)
        i += 1
        if depth == 0:
            return src[start:i]


def _build(tmp_path, name: str, functions, main: str) -> str:
    code = "#include <stdio.h>\n#include <stdlib.h>\n" + "\n".join(functions) + "\n" + main
    src = tmp_path / f"{name}.cpp"
    src.write_text(code)
    exe = tmp_path / name
    subprocess.run(["g++", "-O1", "-w", str(src), "-o", str(exe)], check=True)
    return str(exe)


@pytest.mark.parametrize("h,w", [(3, 3), (5, 7), (16, 9)])
def test_mdf_initial_field_equals_the_references_create_universe(tmp_path, h, w):
    """Called with the arguments its signature asks for (w, h) -- the reference's own call passes (size, 0.15) and
    initialises nothing (MDF_kernel.cu:146 vs :88)."""
    fn = _function_text(os.path.join(REF, "MDF_kernel.cu"), "void create_universe(float *univ, int w, int h)")
    main = ("int main(int c, char** v) { int h = atoi(v[1]), w = atoi(v[2]); float* u = (float*)calloc(h * w, 4);\n"
            "  create_universe(u, w, h); fwrite(u, 4, h * w, stdout); return 0; }\n")
    exe = _build(tmp_path, "mdf_init", [fn], main)
    out = subprocess.run([exe, str(h), str(w)], capture_output=True, check=True).stdout
    ref = np.frombuffer(out, dtype=np.float32).reshape(h, w)
    assert ref.tobytes() == oracle.mdf_init(h, w).tobytes()
    from mpi_cuda_process_b200.config import RunConfig
    assert ref.tobytes() == oracle.init_field(RunConfig(workload="mdf", height=h, width=w)).tobytes()


@pytest.mark.parametrize("h,w", [(4, 4), (12, 10), (33, 21)])
def test_life_initial_field_equals_the_references_create_universe(tmp_path, h, w):
    """Unseeded glibc rand(), consumed in the reference's order (kernel.cu:131-146): the framework's host helper and
    its Python restatement reproduce the same universe cell for cell."""
    fn = _function_text(os.path.join(REF, "kernel.cu"), "void create_universe(int *univ, int h, int w, float prob)")
    main = ("int main(int c, char** v) { int h = atoi(v[1]), w = atoi(v[2]); int* u = (int*)calloc(h * w, 4);\n"
            "  create_universe(u, h, w, 0.15); fwrite(u, 4, h * w, stdout); return 0; }\n")
    exe = _build(tmp_path, "life_init", [fn], main)
    out = subprocess.run([exe, str(h), str(w)], capture_output=True, check=True).stdout
    ref = np.frombuffer(out, dtype=np.int32).reshape(h, w)
    assert (ref == oracle.life_init_reference(h, w)).all()
    from mpi_cuda_process_b200.ops import _lib
    mine = np.zeros((h, w), dtype=np.int32)
    import ctypes as C
    _lib.host_lib().mcp_host_life_init_reference(mine.ctypes.data, h, w, C.c_float(0.15))
    assert (ref == mine).all()


@pytest.mark.parametrize("which", ["kernel.cu", "MDF_kernel.cu"])
def test_print_array_bytes_equal_the_references(tmp_path, which):
    """The only output format the reference defines (kernel.cu:115-129, the same text at MDF_kernel.cu:72-86)."""
    fn = _function_text(os.path.join(REF, which), "void print_array(int arr[], int w, int size)")
    main = ("int main(int c, char** v) { int h = atoi(v[1]), w = atoi(v[2]); int* u = (int*)calloc(h * w, 4);\n"
            "  if (fread(u, 4, h * w, stdin) != (size_t)(h * w)) return 3; print_array(u, w, h * w); return 0; }\n")
    exe = _build(tmp_path, "print_" + re.sub(r"\W", "_", which), [fn], main)
    rng = np.random.default_rng(3)
    for h, w in ((1, 1), (6, 5), (9, 14)):
        u = rng.integers(0, 3, size=(h, w)).astype(np.int32)      # 0, 1 and a value that is neither
        out = subprocess.run([exe, str(h), str(w)], input=u.tobytes(), capture_output=True, check=True).stdout
        assert out == mio.format_array(u)


def test_life_run_printed_like_the_reference_would(tmp_path):
    """End to end through the reference's two host functions: its universe, advanced by this framework's CPU backend,
    printed by its print_array -- equals the framework's own CLI output path (format_array of the oracle run)."""
    fn_c = _function_text(os.path.join(REF, "kernel.cu"), "void create_universe(int *univ, int h, int w, float prob)")
    fn_p = _function_text(os.path.join(REF, "kernel.cu"), "void print_array(int arr[], int w, int size)")
    main = ("int main(int c, char** v) { int h = atoi(v[1]), w = atoi(v[2]); int* u = (int*)calloc(h * w, 4);\n"
            "  if (v[3][0] == 'i') { create_universe(u, h, w, 0.15); fwrite(u, 4, h * w, stdout); return 0; }\n"
            "  if (fread(u, 4, h * w, stdin) != (size_t)(h * w)) return 3; print_array(u, w, h * w); return 0; }\n")
    exe = _build(tmp_path, "life_io", [fn_p, fn_c], main)
    h, w, g = 14, 11, 6
    u0 = np.frombuffer(subprocess.run([exe, str(h), str(w), "i"], capture_output=True, check=True).stdout,
                       dtype=np.int32).reshape(h, w)
    from mpi_cuda_process_b200.config import RunConfig
    from mpi_cuda_process_b200.solver import Solver
    with Solver(RunConfig(workload="life", height=h, width=w, backend="cpu")) as s:
        s.initialize(u0.copy())
        s.run(g)
        mine = s.gather()
    printed = subprocess.run([exe, str(h), str(w), "p"], input=mine.astype(np.int32).tobytes(), capture_output=True,
                             check=True).stdout
    assert printed == mio.format_array(oracle.life_run(oracle.life_init_reference(h, w), g))


DEVICE_MAIN = r"""
int main(int c, char** v) {
    int h = atoi(v[1]), w = atoi(v[2]), size = h * w;
    CELL* u = (CELL*)calloc(size, sizeof(CELL)); CELL* out = (CELL*)calloc(size, sizeof(CELL));
    if (fread(u, sizeof(CELL), size, stdin) != (size_t)size) return 3;
    for (int i = 0; i < size; ++i) out[i] = u[i];
    for (int y = 1; y < h - 1; ++y)            /* interior cells only: the function has no edge handling (SURVEY D7) */
        for (int x = 1; x < w - 1; ++x) UPDATE(u, w, size, y * w + x, out);
    fwrite(out, sizeof(CELL), size, stdout);
    return 0;
}
"""


def _device_function_as_host_code(tmp_path, name, path, signature, cell, call):
    fn = _function_text(path, signature).replace("__device__", "")
    code = (f"#include <stdio.h>\n#include <stdlib.h>\n#define CELL {cell}\n#define UPDATE {call}\n" + fn + "\n" + DEVICE_MAIN)
    src = tmp_path / f"{name}.cpp"
    src.write_text(code)
    exe = tmp_path / name
    # no contraction, no fast math: the host compiler must evaluate the expression as written
    subprocess.run(["g++", "-O1", "-w", "-ffp-contract=off", str(src), "-o", str(exe)], check=True)
    return str(exe)


def test_mdf_update_equals_the_references_run_mdf_bit_for_bit(tmp_path):
    """MDF_kernel.cu:10-22 compiled as host C++ (the function body is plain C): fp32 sums, the double literal 0.25,
    one conversion back to float.  The framework's pinned rounding sequence (stencil_ops.cuh, oracle.mdf_update_cells)
    gives the same bits on every interior cell -- for smooth and for rough fields, over several generations."""
    exe = _device_function_as_host_code(tmp_path, "run_mdf", os.path.join(REF, "MDF_kernel.cu"),
                                        "__device__ void run_mdf(float *univ, int w, int size, int id, float *new_univ)",
                                        "float", "run_mdf")
    rng = np.random.default_rng(11)
    h, w = 19, 23
    fields = [oracle.mdf_init(h, w),
              rng.uniform(0, 100, (h, w)).astype(np.float32),
              (rng.standard_normal((h, w)) * 1e-3).astype(np.float32),
              rng.uniform(0, 1e-38, (h, w)).astype(np.float32)]          # subnormal sums
    for u in fields:
        for _ in range(4):
            out = subprocess.run([exe, str(h), str(w)], input=u.tobytes(), capture_output=True, check=True).stdout
            ref = np.frombuffer(out, dtype=np.float32).reshape(h, w)
            mine = oracle.mdf_step(u)
            assert ref[1:-1, 1:-1].tobytes() == mine[1:-1, 1:-1].tobytes()
            u = mine


def test_life_update_equals_the_references_game_of_life(tmp_path):
    """kernel.cu:10-68 compiled as host C++: every interior cell of a random universe, several generations."""
    exe = _device_function_as_host_code(tmp_path, "gol", os.path.join(REF, "kernel.cu"),
                                        "__device__ void game_of_life(int *univ, int w, int size, int id, int *new_univ)",
                                        "int", "game_of_life")
    h, w = 17, 22
    u = oracle.life_init_reference(h, w, prob=0.4)
    for _ in range(6):
        out = subprocess.run([exe, str(h), str(w)], input=u.astype(np.int32).tobytes(), capture_output=True,
                             check=True).stdout
        ref = np.frombuffer(out, dtype=np.int32).reshape(h, w)
        mine = oracle.life_step(u, bc="fixed-ring")
        assert (ref[1:-1, 1:-1] == mine[1:-1, 1:-1]).all()
        u = mine
