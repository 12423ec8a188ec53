import ctypes as C
import io
import subprocess
import sys

import numpy as np
import pytest

from mpi_cuda_process_b200 import io as mio, oracle
from mpi_cuda_process_b200.cli import PROMPTS, read_ghw
from mpi_cuda_process_b200.config import RunConfig
from mpi_cuda_process_b200.ops import _lib


def test_format_array_matches_cpp_print_array():
    u = oracle.life_init_reference(9, 13)
    cap = 2 + 9 * 14
    buf = C.create_string_buffer(cap)
    n = _lib.host_lib().mcp_host_print_array(u.ctypes.data, 13, u.size, buf, cap)
    assert n == cap
    assert buf.raw[:n] == mio.format_array(u)
    assert (mio.parse_array(mio.format_array(u), 13) == u).all()


def test_format_array_layout():
    u = np.array([[1, 0], [2, 1]], dtype=np.int32)
    assert mio.format_array(u) == b"\n0 \n 0\n\n"      # only cells equal to 1 print as '0' (kernel.cu:121)


def test_read_ghw_prompts_and_scanf_semantics():
    out = io.StringIO()
    assert read_ghw(io.StringIO("100 256\n256\n"), out) == [100, 256, 256]
    assert out.getvalue() == "".join(PROMPTS)
    assert PROMPTS[0] == "Enter desired number of generations:\n"
    with pytest.raises(EOFError):
        read_ghw(io.StringIO("1 2"), io.StringIO())


def test_dump_and_load(tmp_path):
    u = oracle.mdf_run(oracle.mdf_init(6, 5), 3)
    for fmt in ("npy", "raw", "text"):
        p = str(tmp_path / f"f.{fmt}")
        mio.dump_field(u, p + (".npy" if fmt == "npy" else ""), fmt)
    assert (np.load(str(tmp_path / "f.npy.npy")) == u).all()
    assert (mio.load_field(str(tmp_path / "f.raw"), "raw", u.shape, np.float32) == u).all()
    with pytest.raises(ValueError):
        mio.dump_field(u, str(tmp_path / "x"), "ascii")   # print_array cannot take the float field


def test_config_validation():
    with pytest.raises(ValueError):
        RunConfig(workload="nope")
    with pytest.raises(ValueError):
        RunConfig(workload="heat3d", depth=0)
    c = RunConfig(workload="life")
    assert c.dtype == "int32"
    c2 = RunConfig.from_json(RunConfig(workload="mdf", height=8, width=9, tol=1e-3).to_json())
    assert c2.tol == 1e-3 and c2.residual == "linf" and c2.shape == (8, 9)


def test_cli_stdin_compatible_cpu():
    """echo "20 32 32" | python -m mpi_cuda_process_b200 --backend cpu --check  (SURVEY 7.3 minimum slice)"""
    r = subprocess.run([sys.executable, "-m", "mpi_cuda_process_b200", "--workload", "mdf", "--backend", "cpu",
                        "--check"], input="20 32 32\n", capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert r.stdout.startswith("".join(PROMPTS))
    assert "bitwise equal after 20 generations" in r.stdout


def test_cli_life_print(tmp_path):
    out = tmp_path / "life.txt"
    r = subprocess.run([sys.executable, "-m", "mpi_cuda_process_b200", "--workload", "life", "--backend", "cpu",
                        "--no-prompt", "-g", "3", "--height", "12", "--width", "10", "--output", str(out), "--print"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    ref = oracle.life_run(oracle.life_init_reference(12, 10), 3)
    assert out.read_bytes() == mio.format_array(ref)
    assert mio.format_array(ref).decode() in r.stdout


def test_cli_logs_jsonl(tmp_path):
    import json, subprocess, sys
    log = tmp_path / "run.jsonl"
    r = subprocess.run([sys.executable, "-m", "mpi_cuda_process_b200", "--workload", "mdf", "--backend", "cpu", "--no-prompt",
                        "-g", "3", "--height", "12", "--width", "10", "--residual", "linf", "--log-jsonl", str(log)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    rec = json.loads(log.read_text().splitlines()[-1])
    assert rec["workload"] == "mdf" and rec["generations"] == 3 and "ts" in rec


def test_cli_continues_from_an_input_file(tmp_path):
    """--input: start from a field written by --output (the reference can only start from create_universe,
    MDF_kernel.cu:146); 3 + 4 generations through a print_array file equal 7 generations of the oracle."""
    first, second = tmp_path / "g3.txt", tmp_path / "g7.txt"
    base = [sys.executable, "-m", "mpi_cuda_process_b200", "--workload", "life", "--backend", "cpu", "--no-prompt",
            "--height", "12", "--width", "10"]
    r = subprocess.run(base + ["-g", "3", "--output", str(first)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    r = subprocess.run(base + ["-g", "4", "--input", str(first), "--output", str(second), "--check"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "bitwise equal after 4 generations" in r.stdout
    assert second.read_bytes() == mio.format_array(oracle.life_run(oracle.life_init_reference(12, 10), 7))
    # float field through .npy; a file of the wrong shape is refused
    npy = tmp_path / "u.npy"
    np.save(npy, oracle.mdf_run(oracle.mdf_init(9, 11), 2))
    mdf = [sys.executable, "-m", "mpi_cuda_process_b200", "--workload", "mdf", "--backend", "cpu", "--no-prompt",
           "-g", "5", "--input", str(npy), "--output", str(tmp_path / "u7.npy")]
    r = subprocess.run(mdf + ["--height", "9", "--width", "11"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert np.load(tmp_path / "u7.npy").tobytes() == oracle.mdf_run(oracle.mdf_init(9, 11), 7).tobytes()
    r = subprocess.run(mdf + ["--height", "10", "--width", "11"], capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and "shape" in (r.stderr + r.stdout)


def test_cli_prints_and_writes_frames_every_n_generations(tmp_path):
    """--print-every: the reference's per-generation print_array calls (commented out in its loop, kernel.cu:232,265);
    --output-every: the same as files."""
    out = tmp_path / "life.txt"
    r = subprocess.run([sys.executable, "-m", "mpi_cuda_process_b200", "--workload", "life", "--backend", "cpu",
                        "--no-prompt", "-g", "5", "--height", "9", "--width", "8", "--print-every", "2",
                        "--output-every", "2", "--output", str(out), "--check"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    u0 = oracle.life_init_reference(9, 8)
    frames = [mio.format_array(oracle.life_run(u0, g)).decode() for g in (2, 4, 5)]
    pos = [r.stdout.index(f) for f in frames]            # generations 2 and 4 in the loop, 5 at the end, in order
    assert pos == sorted(pos)
    assert (tmp_path / "life.g000002.txt").read_bytes() == frames[0].encode()
    assert (tmp_path / "life.g000004.txt").read_bytes() == frames[1].encode()
    assert out.read_bytes() == frames[2].encode()
    assert "bitwise equal after 5 generations" in r.stdout
