"""Checkpoint / resume (absent in the reference): blocks + manifest, resumable with a different world size."""
import numpy as np

from mpi_cuda_process_b200 import oracle
from mpi_cuda_process_b200.config import RunConfig
from mpi_cuda_process_b200.solver import Solver
from mpi_cuda_process_b200.utils import checkpoint as ckpt

from _mp import run_ranks


def _save_worker(rank, world, cfg_json, gens, path):
    from mpi_cuda_process_b200.config import RunConfig
    from mpi_cuda_process_b200.solver import Solver
    from mpi_cuda_process_b200.utils import checkpoint as ckpt
    with Solver(RunConfig.from_json(cfg_json)) as s:
        s.initialize()
        s.run(gens)
        ckpt.save(s, path)
    return True


def test_save_two_ranks_resume_one_rank(tmp_path):
    cfg = RunConfig(workload="mdf", height=30, width=22, backend="cpu")
    run_ranks(_save_worker, 2, cfg.to_json(), 7, str(tmp_path))
    assert (tmp_path / "manifest.json").exists() and (tmp_path / "block_0001.npy").exists()
    with Solver(cfg) as s:
        assert ckpt.load(s, str(tmp_path)) == 7
        s.run(5)
        assert s.generation == 12
        assert s.gather().tobytes() == oracle.mdf_run(oracle.mdf_init(30, 22), 12).tobytes()


def test_checkpoint_life_roundtrip(tmp_path):
    cfg = RunConfig(workload="life", height=18, width=15, backend="cpu")
    with Solver(cfg) as s:
        s.initialize()
        s.run(4)
        ckpt.save(s, str(tmp_path))
    with Solver(cfg) as s:
        ckpt.load(s, str(tmp_path))
        s.run(3)
        assert (s.gather() == oracle.life_run(oracle.life_init_reference(18, 15), 7)).all()


def test_cli_checkpoint_and_resume(tmp_path):
    import subprocess, sys
    base = [sys.executable, "-m", "mpi_cuda_process_b200", "--workload", "mdf", "--backend", "cpu", "--no-prompt",
            "--height", "20", "--width", "20"]
    r = subprocess.run(base + ["-g", "6", "--checkpoint-dir", str(tmp_path), "--checkpoint-every", "3"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    out = tmp_path / "f.npy"
    r = subprocess.run(base + ["-g", "4", "--resume", str(tmp_path), "--output", str(out)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert np.load(out).tobytes() == oracle.mdf_run(oracle.mdf_init(20, 20), 10).tobytes()
