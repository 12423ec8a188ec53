"""Small single-GPU run of every kernel family, meant to be executed under compute-sanitizer."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_cuda_process_b200 import RunConfig, Solver, oracle  # noqa: E402

cases = [
    dict(workload="mdf", height=70, width=150, backend="fused", time_block=8),
    dict(workload="mdf", height=70, width=150, backend="fused", time_block=2, dtype="fp64", residual="linf"),
    dict(workload="life", height=60, width=90, backend="fused", time_block=4),
    dict(workload="mdf", height=50, width=70, backend="nccl"),
    dict(workload="heat3d", depth=20, height=40, width=70, backend="fused"),
]
for c in cases:
    cfg = RunConfig(**c)
    with Solver(cfg) as s:
        s.initialize()
        s.run(9, residual=cfg.residual != "none")
        s.check_status()
        got = s.gather()
    ref = oracle.run_field(cfg, oracle.init_field(cfg), 9)
    assert got.tobytes() == ref.tobytes(), c
print("sanitize_case ok")
