"""CPU plumbing path: torch CPU + gloo, equal to the single-domain oracle bit for bit.

Covers BASELINE.json config 1 (2-D MDF, 256x256, 100 iterations, world_size 2, CPU) and exactly the case
where the reference hangs: g >= 2 on two ranks (MDF_kernel.cu:168 vs :215).
"""
import numpy as np
import pytest

from mpi_cuda_process_b200 import oracle
from mpi_cuda_process_b200.config import RunConfig
from mpi_cuda_process_b200.solver import Solver

from _mp import run_ranks


def _solve(rank, world, cfg_json, gens, residual=False):
    from mpi_cuda_process_b200.config import RunConfig
    from mpi_cuda_process_b200.solver import Solver
    cfg = RunConfig.from_json(cfg_json)
    with Solver(cfg) as s:
        s.initialize()
        st = s.run(gens, residual=residual)
        out = s.gather()
        return out, st.halo_messages, st.residual_history


@pytest.mark.parametrize("shape", [(3, 3), (8, 5), (33, 47)])
def test_single_rank_mdf(shape):
    cfg = RunConfig(workload="mdf", height=shape[0], width=shape[1], backend="cpu")
    with Solver(cfg) as s:
        s.initialize()
        s.run(7)
        assert s.gather().tobytes() == oracle.mdf_run(oracle.mdf_init(*shape), 7).tobytes()


@pytest.mark.parametrize("bc", ["dead-outside", "fixed-ring"])
def test_single_rank_life(bc):
    cfg = RunConfig(workload="life", height=20, width=17, backend="cpu", bc=bc)
    with Solver(cfg) as s:
        s.initialize()
        s.run(9)
        assert (s.gather() == oracle.life_run(oracle.life_init_reference(20, 17), 9, bc)).all()


def test_single_rank_heat3d():
    cfg = RunConfig(workload="heat3d", depth=6, height=7, width=9, backend="cpu")
    with Solver(cfg) as s:
        s.initialize()
        s.run(5)
        assert s.gather().tobytes() == oracle.heat3d_run(oracle.heat3d_init(6, 7, 9), 5).tobytes()


def test_config1_mdf_256_100_world2():
    cfg = RunConfig(workload="mdf", height=256, width=256, backend="cpu")
    res = run_ranks(_solve, 2, cfg.to_json(), 100)
    ref = oracle.mdf_run(oracle.mdf_init(256, 256), 100)
    assert res[0][0].tobytes() == ref.tobytes()
    assert res[1][0] is None
    # first generation needs no exchange (MDF_kernel.cu:166); afterwards one row message per generation
    assert res[0][1] == 99


@pytest.mark.parametrize("world,g", [(2, 1), (2, 2), (2, 3), (3, 10), (4, 25)])
def test_mdf_multi_rank_generations(world, g):
    cfg = RunConfig(workload="mdf", height=37, width=19, backend="cpu")
    res = run_ranks(_solve, world, cfg.to_json(), g)
    assert res[0][0].tobytes() == oracle.mdf_run(oracle.mdf_init(37, 19), g).tobytes()


def test_mdf_element_messages_like_the_reference():
    """w-2 one-element messages per border row per generation (MDF_kernel.cu:167-169,181-183)."""
    cfg = RunConfig(workload="mdf", height=12, width=10, backend="cpu", halo_message="element")
    res = run_ranks(_solve, 2, cfg.to_json(), 4)
    assert res[0][0].tobytes() == oracle.mdf_run(oracle.mdf_init(12, 10), 4).tobytes()
    assert res[0][1] == 3 * (10 - 2)


@pytest.mark.parametrize("bc", ["dead-outside", "fixed-ring"])
def test_life_glider_crosses_the_cut(bc):
    h, w = 24, 16
    u = np.zeros((h, w), dtype=np.int32)
    u[8:11, 5:8] = np.array([[0, 1, 0], [0, 0, 1], [1, 1, 1]])   # glider heading down-right across row 12
    res = run_ranks(_glider_worker, 2, u, bc)
    assert (res[0] == oracle.life_run(u, 20, bc)).all()
    assert res[0][12:].sum() > 0


def _glider_worker(rank, world, field, bc):
    from mpi_cuda_process_b200.config import RunConfig
    from mpi_cuda_process_b200.solver import Solver
    cfg = RunConfig(workload="life", height=24, width=16, backend="cpu", bc=bc)
    with Solver(cfg) as s:
        s.initialize(field)
        s.run(20)
        return s.gather()


def test_heat3d_pencil_4_ranks():
    cfg = RunConfig(workload="heat3d", depth=8, height=10, width=7, backend="cpu", decomposition="pencil", pencil=(2, 2))
    res = run_ranks(_solve, 4, cfg.to_json(), 6)
    assert res[0][0].tobytes() == oracle.heat3d_run(oracle.heat3d_init(8, 10, 7), 6).tobytes()


def test_residual_and_tolerance_world2():
    cfg = RunConfig(workload="mdf", height=16, width=16, backend="cpu", residual="linf")
    res = run_ranks(_solve, 2, cfg.to_json(), 5, True)
    u = oracle.mdf_init(16, 16)
    exp = []
    for _ in range(5):
        v = oracle.mdf_step(u)
        exp.append(oracle.residual(v, u, "linf"))
        u = v
    got = [r for _, r in res[0][2]]
    assert got == exp
    assert res[1][2] == res[0][2]          # same bits on every rank


def test_run_to_tolerance_cpu():
    cfg = RunConfig(workload="mdf", height=12, width=12, backend="cpu", tol=1e-2, max_generations=2000, check_every=10)
    with Solver(cfg) as s:
        s.initialize()
        st = s.run_to_tolerance()
        assert st.residual_history[-1][1] <= 1e-2
        assert s.generation < 2000 and s.generation % 10 == 0
        assert st.stop_reason == "converged"
    with Solver(RunConfig(workload="mdf", height=12, width=12, backend="cpu", tol=1e-9, max_generations=20)) as s:
        s.initialize()
        assert s.run_to_tolerance().stop_reason == "max_generations"


def test_checksum_is_decomposition_independent():
    from mpi_cuda_process_b200 import oracle as orc
    f = orc.mdf_run(orc.mdf_init(23, 17), 6)
    whole = orc.checksum_field(f)
    parts = (orc.checksum_rows(f[:5], 0, 17) + orc.checksum_rows(f[5:16], 5, 17) + orc.checksum_rows(f[16:], 16, 17))
    assert whole == parts & 0xFFFFFFFFFFFFFFFF
    g = f.copy()
    g[11, 3] = np.nextafter(g[11, 3], np.float32(1e9))
    assert orc.checksum_field(g) != whole
    cfg = RunConfig(workload="mdf", height=23, width=17, backend="cpu")
    with Solver(cfg) as s:
        s.initialize()
        s.run(6)
        assert s.checksum() == whole


def test_step_is_one_trip_through_the_reference_loop():
    """Solver.step(): one generation (MDF_kernel.cu:157-187); n steps equal run(n) and count generations."""
    cfg = RunConfig(workload="mdf", height=11, width=13, backend="cpu", residual="linf")
    with Solver(cfg) as a, Solver(cfg) as b:
        a.initialize()
        b.initialize()
        for _ in range(5):
            st = a.step(residual=True)
        b.run(5)
        assert a.generation == b.generation == 5 and st.generations == 1 and len(st.residual_history) == 1
        assert a.gather().tobytes() == b.gather().tobytes() == oracle.mdf_run(oracle.mdf_init(11, 13), 5).tobytes()
