"""GPU numerics: every sm_100a kernel against the plain NumPy/PyTorch reference of the same op (oracle.py)."""
import ctypes as C

import numpy as np
import pytest
import torch

from mpi_cuda_process_b200 import oracle
from mpi_cuda_process_b200.config import RunConfig
from mpi_cuda_process_b200.solver import Solver

pytestmark = pytest.mark.gpu


def _lib():
    from mpi_cuda_process_b200.ops import _lib
    return _lib, _lib.cuda_lib()


def test_native_library_is_loaded():
    from mpi_cuda_process_b200.ops import _lib
    lib = _lib.cuda_lib()
    assert lib.mcp_version() >= 100
    sm, maj, mnr = C.c_int(), C.c_int(), C.c_int()
    mem = C.c_size_t()
    _lib.check(lib.mcp_device_info(C.byref(sm), C.byref(maj), C.byref(mnr), C.byref(mem)))
    assert sm.value > 0
    with open("/proc/self/maps") as f:
        assert "libmcp_b200.so" in f.read()


@pytest.mark.parametrize("h,w", [(3, 3), (5, 4), (64, 96), (257, 130), (100, 1000)])
def test_device_init_ring(h, w):
    with Solver(RunConfig(workload="mdf", height=h, width=w, backend="fused")) as s:
        s.initialize()
        assert s.gather().tobytes() == oracle.mdf_init(h, w).tobytes()
        # pad columns and out-of-domain ghost rows are zero
        full = s.bufs[0].cpu().numpy()
        assert (full[:s.ghost] == 0).all() and (full[:, w:] == 0).all()


def test_device_init_life_hash():
    cfg = RunConfig(workload="life", height=77, width=91, backend="fused", seed=5)
    with Solver(cfg) as s:
        s.initialize()
        assert (s.gather() == oracle.life_init_hash(77, 91, 0.15, 5)).all()


@pytest.mark.parametrize("dtype", ["fp32", "fp64"])
@pytest.mark.parametrize("h,w,g", [(3, 3, 2), (9, 7, 5), (130, 259, 17), (512, 512, 33)])
def test_plain_mdf_bitwise(dtype, h, w, g):
    cfg = RunConfig(workload="mdf", height=h, width=w, backend="nccl", dtype=dtype)
    npdt = np.float32 if dtype == "fp32" else np.float64
    with Solver(cfg) as s:
        s.initialize()
        s.run(g)
        assert s.gather().tobytes() == oracle.mdf_run(oracle.mdf_init(h, w, dtype=npdt), g).tobytes()


def test_plain_mdf_strict_fp64_accum_is_bit_identical():
    a = RunConfig(workload="mdf", height=200, width=300, backend="nccl")
    b = RunConfig(workload="mdf", height=200, width=300, backend="nccl", strict_fp64_accum=True)
    outs = []
    rng = np.random.default_rng(0)
    f = oracle.mdf_init(200, 300)
    f[1:-1, 1:-1] = rng.uniform(0, 100, (198, 298)).astype(np.float32)
    for cfg in (a, b):
        with Solver(cfg) as s:
            s.initialize(f)
            s.run(25)
            outs.append(s.gather())
    assert outs[0].tobytes() == outs[1].tobytes() == oracle.mdf_run(f, 25).tobytes()


@pytest.mark.parametrize("tb", [1, 2, 4, 8])
@pytest.mark.parametrize("h,w,g", [(3, 3, 3), (16, 16, 8), (37, 130, 13), (300, 500, 24), (1030, 770, 41)])
def test_fused_mdf_f32_bitwise(tb, h, w, g):
    cfg = RunConfig(workload="mdf", height=h, width=w, backend="fused", time_block=tb)
    with Solver(cfg) as s:
        s.initialize()
        s.run(g)
        s.check_status()
        assert s.gather().tobytes() == oracle.mdf_run(oracle.mdf_init(h, w), g).tobytes()


def test_fused_mdf_random_field_many_chunks():
    h, w, g = 700, 333, 19
    rng = np.random.default_rng(4)
    f = oracle.mdf_init(h, w)
    f[1:-1, 1:-1] = rng.uniform(-50, 150, (h - 2, w - 2)).astype(np.float32)
    cfg = RunConfig(workload="mdf", height=h, width=w, backend="fused", time_block=4, extra={"chunk_rows": 48})
    with Solver(cfg) as s:
        s.initialize(f)
        s.run(g)
        s.check_status()
        assert s.gather().tobytes() == oracle.mdf_run(f, g).tobytes()


@pytest.mark.parametrize("tb", [1, 2, 4, 8])
def test_fused_mdf_f64_bitwise(tb):
    h, w, g = 211, 390, 14
    cfg = RunConfig(workload="mdf", height=h, width=w, backend="fused", time_block=tb, dtype="fp64")
    with Solver(cfg) as s:
        s.initialize()
        s.run(g)
        s.check_status()
        assert s.gather().tobytes() == oracle.mdf_run(oracle.mdf_init(h, w, dtype=np.float64), g).tobytes()


@pytest.mark.parametrize("backend", ["fused", "nccl"])
@pytest.mark.parametrize("bc", ["dead-outside", "fixed-ring"])
@pytest.mark.parametrize("tb", [1, 4])
def test_life_exact(backend, bc, tb):
    h, w, g = 150, 275, 23
    cfg = RunConfig(workload="life", height=h, width=w, backend=backend, bc=bc, time_block=tb)
    with Solver(cfg) as s:
        s.initialize()
        s.run(g)
        assert (s.gather() == oracle.life_run(oracle.life_init_reference(h, w), g, bc)).all()


def test_life_full_density_edges():
    """dead-outside: edge cells evolve and cells beyond the grid (pad columns, ghost rows) stay dead."""
    h, w, g = 40, 70, 12
    f = (np.random.default_rng(9).random((h, w)) < 0.5).astype(np.int32)
    cfg = RunConfig(workload="life", height=h, width=w, backend="fused", time_block=4)
    with Solver(cfg) as s:
        s.initialize(f)
        s.run(g)
        assert (s.gather() == oracle.life_run(f, g)).all()


@pytest.mark.parametrize("backend", ["nccl", "fused"])
@pytest.mark.parametrize("dtype", ["fp32", "fp64"])
@pytest.mark.parametrize("d,h,w,g", [(21, 34, 70, 6), (3, 3, 3, 2), (70, 67, 300, 5), (9, 130, 129, 7)])
def test_heat3d_bitwise(backend, dtype, d, h, w, g):
    npdt = np.float32 if dtype == "fp32" else np.float64
    cfg = RunConfig(workload="heat3d", depth=d, height=h, width=w, backend=backend, dtype=dtype,
                    extra={"zchunk": 16})
    with Solver(cfg) as s:
        s.initialize()
        s.run(g)
        assert s.gather().tobytes() == oracle.heat3d_run(oracle.heat3d_init(d, h, w, dtype=npdt), g).tobytes()


def test_fused_residual_matches_oracle():
    h, w, tb = 260, 390, 4
    cfg = RunConfig(workload="mdf", height=h, width=w, backend="fused", time_block=tb, residual="linf")
    with Solver(cfg) as s:
        s.initialize()
        st = s.run(12, residual=True)
        u = oracle.mdf_init(h, w)
        exp = []
        for gen in range(1, 13):
            v = oracle.mdf_step(u)
            if gen % tb == 0:
                exp.append((oracle.residual(v, u, "linf"), oracle.residual(v, u, "l2")))
            u = v
        got = [r for _, r in st.residual_history]
        assert got == [e[0] for e in exp]                      # max |d| is exact
        ss, mx = s._last_residual
        assert abs(np.sqrt(ss) - exp[-1][1]) <= 1e-6 * exp[-1][1]     # 4-cell groups are squared and summed in fp32
        # deterministic: same bits on a second run from the same state
        s.initialize()
        st2 = s.run(12, residual=True)
        assert st2.residual_history == st.residual_history and s._last_residual == (ss, mx)


def test_plain_residual_kernel():
    lib_mod, lib = _lib()
    h, w, pitch = 97, 130, 160
    rng = np.random.default_rng(1)
    a = torch.from_numpy(rng.normal(size=(h, pitch)).astype(np.float32)).cuda()
    b = torch.from_numpy(rng.normal(size=(h, pitch)).astype(np.float32)).cuda()
    scratch = torch.empty(lib.mcp_residual2d_scratch_bytes() // 8, dtype=torch.float64, device="cuda")
    out = torch.zeros(2, dtype=torch.float64, device="cuda")
    lib_mod.check(lib.mcp_residual2d(0, a.data_ptr(), b.data_ptr(), pitch, w, 3, 90, scratch.data_ptr(),
                                     out.data_ptr(), None))
    d = (a[3:90, :w] - b[3:90, :w]).double()
    ref_ss, ref_mx = float((d * d).sum()), float(d.abs().max())
    got = out.cpu().numpy()
    assert abs(got[0] - ref_ss) <= 1e-12 * ref_ss and got[1] == ref_mx


def test_pack_rows_ascii_and_cast():
    from mpi_cuda_process_b200 import io as mio
    lib_mod, lib = _lib()
    h, w, pitch = 11, 37, 64
    u = (np.random.default_rng(2).random((h, w)) < 0.3).astype(np.int32)
    src = torch.zeros((h + 2, pitch), dtype=torch.int32, device="cuda")
    src[1:h + 1, :w] = torch.from_numpy(u).cuda()
    dst = torch.zeros(h * (w + 1), dtype=torch.uint8, device="cuda")
    lib_mod.check(lib.mcp_pack_rows(2, 2, src.data_ptr(), pitch, w, 1, h + 1, dst.data_ptr(), 0, None))
    assert b"\n" + dst.cpu().numpy().tobytes() + b"\n" == mio.format_array(u)
    s64 = torch.rand((h, pitch), dtype=torch.float64, device="cuda")
    d32 = torch.zeros((h, w), dtype=torch.float32, device="cuda")
    lib_mod.check(lib.mcp_pack_rows(1, 1, s64.data_ptr(), pitch, w, 0, h, d32.data_ptr(), 0, None))
    assert torch.equal(d32, s64[:, :w].float())


def test_cli_on_gpu():
    import subprocess, sys
    r = subprocess.run([sys.executable, "-m", "mpi_cuda_process_b200", "--workload", "mdf", "--check", "--json"],
                       input="100 256 256\n", capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr + r.stdout
    assert "bitwise equal after 100 generations" in r.stdout and '"backend": "fused"' in r.stdout


def test_fused_run_to_tolerance_single_launch():
    """The kernel stops itself from the fused residual: one launch, one time block after convergence."""
    h, w, tb, tol = 64, 96, 4, 0.05
    cfg = RunConfig(workload="mdf", height=h, width=w, backend="fused", time_block=tb, tol=tol, max_generations=4000)
    with Solver(cfg) as s:
        s.initialize()
        st = s.run_to_tolerance()
        s.check_status()
        assert st.launches == 1
        # oracle: first block whose last generation has max|d| <= tol; the run ends one block later
        u = oracle.mdf_init(h, w)
        gen, conv = 0, None
        while conv is None:
            v = oracle.mdf_step(u)
            gen += 1
            if gen % tb == 0 and oracle.residual(v, u, "linf") <= tol:
                conv = gen
            u = v
        assert st.generations == conv + tb
        assert s.generation == st.generations
        assert s.gather().tobytes() == oracle.mdf_run(oracle.mdf_init(h, w), st.generations).tobytes()
        hist = dict(st.residual_history)
        assert hist[conv] <= tol and all(r > tol for g_, r in st.residual_history if g_ < conv)
        # the solver can keep going afterwards (counters stayed consistent)
        s.run(8)
        assert s.gather().tobytes() == oracle.mdf_run(oracle.mdf_init(h, w), st.generations + 8).tobytes()


def test_heat3d_fused_random_field_and_residual():
    d, h, w, g = 40, 70, 200, 6
    rng = np.random.default_rng(5)
    f = oracle.heat3d_init(d, h, w)
    f[1:-1, 1:-1, 1:-1] = rng.uniform(0, 100, (d - 2, h - 2, w - 2)).astype(np.float32)
    cfg = RunConfig(workload="heat3d", depth=d, height=h, width=w, backend="fused", residual="linf")
    with Solver(cfg) as s:
        s.initialize(f)
        st = s.run(g, residual=True)
        s.check_status()
        ref_prev = oracle.heat3d_run(f, g - 1)
        ref = oracle.heat3d_step(ref_prev)
        assert s.gather().tobytes() == ref.tobytes()
        assert st.residual_history[-1][1] == oracle.residual(ref, ref_prev, "linf")


def test_gather_ascii_and_fp32_cast_on_device():
    from mpi_cuda_process_b200 import io as mio
    cfg = RunConfig(workload="life", height=33, width=47, backend="fused", time_block=2)
    with Solver(cfg) as s:
        s.initialize()
        s.run(5)
        ref = oracle.life_run(oracle.life_init_reference(33, 47), 5)
        assert s.gather_ascii() == mio.format_array(ref)
    cfg = RunConfig(workload="mdf", height=40, width=50, backend="fused", dtype="fp64")
    with Solver(cfg) as s:
        s.initialize()
        s.run(9)
        ref = oracle.mdf_run(oracle.mdf_init(40, 50, dtype=np.float64), 9)
        got = s.gather(as_fp32=True)
        assert got.dtype == np.float32 and got.tobytes() == ref.astype(np.float32).tobytes()


@pytest.mark.parametrize("init", ["hash", "box"])
@pytest.mark.parametrize("dtype", ["fp32", "fp64"])
def test_device_init_shapes_and_run(init, dtype):
    npdt = np.float32 if dtype == "fp32" else np.float64
    cfg = RunConfig(workload="mdf", height=150, width=203, backend="fused", init=init, seed=3, dtype=dtype, time_block=4)
    with Solver(cfg) as s:
        s.initialize()
        f0 = oracle.mdf_init_shape(150, 203, init, seed=3, dtype=npdt)
        assert s.gather().tobytes() == f0.tobytes()
        s.run(13)
        assert s.gather().tobytes() == oracle.mdf_run(f0, 13).tobytes()


@pytest.mark.parametrize("bc", ["dead-outside", "fixed-ring"])
@pytest.mark.parametrize("tb", [1, 2, 4])
@pytest.mark.parametrize("h,w,g", [(40, 70, 12), (150, 275, 23), (64, 4000, 9), (300, 4100, 14)])
def test_life_bit_packed_exact(bc, tb, h, w, g):
    """One cell per bit, 128 cells per lane: same generations as the int32 oracle, cell for cell."""
    f = (np.random.default_rng(h + w).random((h, w)) < 0.4).astype(np.int32)
    cfg = RunConfig(workload="life", dtype="bits", height=h, width=w, backend="fused", bc=bc, time_block=tb)
    with Solver(cfg) as s:
        s.initialize(f)
        st = s.run(g, residual=True)
        s.check_status()
        ref_prev = oracle.life_run(f, g - 1, bc)
        ref = oracle.life_step(ref_prev, bc)
        assert (s.gather() == ref).all()
        assert s._last_residual[0] == float((ref != ref_prev).sum())         # changed cells in the last generation


def test_life_bit_packed_device_init_matches_hash_oracle():
    cfg = RunConfig(workload="life", dtype="bits", height=77, width=391, backend="fused", seed=5, time_block=2)
    with Solver(cfg) as s:
        s.initialize()
        assert (s.gather() == oracle.life_init_hash(77, 391, 0.15, 5)).all()
        s.run(6)
        assert (s.gather() == oracle.life_run(oracle.life_init_hash(77, 391, 0.15, 5), 6)).all()


def test_checkpoint_roundtrip_bit_packed_life(tmp_path):
    """save() writes unpacked cells, load() must pack them again (w > pitch_words used to raise, small w
    silently stored cells as words)."""
    from mpi_cuda_process_b200.utils import checkpoint as ckpt
    for h, w in ((40, 70), (33, 1100)):
        cfg = RunConfig(workload="life", dtype="bits", height=h, width=w, backend="fused", time_block=2)
        d = tmp_path / f"c{w}"
        with Solver(cfg) as s:
            s.initialize()
            s.run(4)
            ckpt.save(s, str(d))
        with Solver(cfg) as s:
            assert ckpt.load(s, str(d)) == 4
            s.run(6)
            got = s.gather()
        ref = oracle.life_run(oracle.life_init_reference(h, w), 10)
        assert (got == ref).all()


@pytest.mark.parametrize("dtype", ["fp32", "fp64"])
def test_device_checksum_matches_numpy_restatement(dtype):
    cfg = RunConfig(workload="mdf", height=150, width=333, backend="fused", time_block=4, dtype=dtype, init="hash", seed=3)
    with Solver(cfg) as s:
        s.initialize()
        s.run(12)
        got = s.checksum()
        field = s.gather()
    assert got == oracle.checksum_field(field)
    assert field.tobytes() == oracle.run_field(cfg, oracle.init_field(cfg), 12).tobytes()


def test_gather_bits_unpacks_on_the_device():
    cfg = RunConfig(workload="life", dtype="bits", height=70, width=1000, backend="fused", time_block=2)
    with Solver(cfg) as s:
        s.initialize()
        s.run(5)
        got = s.gather()
        assert s.last_gather_stats["bytes_into_rank0"] == 70 * 1000 * 4
    assert (got == oracle.life_run(oracle.life_init_reference(70, 1000), 5)).all()


def test_final_residual_single_launch_matches_history():
    """residual="final": one launch, residual-free row loop for all but the last time block, same value as the
    per-block history's last entry and as the oracle."""
    cfg = RunConfig(workload="mdf", height=200, width=300, backend="fused", time_block=4, residual="linf")
    with Solver(cfg) as s:
        s.initialize()
        st = s.run(40, residual="final")
        assert st.launches == 1 and len(st.residual_history) == 1
        final = st.residual_history[-1][1]
    with Solver(cfg) as s:
        s.initialize()
        hist = s.run(40, residual=True).residual_history
    u = oracle.mdf_run(oracle.mdf_init(200, 300), 39)
    assert final == hist[-1][1] == oracle.residual(oracle.mdf_step(u), u, "linf")


@pytest.mark.parametrize("dtype", ["fp32", "fp64"])
def test_heat3d_fused_residual_in_kernel(dtype):
    """3-D: the residual of every generation comes out of the sweep kernel itself (one launch, no second pass)."""
    d, h, w, g = 37, 45, 150, 9
    cfg = RunConfig(workload="heat3d", depth=d, height=h, width=w, backend="fused", dtype=dtype, residual="linf")
    npdt = np.float32 if dtype == "fp32" else np.float64
    with Solver(cfg) as s:
        s.initialize()
        st = s.run(g, residual=True)
        assert st.launches == 1 and len(st.residual_history) == g
        got = s.gather()
        l2_last = float(np.sqrt(s._last_residual[0]))
    u = oracle.heat3d_init(d, h, w, dtype=npdt)
    exp = []
    for _ in range(g):
        v = oracle.heat3d_step(u)
        exp.append(oracle.residual(v, u, "linf"))
        prev, u = u, v
    assert [r for _, r in st.residual_history] == exp
    assert got.tobytes() == u.tobytes()
    assert l2_last == pytest.approx(oracle.residual(u, prev, "l2"), rel=1e-6)
    with Solver(cfg) as s:                      # "final": only the last generation carries it
        s.initialize()
        st = s.run(g, residual="final")
        assert st.launches == 1 and [r for _, r in st.residual_history] == exp[-1:]


def test_heat3d_run_to_tolerance_single_launch():
    d, h, w, tol = 20, 24, 40, 0.5
    cfg = RunConfig(workload="heat3d", depth=d, height=h, width=w, backend="fused", tol=tol, max_generations=3000)
    with Solver(cfg) as s:
        s.initialize()
        st = s.run_to_tolerance()
        s.check_status()
        got = s.gather()
    u = oracle.heat3d_init(d, h, w)
    conv = None
    for gen in range(1, 3001):
        v = oracle.heat3d_step(u)
        if oracle.residual(v, u, "linf") <= tol:
            conv = gen
        u = v
        if conv is not None:
            break
    assert st.launches == 1 and st.stop_reason == "converged"
    assert st.generations == conv + 1               # the run ends one generation after the one that converged
    assert got.tobytes() == oracle.heat3d_run(oracle.heat3d_init(d, h, w), st.generations).tobytes()


@pytest.mark.parametrize("dtype,shape", [("fp32", (20, 24, 40)), ("fp32", (37, 77, 260)), ("fp64", (19, 100, 150)),
                                         ("fp32", (40, 33, 129))])
def test_heat3d_cluster_dsmem_kernel_bitwise(dtype, shape):
    """Thread-block clusters: y-adjacent CTAs share the tile-boundary row through DSMEM (mcp_step3d_cluster)."""
    d, h, w = shape
    cfg = RunConfig(workload="heat3d", depth=d, height=h, width=w, backend="nccl", dtype=dtype, extra={"cluster3d": True})
    with Solver(cfg) as s:
        s.initialize()
        s.run(5)
        got = s.gather()
    npdt = np.float32 if dtype == "fp32" else np.float64
    assert got.tobytes() == oracle.heat3d_run(oracle.heat3d_init(d, h, w, dtype=npdt), 5).tobytes()
