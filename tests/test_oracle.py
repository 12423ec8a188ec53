"""Oracle self-consistency: the pinned arithmetic of MDF_kernel.cu:20, life rule, glibc rand stream."""
import ctypes as C

import numpy as np
import pytest

from mpi_cuda_process_b200 import oracle
from mpi_cuda_process_b200.ops import _lib


def test_mdf_init_matches_create_universe():
    u = oracle.mdf_init(5, 7)
    assert u.dtype == np.float32
    assert (u[0] == 100).all() and (u[-1] == 100).all() and (u[:, 0] == 100).all() and (u[:, -1] == 100).all()
    assert (u[1:-1, 1:-1] == 0).all()


def test_mdf_one_step_by_hand():
    u = oracle.mdf_init(4, 4)
    v = oracle.mdf_step(u)
    # interior cell (1,1): E=0 W=100 N=100 S=0 C=0 -> 0.25*200 = 50
    assert v[1, 1] == np.float32(50.0)
    assert (v[0] == 100).all() and (v[:, 0] == 100).all()


def test_mdf_fp32_fma_equals_reference_fp64_fma():
    """dropping the reference's fp64 DFMA does not change a single bit."""
    rng = np.random.default_rng(0)
    C_, N, S, W, E = (rng.uniform(0, 100, 1 << 20).astype(np.float32) for _ in range(5))
    a = oracle.mdf_update_cells(C_, N, S, W, E)
    b = oracle.mdf_update_cells_ref_fp64(C_, N, S, W, E)
    assert a.tobytes() == b.tobytes()


def test_simplified_algebra_is_not_bit_equal():
    rng = np.random.default_rng(1)
    C_, N, S, W, E = (rng.uniform(0, 100, 1 << 16).astype(np.float32) for _ in range(5))
    a = oracle.mdf_update_cells(C_, N, S, W, E)
    b = oracle.mdf_update_cells_simplified(C_, N, S, W, E)
    frac = np.mean(a != b)
    assert frac > 0.01


def test_mdf_converges_to_ring_value():
    u = oracle.mdf_run(oracle.mdf_init(8, 8), 400)
    assert np.allclose(u, 100.0, atol=1e-3)


def test_host_cpp_mdf_matches_numpy():
    lib = _lib.host_lib()
    u = oracle.mdf_init(33, 47)
    u[1:-1, 1:-1] = np.random.default_rng(2).uniform(0, 100, (31, 45)).astype(np.float32)
    for strict in (0, 1):
        out = np.empty_like(u)
        lib.mcp_host_mdf_step_f32(u.ctypes.data, out.ctypes.data, 33, 47, strict)
        assert out.tobytes() == oracle.mdf_step(u).tobytes()
    u64 = u.astype(np.float64)
    out64 = np.empty_like(u64)
    lib.mcp_host_mdf_step_f64(u64.ctypes.data, out64.ctypes.data, 33, 47)
    assert out64.tobytes() == oracle.mdf_step(u64).tobytes()


def test_glibc_rand_stream():
    r = oracle.GlibcRand()
    assert r.next() == 1804289383
    assert r.next() == 846930886
    assert r.next() == 1681692777


def test_life_init_matches_host_rand():
    lib = _lib.host_lib()
    h, w = 19, 23
    u = np.empty((h, w), dtype=np.int32)
    lib.mcp_host_life_init_reference(u.ctypes.data, h, w, C.c_float(0.15))
    assert (u == oracle.life_init_reference(h, w, 0.15)).all()
    assert u[1:-1, 1:-1].sum() > 0 and (u[0] == 0).all()


@pytest.mark.parametrize("bc", ["dead-outside", "fixed-ring"])
def test_life_step_matches_host(bc):
    lib = _lib.host_lib()
    rng = np.random.default_rng(3)
    u = (rng.random((17, 21)) < 0.4).astype(np.int32)
    out = np.empty_like(u)
    lib.mcp_host_life_step(u.ctypes.data, out.ctypes.data, 17, 21, 1 if bc == "fixed-ring" else 0)
    assert (out == oracle.life_step(u, bc)).all()


def test_life_blinker_and_block():
    u = np.zeros((5, 5), dtype=np.int32)
    u[2, 1:4] = 1
    v = oracle.life_step(u)
    assert (v[1:4, 2] == 1).all() and v.sum() == 3
    assert (oracle.life_step(v) == u).all()
    b = np.zeros((4, 4), dtype=np.int32)
    b[1:3, 1:3] = 1
    assert (oracle.life_step(b) == b).all()


def test_heat3d_step_by_hand():
    u = oracle.heat3d_init(3, 3, 3)
    v = oracle.heat3d_step(u)
    # centre: six neighbours at 100, C = 0 -> 0 + (1/6)*600
    assert v[1, 1, 1] == np.float32(0.0) + np.float32(1.0 / 6.0) * np.float32(600.0)


def test_residual():
    a = oracle.mdf_init(6, 6)
    b = oracle.mdf_step(a)
    assert oracle.residual(b, a, "linf") == 50.0
    assert oracle.residual(b, a, "l2") > 50.0


def test_hash_init_density():
    u = oracle.life_init_hash(200, 200, 0.15, seed=7)
    d = u[1:-1, 1:-1].mean()
    assert 0.12 < d < 0.18 and (u[0] == 0).all()


def test_mdf_subnormal_front_matches_true_fma():
    """The diffusion front passes through subnormal values, where 0.25*s is not exact: the oracle must
    behave like a real fused multiply-add (what the GPU and the reference's DFMA do)."""
    lib = _lib.host_lib()
    u = oracle.mdf_init(200, 200)
    v = u.copy()
    out = np.empty_like(u)
    for _ in range(80):
        lib.mcp_host_mdf_step_f32(v.ctypes.data, out.ctypes.data, 200, 200, 0)
        v, out = out, v
        u = oracle.mdf_step(u)
    assert np.any((np.abs(u) > 0) & (np.abs(u) < np.finfo(np.float32).tiny))   # subnormals are present
    assert u.tobytes() == v.tobytes()
