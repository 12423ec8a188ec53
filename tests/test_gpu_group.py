"""The cross-rank protocol of the fused 2-D kernel on ONE GPU: several ranks advance inside one cooperative launch
(mcp_fused2d_run_group) and wait on each other's flags exactly as kernels on different GPUs do.  Covers what a 1-GPU
box otherwise never runs: per-strip halo flags, the residual exchange, the drain at the end of a launch, launches with
different time blocks in a row, injected delays (the write-after-read argument of DESIGN.md) and the timeout path.
Reference counterpart of the hazard: the unsynchronised D2H -> MPI_Send of MDF_kernel.cu:177-183."""
import ctypes as C

import numpy as np
import pytest
import torch

from mpi_cuda_process_b200 import oracle

pytestmark = pytest.mark.gpu

MCP_OPT_TIMEOUT_MS, MCP_OPT_DEBUG_FLAGS, MCP_OPT_SYNC_MODE = 1, 2, 3


def _field(h, w, seed=3):
    f = oracle.mdf_init(h, w)
    f[1:-1, 1:-1] = np.random.default_rng(seed).uniform(0, 100, (h - 2, w - 2)).astype(np.float32)
    return f


@pytest.mark.parametrize("n,tb,h,w,strip_flags", [(2, 4, 90, 200, True), (3, 8, 200, 700, True), (4, 8, 333, 389, True),
                                                  (2, 4, 90, 200, False), (8, 8, 1100, 500, True)])
def test_emulated_ranks_bitwise(n, tb, h, w, strip_flags):
    from _emu import EmulatedRanks
    f0 = _field(h, w)
    emu = EmulatedRanks(f0, n, tb, strip_flags=strip_flags)
    try:
        g = 0
        for sweeps in (1, 3, 2):                    # several launches: sequence numbers continue across them
            emu.run(sweeps * tb)
            g += sweeps * tb
        assert emu.status() == [0] * n
        assert emu.field().tobytes() == oracle.mdf_run(f0, g).tobytes()
        ns = -(-w // ((32 - 2 * (-(-tb // 4))) * 4))
        for r in range(n):                          # every halo counter saw exactly the accounted arrivals
            fl = emu.flags[r].cpu().tolist()
            if r > 0:
                assert fl[0] == (g // tb) * ns
            if r < n - 1:
                assert fl[1] == (g // tb) * ns
    finally:
        emu.close()


@pytest.mark.parametrize("sync_mode", [1, 2])
def test_emulated_ranks_with_injected_delays(sync_mode):
    """debug bit 8: pseudo-random 5-200 us stalls in front of halo pushes and publishes, per (rank, sweep, strip)."""
    from _emu import EmulatedRanks
    h, w, n, tb = 400, 900, 4, 8
    f0 = _field(h, w, seed=11)
    emu = EmulatedRanks(f0, n, tb, chunk_rows=24, options={MCP_OPT_DEBUG_FLAGS: 8, MCP_OPT_SYNC_MODE: sync_mode})
    try:
        emu.run(6 * tb)
        emu.run(5 * tb)
        assert emu.status() == [0] * n
        assert emu.field().tobytes() == oracle.mdf_run(f0, 11 * tb).tobytes()
    finally:
        emu.close()


def test_emulated_final_residual_is_global_and_identical():
    """MCP_RES_FINAL on the barrier-free path: the last tile to finish reduces, the ranks exchange over their inboxes,
    every rank reads the same bits through the one-copy result block."""
    from _emu import EmulatedRanks
    h, w, n, tb = 300, 500, 3, 4
    f0 = _field(h, w, seed=5)
    emu = EmulatedRanks(f0, n, tb)
    try:
        emu.run(5 * tb, residual_mode=2)
        res = [emu.result(r) for r in range(n)]
        assert all(x["status"] == 0 and x["sweeps_done"] == 5 for x in res)
        assert len({(x["sumsq"], x["max"]) for x in res}) == 1
        prev, last = oracle.mdf_run(f0, 5 * tb - 1), oracle.mdf_run(f0, 5 * tb)
        d = (last - prev).astype(np.float32)
        assert res[0]["max"] == float(np.abs(d).max())
        assert res[0]["sumsq"] == pytest.approx(float((d.astype(np.float64) ** 2).sum()), rel=1e-6)
        assert emu.field().tobytes() == last.tobytes()
    finally:
        emu.close()


def test_timeout_when_a_neighbour_never_runs():
    """A rank whose down neighbour never starts: the second sweep's boundary tiles wait for halo flags that never come.
    The bounded wait gives up after the configured 300 ms, every CTA drains, the status word says MCP_ERR_TIMEOUT and
    the plan can be destroyed."""
    from _emu import EmulatedRanks
    import time
    from mpi_cuda_process_b200.ops import _lib as lib_mod
    h, w, tb = 120, 300, 4
    emu = EmulatedRanks(_field(h, w), 2, tb, options={MCP_OPT_TIMEOUT_MS: 300})
    try:
        lib = emu.lib
        t0 = time.perf_counter()
        rc = lib.mcp_fused2d_run(emu.plans[0], 3 * tb, tb, 0, None)      # rank 0 alone; rank 1 is never launched
        assert rc == 0
        torch.cuda.synchronize()
        assert time.perf_counter() - t0 < 10.0
        assert lib.mcp_fused2d_status(emu.plans[0]) == -4                  # MCP_ERR_TIMEOUT
        assert b"timed out" in lib.mcp_last_error()
        lib_mod.check(lib.mcp_fused2d_clear_status(emu.plans[0]))
        assert lib.mcp_fused2d_status(emu.plans[0]) == 0
    finally:
        emu.close()
    torch.cuda.synchronize()                                               # the context is healthy afterwards
    assert float(torch.ones(4, device="cuda").sum()) == 4.0


def test_run_to_tolerance_refuses_a_local_stop_decision():
    """neighbours but no residual exchange -> ranks would stop in different sweeps; the C ABI says no."""
    from _emu import EmulatedRanks
    emu = EmulatedRanks(_field(90, 200), 2, 4, residual=False)
    try:
        done = C.c_int(0)
        rc = emu.lib.mcp_fused2d_run_to_tolerance(emu.plans[0], 400, 4, 1, C.c_double(0.1), C.byref(done), None)
        assert rc == -2 and b"residual exchange" in emu.lib.mcp_last_error()
    finally:
        emu.close()


def test_coarse_halo_counters_keep_the_grid_barrier():
    """Found by scripts/stress_fused.py: with ONE arrival counter per neighbour (no per-strip flags) a barrier-free
    neighbour's early strips of sweep k+1 were counted for a late strip of sweep k.  Such plans now keep the grid
    barrier; this is the case that failed (3 ranks, injected delays, two launches, final residual on the first)."""
    from _emu import EmulatedRanks
    h, w, n, tb = 136, 1422, 3, 8
    f0 = _field(h, w, seed=4)
    emu = EmulatedRanks(f0, n, tb, strip_flags=False, chunk_rows=24,
                        options={MCP_OPT_DEBUG_FLAGS: 8, MCP_OPT_SYNC_MODE: 2})
    try:
        emu.run(5 * tb, residual_mode=2)
        emu.run(1 * tb)
        assert emu.status() == [0] * n
        assert emu.field().tobytes() == oracle.mdf_run(f0, 6 * tb).tobytes()
    finally:
        emu.close()
