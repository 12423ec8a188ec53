"""The scalar fp32 operator behind the grid barrier -- the instantiation the plan picks for fp32 MDF at time block 8
when a warp has at least five tiles per sweep and the plan has no neighbours (stencil2d_fused.cu:
prefers_scalar_barrier), i.e. what the 16384 x 16384 single-GPU headline runs.  Grids of unit-test size never reach five
tiles per warp on 148 CTAs x 8 warps, so here the grid is held to ONE CTA (``max_ctas``) to get there on a small field,
and the result is compared bit for bit with the oracle (the reference's arithmetic, MDF_kernel.cu:20).

Written after the round's GPU time was spent (the operator itself was validated at benchmark scale: same final residual
as the packed operator at 16384 x 16384, profiles/r2/bench_n1_final.json), hence xfail(strict=False): XPASS where it
works, never a failure of the suite."""
import ctypes as C

import numpy as np
import pytest

from mpi_cuda_process_b200 import RunConfig, Solver, oracle

pytestmark = [pytest.mark.gpu,
              pytest.mark.xfail(strict=False, reason="written after the round's GPU budget was spent: not yet run")]


def _tiles_per_warp(s) -> float:
    ns, nc, grid, smem = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    s._lib.mcp_fused2d_geometry(s._plan, s.time_block, C.byref(ns), C.byref(nc), C.byref(grid), C.byref(smem))
    return ns.value * nc.value / (grid.value * 8.0)


@pytest.mark.parametrize("h,w,g", [(1200, 900, 24), (1500, 1031, 40)])
def test_scalar_barrier_operator_bitwise(h, w, g):
    rng = np.random.default_rng(7)
    f = oracle.mdf_init(h, w)
    f[1:-1, 1:-1] = rng.uniform(-50, 150, (h - 2, w - 2)).astype(np.float32)
    cfg = RunConfig(workload="mdf", height=h, width=w, backend="fused", time_block=8, residual="linf",
                    extra={"max_ctas": 1})
    with Solver(cfg) as s:
        s.initialize(f)
        st = s.run(g, residual="final")
        s.check_status()
        assert _tiles_per_warp(s) >= 5.0            # the rule's condition: this run used the scalar operator
        assert s.gather().tobytes() == oracle.mdf_run(f, g).tobytes()
        prev = oracle.mdf_run(f, g - 1)
        assert st.residual_history[-1][1] == oracle.residual(oracle.mdf_step(prev), prev, "linf")
