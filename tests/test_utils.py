import json

import numpy as np

from mpi_cuda_process_b200 import models
from mpi_cuda_process_b200.utils import metrics
from mpi_cuda_process_b200.utils.clocks import ClockSampler, NvmlClockSampler


def test_roofline_arithmetic():
    peaks = {"hbm_gbs": 6400.0, "source": "measured"}
    # 64 Mi cells, 8 B/update, 10 generations in 1 ms -> 5.4 TB/s effective
    r = metrics.roofline(1 << 26, 10, 1.0, 8, 0.0, peaks)
    assert r["bound"] == "hbm"
    assert abs(r["one_pass_equivalent_gbs"] - (1 << 26) * 8 * 10 / 1e-3 / 1e9) < 1e-6
    assert abs(r["fraction_of_roofline"] - r["one_pass_equivalent_gbs"] / 6400.0) < 1e-9
    assert "of measured" in r["peaks_source"]
    # a tiny grid with a huge halo is NVLink-bound
    r2 = metrics.roofline(1024, 1, 1.0, 8, 1e9, peaks)
    assert r2["bound"] == "nvlink"
    assert metrics.glups(10**9, 2, 1000.0) == 2.0


def test_measured_peaks_fallback(tmp_path):
    assert metrics.measured_peaks(str(tmp_path))["source"] == "fallback"
    (tmp_path / "MEASURED_PEAKS.json").write_text(json.dumps({"hbm_gbs": 6454.6}))
    p = metrics.measured_peaks(str(tmp_path))
    assert p["source"] == "measured" and p["hbm_gbs"] == 6454.6


def test_clock_samplers_degrade_without_a_gpu():
    for s in (ClockSampler(0), NvmlClockSampler(0)):
        out = s.start().stop()
        assert "reasons" in out and "samples" in out


def test_workload_registry():
    assert set(models.REGISTRY) == {"mdf", "life", "heat3d"}
    m = models.get("mdf")
    assert m.bytes_per_update("fp32") == 8 and m.bytes_per_update("fp64") == 16
    u = m.step(m.init(6, 6))
    assert u[1, 1] == np.float32(50.0)
    import pytest
    with pytest.raises(ValueError):
        models.get("nope")


def test_solver_dispatches_through_the_workload_registry():
    from mpi_cuda_process_b200 import RunConfig, Solver
    assert models.get("mdf").time_block_limit("fp32") == 8 and models.get("life").time_block_limit("bits") == 2
    assert models.get("life").op_code == 1 and models.get("heat3d").op_code == -1
    cfg = RunConfig(workload="life", height=9, width=7, backend="cpu", bc="dead-outside")
    with Solver(cfg) as s:
        assert s.model is models.get("life") and s._fix_bounds == (0, 8, 0, 6)
    cfg = RunConfig(workload="mdf", height=9, width=7, backend="cpu")
    with Solver(cfg) as s:
        assert s._fix_bounds == (1, 7, 1, 5)


def test_result_ring_settles_the_handle_that_still_owns_a_slot():
    """Solver.run_async hands out 8 pinned result slots in a ring; a ninth enqueue must read the oldest handle's value
    out before the slot is overwritten (solver._ResultRing; no counterpart in the reference, which has no results)."""
    from mpi_cuda_process_b200.solver import _ResultRing

    class Handle:
        def __init__(self, log, name):
            self.log, self.name, self.waited = log, name, False

        def wait(self):
            if not self.waited:
                self.waited = True
                self.log.append(self.name)

    log, ring, handles = [], _ResultRing(4), []
    for i in range(10):
        idx = ring.claim()
        assert idx == i % 4
        handles.append(ring.bind(idx, Handle(log, i)))
        if i == 1:
            handles[1].wait()                       # a handle the caller has already waited for is not touched again
    assert log == [1, 0, 2, 3, 4, 5]                # 0 settled when slot 0 came round (i=4), 2 at i=6, ...
    assert [h.waited for h in handles] == [True] * 6 + [False] * 4


def test_numa_binding_from_fake_sysfs(tmp_path):
    """utils/numa.py: CPU affinity from the GPU's PCI locality (best effort, never raises).  No counterpart in the
    reference, where every rank shares GPU 0 (MDF_kernel.cu:114-121)."""
    import os
    from mpi_cuda_process_b200.utils import numa
    assert numa.parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert numa.parse_cpulist("") == set()
    allowed = sorted(os.sched_getaffinity(0))
    dev = tmp_path / "0000:1b:00.0"
    dev.mkdir()
    (dev / "numa_node").write_text("1\n")
    try:
        # a cpulist that holds exactly one allowed CPU: the process ends up bound to it
        (dev / "local_cpulist").write_text(f"{allowed[0]}\n")
        rec = numa.bind_to_gpu(0, sysfs=str(tmp_path), pci="0000:1b:00.0")
        assert rec["bound"] and rec["numa_node"] == 1 and rec["cpus"] == 1
        assert os.sched_getaffinity(0) == {allowed[0]}
    finally:
        os.sched_setaffinity(0, allowed)
    # CPUs outside the cpuset, a missing device and a missing file are reported, not raised
    (dev / "local_cpulist").write_text("100000-100003\n")
    assert numa.bind_to_gpu(0, sysfs=str(tmp_path), pci="0000:1b:00.0")["bound"] is False
    assert numa.bind_to_gpu(0, sysfs=str(tmp_path), pci="0000:ff:00.0")["bound"] is False
    assert os.sched_getaffinity(0) == set(allowed)
    # every CPU local (single node): nothing to change
    (dev / "local_cpulist").write_text(",".join(str(c) for c in allowed) + "\n")
    rec = numa.bind_to_gpu(0, sysfs=str(tmp_path), pci="0000:1b:00.0")
    assert rec["bound"] and "already local" in rec["why"]
