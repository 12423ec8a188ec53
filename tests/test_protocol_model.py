"""Protocol checker: the flag rules of the fused 2-D kernel restated on an abstract machine and run under an
adversarial scheduler (mpi_cuda_process_b200/utils/protocol_model.py).  CPU only.

The reference has nothing to compare with: blocking MPI calls and cudaDeviceSynchronize are its whole protocol
(MDF_kernel.cu:166-183).  What is pinned here is (a) the two combinations the kernel uses never read a stale or an
overwritten region whatever the interleaving, and (b) the combination the kernel refuses -- one arrival counter per
neighbour without a grid barrier -- is refused for a reason the model can show.
"""
import itertools

import pytest

from mpi_cuda_process_b200.utils.protocol_model import Model, ProtocolViolation, explore


@pytest.mark.parametrize("flags,sync", [("strip", "dataflow"), ("strip", "barrier"), ("coarse", "barrier")])
def test_sound_combinations_never_violate(flags, sync):
    r = explore(seeds=120, ranks=3, chunks=3, strips=4, sweeps=5, warps=3, flags=flags, sync=sync)
    assert r["violations"] == 0, r.get("first")
    assert r["ok"] == 120


def test_small_and_degenerate_geometries():
    # one chunk (a tile that is both the upper and the lower edge), one strip, one warp, one rank
    for R, C, S, W in itertools.product((1, 2, 4), (1, 2, 4), (1, 3), (1, 4)):
        for flags, sync in (("strip", "dataflow"), ("coarse", "barrier")):
            r = explore(seeds=8, ranks=R, chunks=C, strips=S, sweeps=4, warps=W, flags=flags, sync=sync)
            assert r["violations"] == 0, (R, C, S, W, flags, sync, r.get("first"))


def test_coarse_counter_without_barrier_is_unsound():
    """The case the injected-delay stress run caught on the GPU: the kernel now forces such plans to barrier mode
    (stencil2d_fused.cu choose_mode); the model shows why."""
    r = explore(seeds=200, ranks=3, chunks=3, strips=4, sweeps=5, warps=3, flags="coarse", sync="dataflow")
    assert r["violations"] > 0
    assert "ghost" in r["first"] and "stale" in r["first"]


def test_single_rank_coarse_dataflow_is_fine():
    # no neighbours -> no counters involved: the barrier-free schedule alone is sound
    r = explore(seeds=60, ranks=1, chunks=4, strips=4, sweeps=5, warps=4, flags="coarse", sync="dataflow")
    assert r["violations"] == 0, r.get("first")


def test_violation_is_reproducible_from_its_seed():
    r = explore(seeds=200, ranks=3, chunks=3, strips=4, sweeps=5, warps=3, flags="coarse", sync="dataflow")
    seed = int(r["first"].rsplit("seed=", 1)[1])
    with pytest.raises(ProtocolViolation) as e1:
        Model(ranks=3, chunks=3, strips=4, sweeps=5, warps=3, flags="coarse", sync="dataflow", seed=seed).run()
    assert str(e1.value) == r["first"]


def test_without_long_stalls_the_bug_hides():
    """Documents why the stress test injects long delays: with only fine-grained interleaving a neighbour's sweeps
    never separate and the unsound combination passes -- short random schedules are not evidence."""
    r = explore(seeds=60, ranks=3, chunks=3, strips=4, sweeps=5, warps=3, flags="coarse", sync="dataflow",
                stall_rate=0.0)
    assert r["ok"] + r["violations"] == 60
    assert r["violations"] <= 2


MULTI = dict(ranks=3, chunks=3, strips=4, sweeps=3, warps=3, launches=3)


@pytest.mark.parametrize("flags,sync", [("strip", "dataflow"), ("coarse", "barrier")])
def test_back_to_back_launches_with_host_uploads_need_no_host_barrier(flags, sync):
    """Solver.load_local / commit_input overwrite the slab between launches with no barrier between the ranks
    (solver.py); drain_incoming at the end of every launch is what makes that safe."""
    r = explore(seeds=100, flags=flags, sync=sync, **MULTI)
    assert r["violations"] == 0, r.get("first")


@pytest.mark.parametrize("flags,sync", [("strip", "dataflow"), ("coarse", "barrier")])
def test_without_the_drain_an_upload_races_with_late_halo_pushes(flags, sync):
    r = explore(seeds=60, flags=flags, sync=sync, drain=False, **MULTI)
    assert r["violations"] > 0
    assert "ghost" in r["first"]


@pytest.mark.parametrize("flags,sync", [("strip", "dataflow"), ("coarse", "barrier")])
def test_the_drain_is_only_needed_for_uploads(flags, sync):
    # launches that continue from the previous state are ordered by the sequence flags alone
    r = explore(seeds=100, flags=flags, sync=sync, drain=False, upload=False, **MULTI)
    assert r["violations"] == 0, r.get("first")


def test_random_geometries_property():
    """Hypothesis over ranks x chunks x strips x warps x sweeps x launches: the two combinations the kernel runs stay
    sound; shrinks to a minimal geometry + seed if they ever do not."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=60, deadline=None)
    @given(ranks=st.integers(1, 4), chunks=st.integers(1, 5), strips=st.integers(1, 5), warps=st.integers(1, 5),
           sweeps=st.integers(1, 4), launches=st.integers(1, 3), seed=st.integers(0, 10**6),
           combo=st.sampled_from([("strip", "dataflow"), ("coarse", "barrier"), ("strip", "barrier")]))
    def run(ranks, chunks, strips, warps, sweeps, launches, seed, combo):
        Model(ranks=ranks, chunks=chunks, strips=strips, warps=warps, sweeps=sweeps, launches=launches,
              flags=combo[0], sync=combo[1], seed=seed, stall_rate=0.05, max_stall=200).run()

    run()


def test_prefetch_must_skip_tiles_that_read_ghost_rows():
    """The next-tile prefetch probes the nine LOCAL completion words only, so the kernel never prefetches a tile that
    reads ghost rows (stencil2d_fused.cu: NextTile.valid); the model shows what happens otherwise."""
    kw = dict(ranks=3, chunks=3, strips=4, sweeps=5, warps=3)
    assert explore(seeds=80, **kw)["violations"] == 0                       # look-ahead + prefetch as built
    assert explore(seeds=80, lookahead=False, prefetch=False, **kw)["violations"] == 0
    bad = explore(seeds=80, prefetch_ghost=True, **kw)
    assert bad["violations"] > 0 and "ghost" in bad["first"]


def test_lookahead_tickets_do_not_deadlock():
    """A warp holds its next ticket while it runs the current one (grab() behind the tile).  The lowest unfinished
    ticket is still always somebody's CURRENT tile, so it can run: no schedule may end in a deadlock (Model.run raises
    RuntimeError on one), also with a single warp per rank and more tiles than warps."""
    for warps in (1, 2, 7):
        for seed in range(25):
            Model(ranks=3, chunks=4, strips=3, sweeps=4, warps=warps, launches=2, seed=seed).run()


def test_proposed_early_upward_signal_is_sound():
    """Design check for an optimisation that is NOT in the kernel yet (NOTES_NEXT_ROUND.md): a top boundary tile pushes
    its first rows upwards and raises the neighbour's flag as soon as they exist -- by then it has consumed its upper
    ghost rows, which are the first rows it streams -- and only then computes the rest of the tile.  The neighbour's
    wait ends most of a tile earlier; the model finds no hazard, also with one chunk per rank and across launches."""
    for kw in (dict(ranks=3, chunks=3, strips=4, sweeps=5, warps=3), dict(ranks=4, chunks=1, strips=3, sweeps=4, warps=2),
               dict(ranks=3, chunks=2, strips=4, sweeps=4, warps=3, launches=3)):
        r = explore(seeds=80, early_up_signal=True, **kw)
        assert r["violations"] == 0, r.get("first")
