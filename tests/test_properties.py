"""Property-based checks (hypothesis) of the host-side logic: decomposition, oracle invariants, I/O round trips."""
import numpy as np
from hypothesis import given, settings, strategies as st

from mpi_cuda_process_b200 import io as mio, oracle
from mpi_cuda_process_b200.parallel.decomp import Decomposition, split_axis


@settings(max_examples=60, deadline=None)
@given(n=st.integers(1, 5000), parts=st.integers(1, 64))
def test_split_axis_partition(n, parts):
    if parts > n:
        parts = n
    r = split_axis(n, parts)
    assert r[0][0] == 0 and r[-1][1] == n and len(r) == parts
    assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
    sizes = [e - s for s, e in r]
    assert max(sizes) - min(sizes) <= 1 and min(sizes) >= 1


@settings(max_examples=40, deadline=None)
@given(h=st.integers(8, 200), w=st.integers(3, 50), world=st.integers(1, 8))
def test_slab_blocks_tile_the_grid_and_neighbours_are_mutual(h, w, world):
    d = Decomposition.slab((h, w), world)
    cover = np.zeros((h, w), dtype=np.int32)
    for r in range(world):
        b = d.block(r)
        cover[d.global_slices(r)] += 1
        lo, hi = b.neighbors[0]
        if lo is not None:
            assert d.block(lo).neighbors[0][1] == r and d.block(lo).stop[0] == b.start[0]
        if hi is not None:
            assert d.block(hi).neighbors[0][0] == r
    assert (cover == 1).all()


@settings(max_examples=25, deadline=None)
@given(pz=st.integers(1, 4), py=st.integers(1, 4), d=st.integers(8, 40), h=st.integers(8, 40))
def test_pencil_blocks_tile_the_grid(pz, py, d, h):
    dec = Decomposition.pencil((d, h, 5), pz * py, (pz, py))
    cover = np.zeros((d, h, 5), dtype=np.int32)
    for r in range(pz * py):
        cover[dec.global_slices(r)] += 1
        assert dec.rank_of(dec.coords_of(r)) == r
    assert (cover == 1).all()


@settings(max_examples=25, deadline=None)
@given(h=st.integers(3, 24), w=st.integers(3, 24), seed=st.integers(0, 10**6))
def test_mdf_maximum_principle_and_fixed_ring(h, w, seed):
    rng = np.random.default_rng(seed)
    u = oracle.mdf_init(h, w)
    u[1:-1, 1:-1] = rng.uniform(0, 100, (h - 2, w - 2)).astype(np.float32)
    v = oracle.mdf_step(u)
    assert (v[0] == u[0]).all() and (v[-1] == u[-1]).all() and (v[:, 0] == u[:, 0]).all() and (v[:, -1] == u[:, -1]).all()
    # a convex combination of the four neighbours (up to rounding): stays inside the data range
    assert v.min() >= u.min() - 1e-3 and v.max() <= u.max() + 1e-3


@settings(max_examples=25, deadline=None)
@given(h=st.integers(1, 20), w=st.integers(1, 20), seed=st.integers(0, 10**6))
def test_life_matches_bruteforce_and_print_array_roundtrip(h, w, seed):
    rng = np.random.default_rng(seed)
    u = (rng.random((h, w)) < 0.45).astype(np.int32)
    v = oracle.life_step(u)
    for i in range(h):
        for j in range(w):
            n = sum(u[y, x] for y in range(max(0, i - 1), min(h, i + 2)) for x in range(max(0, j - 1), min(w, j + 2))) - u[i, j]
            assert v[i, j] == (1 if (n == 3 or (n == 2 and u[i, j])) else 0)
    assert (mio.parse_array(mio.format_array(u), w) == u).all()


@settings(max_examples=30, deadline=None)
@given(rows=st.integers(1, 6), w=st.integers(1, 300), seed=st.integers(0, 10**6))
def test_bit_packing_roundtrip(rows, w, seed):
    from mpi_cuda_process_b200.solver import Solver

    class Shim:
        pass
    f = Shim()
    f.cfg = Shim()
    f.cfg.width = w
    f.pitch = ((w + 31) // 32 + 31) // 32 * 32
    cells = (np.random.default_rng(seed).random((rows, w)) < 0.5).astype(np.int32)
    words = Solver._pack_bits(f, cells)
    assert words.shape == (rows, f.pitch)
    assert (Solver._unpack_bits(f, words) == cells).all()
    x = int(np.random.default_rng(seed + 1).integers(0, w))
    assert ((int(words[0, x // 32]) >> (x % 32)) & 1) == cells[0, x]


@settings(max_examples=80, deadline=None)
@given(rows=st.integers(1, 40000), w=st.integers(1, 40000), tb=st.sampled_from([1, 2, 4, 8]),
       dtype=st.sampled_from([0, 1, 2, 3]), grid=st.sampled_from([1, 37, 148]), chunk_rows=st.sampled_from([0, 0, 7, 48, 500]))
def test_tile_schedule_invariants_without_a_gpu(rows, w, tb, dtype, grid, chunk_rows):
    """The fused kernel's host-side planning (pure arithmetic in the native library) for arbitrary slabs."""
    import ctypes as C
    from mpi_cuda_process_b200.ops import _lib
    lib = _lib.cuda_lib()                       # loads without a GPU; only host code is called
    op = 1 if dtype in (2, 3) else 0
    if op == 1 and tb == 8:
        tb = 4
    ghost = tb
    if rows < ghost:
        rows = ghost
    out = (C.c_int * 8192)()
    ns = C.c_int(0)
    n = lib.mcp_fused2d_plan_schedule(op, dtype, w, rows, ghost, tb, grid, chunk_rows, out, 8192, C.byref(ns))
    assert 1 <= n <= 4096
    hts = [out[i] for i in range(n)]
    assert sum(hts) == rows                                  # the chunks tile the slab exactly
    assert min(hts) >= min(ghost, rows)                      # every chunk holds the halo rows a neighbour needs
    cells_per_lane = {0: 4, 1: 2, 2: 4, 3: 128}[dtype]
    margin = -(-tb // cells_per_lane)
    outw = (32 - 2 * margin) * cells_per_lane
    assert ns.value == -(-w // outw) and ns.value <= 2048 or w > 2048 * outw


@settings(max_examples=40, deadline=None)
@given(ns=st.integers(1, 40), nc=st.integers(1, 40))
def test_tile_order_is_a_bijection_and_runs_outside_in(ns, nc):
    """The barrier-free sweeps rely on this order: positions map one-to-one to tiles, slab-cut chunks come first,
    and the chunks a tile depends on (c-1, c, c+1) sit at most two chunk-groups away in the queue."""
    import ctypes as C
    from mpi_cuda_process_b200.ops import _lib
    lib = _lib.cuda_lib()
    seen = set()
    pos_of_chunk = {}
    for q in range(ns * nc):
        c, s_ = C.c_int(), C.c_int()
        back = lib.mcp_fused2d_tile_at(ns, nc, q, C.byref(c), C.byref(s_))
        assert back == q
        assert 0 <= c.value < nc and 0 <= s_.value < ns
        seen.add((c.value, s_.value))
        pos_of_chunk.setdefault(c.value, q // ns)
    assert len(seen) == ns * nc
    assert pos_of_chunk[0] == 0 and (nc == 1 or pos_of_chunk[nc - 1] == 1)
    for c in range(nc - 1):
        assert abs(pos_of_chunk[c] - pos_of_chunk[c + 1]) <= 2


@settings(max_examples=60, deadline=None)
@given(rows=st.integers(8, 40000), w=st.integers(64, 40000), tb=st.sampled_from([1, 2, 4, 8]), grid=st.sampled_from([18, 148]))
def test_dataflow_schedule_invariants_without_a_gpu(rows, w, tb, grid):
    """Barrier-free schedule (mode 1): constant chunk height with r + 2*TB a whole number of 6-row TMA stages, the last
    chunk between half and 1.5x of it, and -- when the slab allows it -- at least two tiles per warp and sweep."""
    import ctypes as C
    from mpi_cuda_process_b200.ops import _lib
    lib = _lib.cuda_lib()
    ghost = tb
    rows = max(rows, ghost)
    out = (C.c_int * 8192)()
    ns = C.c_int(0)
    n = lib.mcp_fused2d_plan_schedule2(0, 0, w, rows, ghost, tb, grid, 0, 1, out, 8192, C.byref(ns))
    assert 1 <= n <= 4096
    hts = [out[i] for i in range(n)]
    assert sum(hts) == rows and min(hts) >= min(ghost, rows)
    if 1 < n < 4096:                           # (at the 4096-chunk cap the height is rows / 4096, unaligned)
        r = hts[0]
        assert all(x == r for x in hts[:-1]) and (r + 2 * tb) % 6 == 0
        assert r // 2 <= hts[-1] < 2 * r
        warps = grid * (8 if tb == 8 else 16)
        if r > max(ghost, 4 * tb) + 6:          # not already at the minimum tile height
            assert n * ns.value >= 2 * warps * 0.9
