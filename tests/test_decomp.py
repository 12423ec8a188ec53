import numpy as np
import pytest

from mpi_cuda_process_b200.parallel.decomp import Decomposition, choose_pencil_grid, split_axis


def test_split_axis_covers_everything():
    for n, p in [(10, 3), (7, 7), (16384, 8), (5, 1)]:
        r = split_axis(n, p)
        assert r[0][0] == 0 and r[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
        sizes = [e - s for s, e in r]
        assert max(sizes) - min(sizes) <= 1


def test_slab_neighbours_open_chain():
    d = Decomposition.slab((16, 8), 4)
    b = [d.block(r) for r in range(4)]
    assert b[0].neighbors[0] == (None, 1) and b[3].neighbors[0] == (2, None)
    assert b[1].neighbors[0] == (0, 2)
    assert all(x.shape == (4, 8) for x in b)
    assert d.split_axes() == [0]


def test_slab_odd_height_is_row_aligned():
    # the reference cuts at linear index size/2, mid-row for odd h; here cuts are row aligned
    d = Decomposition.slab((7, 5), 2)
    assert d.block(0).shape == (4, 5) and d.block(1).shape == (3, 5)


def test_global_slices_reassemble():
    d = Decomposition.slab((13, 6), 3)
    g = np.arange(13 * 6).reshape(13, 6)
    out = np.zeros_like(g)
    for r in range(3):
        out[d.global_slices(r)] = g[d.global_slices(r)]
    assert (out == g).all()


def test_pencil_grid():
    assert choose_pencil_grid(8, 1024, 1024) in [(2, 4), (4, 2)]
    d = Decomposition.pencil((16, 16, 8), 8, (2, 4))
    assert d.world == 8 and d.pgrid == (2, 4, 1)
    b = d.block(5)  # coords (1, 1)
    assert b.coords == (1, 1, 0) and b.start == (8, 4, 0) and b.shape == (8, 4, 8)
    assert b.neighbors[0] == (1, None) and b.neighbors[1] == (4, 6)
    assert d.split_axes() == [0, 1]


def test_too_thin_rejected():
    with pytest.raises(ValueError):
        Decomposition.slab((8, 8), 4, ghost=4)
    with pytest.raises(ValueError):
        Decomposition.slab((2, 8), 4)


def test_halo_bytes():
    d = Decomposition.slab((16384, 16384), 8)
    assert d.halo_bytes_per_exchange(3, 4) == 2 * 16384 * 4
    assert d.halo_bytes_per_exchange(0, 4) == 16384 * 4
