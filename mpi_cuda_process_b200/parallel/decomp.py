"""Domain decomposition: 1-D row slabs and 2-D pencils, with partitioned storage.

The reference hard-codes exactly two row slabs as ``p_id == 0 / p_id == 1`` predicates
inside its kernels, cuts at linear index ``size/2`` (not row-aligned for odd ``h``) and
replicates the whole grid on every rank (MDF_kernel.cu:29-68,137-144,159).  Here the
split is a host-side object: any number of ranks, cuts aligned to rows/planes, each rank
stores only its block plus ``ghost`` layers, and the neighbour table replaces the
hard-wired peer ``1`` of MDF_kernel.cu:168,182,201,215.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple


def split_axis(n: int, parts: int) -> List[Tuple[int, int]]:
    """Contiguous near-equal ranges [start, stop) of ``n`` items over ``parts`` owners."""
    if parts < 1:
        raise ValueError("parts must be >= 1")
    base, rem = divmod(n, parts)
    out, s = [], 0
    for p in range(parts):
        e = s + base + (1 if p < rem else 0)
        out.append((s, e))
        s = e
    return out


def choose_pencil_grid(world: int, depth: int, height: int) -> Tuple[int, int]:
    """(Pz, Py) with Pz*Py == world, minimising exchanged face area for a d x h x w grid."""
    best = None
    for pz in range(1, world + 1):
        if world % pz:
            continue
        py = world // pz
        if pz > depth or py > height:
            continue
        # face area per rank ~ (#z-cuts faces)*(h/py) + (#y-cut faces)*(d/pz)   (times w)
        cost = (0 if pz == 1 else 2) * (height / py) + (0 if py == 1 else 2) * (depth / pz)
        key = (cost, abs(pz - py))
        if best is None or key < best[0]:
            best = (key, (pz, py))
    if best is None:
        raise ValueError(f"cannot place {world} ranks on a {depth}x{height} cross-section")
    return best[1]


@dataclass(frozen=True)
class Block:
    """What one rank owns: per-axis [start, stop) in global coordinates and its neighbours.

    ``neighbors[axis] = (rank_minus, rank_plus)``; None where the block touches the
    physical boundary (the chain is open: the reference has no periodic wrap either).
    """

    rank: int
    coords: Tuple[int, ...]
    start: Tuple[int, ...]
    stop: Tuple[int, ...]
    neighbors: Tuple[Tuple[Optional[int], Optional[int]], ...]

    @property
    def shape(self) -> Tuple[int, ...]:
        return tuple(e - s for s, e in zip(self.start, self.stop))

    @property
    def cells(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n


class Decomposition:
    """Cartesian block decomposition of ``shape`` over a process grid ``pgrid``.

    ``pgrid`` has one entry per axis (1 = axis not split).  Slab = ``(N, 1)`` for 2-D or
    ``(N, 1, 1)`` for 3-D; pencil = ``(Pz, Py, 1)``.  The fastest axis (x / columns) is never
    split: rows stay contiguous so halo rows are single contiguous NVLink stores.
    Ranks are numbered row-major over ``pgrid``.
    """

    def __init__(self, shape: Sequence[int], pgrid: Sequence[int], ghost: int = 1):
        shape, pgrid = tuple(int(s) for s in shape), tuple(int(p) for p in pgrid)
        if len(shape) != len(pgrid):
            raise ValueError("shape and pgrid rank mismatch")
        if pgrid[-1] != 1:
            raise ValueError("the contiguous (last) axis is never split")
        for n, p in zip(shape, pgrid):
            if p < 1 or p > max(n, 1):
                raise ValueError(f"cannot split extent {n} over {p} ranks")
        self.shape, self.pgrid, self.ghost = shape, pgrid, int(ghost)
        self.world = 1
        for p in pgrid:
            self.world *= p
        self._ranges = [split_axis(n, p) for n, p in zip(shape, pgrid)]
        for axis, (rngs, p) in enumerate(zip(self._ranges, pgrid)):
            if p > 1 and min(e - s for s, e in rngs) < self.ghost:
                raise ValueError(
                    f"axis {axis}: a block is thinner than the ghost depth {self.ghost}; "
                    f"use fewer ranks or a smaller time_block")

    # -- constructors ------------------------------------------------------------------
    @classmethod
    def slab(cls, shape: Sequence[int], world: int, ghost: int = 1) -> "Decomposition":
        return cls(shape, (world,) + (1,) * (len(shape) - 1), ghost)

    @classmethod
    def pencil(cls, shape: Sequence[int], world: int, pz_py: Tuple[int, int] = (0, 0), ghost: int = 1) -> "Decomposition":
        if len(shape) != 3:
            raise ValueError("pencil decomposition needs a 3-D grid")
        pz, py = pz_py
        if pz <= 0 or py <= 0:
            pz, py = choose_pencil_grid(world, shape[0], shape[1])
        if pz * py != world:
            raise ValueError(f"pencil grid {pz}x{py} != world {world}")
        return cls(shape, (pz, py, 1), ghost)

    # -- rank <-> coords ---------------------------------------------------------------
    def coords_of(self, rank: int) -> Tuple[int, ...]:
        c = []
        for p in reversed(self.pgrid):
            c.append(rank % p)
            rank //= p
        return tuple(reversed(c))

    def rank_of(self, coords: Sequence[int]) -> int:
        r = 0
        for c, p in zip(coords, self.pgrid):
            r = r * p + c
        return r

    def block(self, rank: int) -> Block:
        if not 0 <= rank < self.world:
            raise ValueError(f"rank {rank} outside world {self.world}")
        coords = self.coords_of(rank)
        start = tuple(self._ranges[a][c][0] for a, c in enumerate(coords))
        stop = tuple(self._ranges[a][c][1] for a, c in enumerate(coords))
        nbrs = []
        for a, c in enumerate(coords):
            lo = hi = None
            if c > 0:
                cc = list(coords); cc[a] = c - 1; lo = self.rank_of(cc)
            if c < self.pgrid[a] - 1:
                cc = list(coords); cc[a] = c + 1; hi = self.rank_of(cc)
            nbrs.append((lo, hi))
        return Block(rank, coords, start, stop, tuple(nbrs))

    def blocks(self) -> List[Block]:
        return [self.block(r) for r in range(self.world)]

    def split_axes(self) -> List[int]:
        return [a for a, p in enumerate(self.pgrid) if p > 1]

    def local_alloc_shape(self, rank: int) -> Tuple[int, ...]:
        """Owned extent plus ``ghost`` layers on both sides of every split axis."""
        b = self.block(rank)
        return tuple(n + (2 * self.ghost if self.pgrid[a] > 1 else 0) for a, n in enumerate(b.shape))

    def ghost_offset(self) -> Tuple[int, ...]:
        return tuple(self.ghost if p > 1 else 0 for p in self.pgrid)

    def global_slices(self, rank: int, with_ghost: bool = False):
        """Slices into the global array of this rank's block (optionally widened by ghosts, clipped)."""
        b = self.block(rank)
        out = []
        for a, (s, e) in enumerate(zip(b.start, b.stop)):
            if with_ghost and self.pgrid[a] > 1:
                s, e = max(0, s - self.ghost), min(self.shape[a], e + self.ghost)
            out.append(slice(s, e))
        return tuple(out)

    def halo_bytes_per_exchange(self, rank: int, itemsize: int) -> int:
        """Bytes this rank sends in one ghost exchange (all faces, depth = ghost)."""
        b = self.block(rank)
        total = 0
        for a, (lo, hi) in enumerate(b.neighbors):
            face = 1
            for a2, n in enumerate(b.shape):
                if a2 != a:
                    face *= n
            total += face * self.ghost * itemsize * ((lo is not None) + (hi is not None))
        return total

    def describe(self) -> str:
        return f"Decomposition(shape={self.shape}, pgrid={self.pgrid}, ghost={self.ghost})"
