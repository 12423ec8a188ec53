"""NumPy oracles: the *intended* semantics of the reference, restated on the CPU.

The reference as shipped never advances its field (no buffer swap, MDF_kernel.cu:164,177)
and deadlocks for g >= 2 (MDF_kernel.cu:168 vs :215), so "what the reference computes"
is defined here from the parts of its text that express intent:

* MDF  (MDF_kernel.cu:10-22, :88-99): fp32, Dirichlet ring = 100 held fixed, interior
  ``u' = u + 0.25*(uE + uW + uN + uS - 4u)`` with the operand order of MDF_kernel.cu:20,
  Jacobi (two buffers swapped every generation).
* Life (kernel.cu:10-68, :131-146): int32 cells, B3/S23 on the 8-neighbourhood, cells
  outside the grid are dead.

Every function is pure NumPy, elementwise, in the cell dtype, so the rounding sequence is
explicit and the GPU kernels can be compared bit for bit.
"""
from __future__ import annotations

import numpy as np

# --------------------------------------------------------------------------- MDF 2-D


def mdf_init(h: int, w: int, ring: float = 100.0, interior: float = 0.0, dtype=np.float32) -> np.ndarray:
    """``create_universe`` of MDF_kernel.cu:88-99: edge cells = ring, others = interior."""
    u = np.full((h, w), interior, dtype=dtype)
    u[0, :] = ring
    u[-1, :] = ring
    u[:, 0] = ring
    u[:, -1] = ring
    return u


def mdf_init_shape(h: int, w: int, kind: str, ring: float = 100.0, interior: float = 0.0, seed: int = 0,
                   dtype=np.float32) -> np.ndarray:
    """Synthetic MDF fields (extension): "hash" = interior uniform in [interior, ring) from hash_u32(i*w+j, seed)
    (24 bits), "box" = hot box over the central third.  Evaluated in fp64 and cast, exactly like the device kernel."""
    if kind == "ring":
        return mdf_init(h, w, ring, interior, dtype)
    if kind == "hash":
        idx = (np.arange(h, dtype=np.int64)[:, None] * w + np.arange(w, dtype=np.int64)[None, :])
        hsh = hash_u32(idx.astype(np.uint32), seed)
        v = interior + (ring - interior) * ((hsh >> np.uint32(8)).astype(np.float64) * (1.0 / 16777216.0))
    elif kind == "box":
        v = np.full((h, w), float(interior))
        v[h // 3:h - h // 3, w // 3:w - w // 3] = ring
    else:
        raise ValueError(f"unknown init {kind!r}")
    v[0, :] = ring
    v[-1, :] = ring
    v[:, 0] = ring
    v[:, -1] = ring
    return v.astype(dtype)


def mdf_update_cells(C, N, S, W, E):
    """The pinned rounding sequence of MDF_kernel.cu:20 for arrays of one float dtype.

    Reference SASS: FADD(E,W); FADD(+N); FADD(+S); FFMA(C,-4,sum); then an fp64
    DFMA(sum, 0.25, C) rounded back to fp32.  The GPU kernels use true fused multiply-adds:
    ``s = fma(C, -4, s); out = fma(s, 0.25, C)``.

    ``-4*C`` is exact (power-of-two scaling up), so the first fma equals a plain add of the
    exact product.  ``0.25*s`` is exact too *except when it lands in the subnormal range*,
    where the product loses bits while a real fma does not -- and the diffusion front of a
    cold field does pass through subnormal values.  So the last step is evaluated the way
    the reference does it: in fp64 (exact product, exact-enough sum: two 24-bit mantissas
    cannot produce a double-rounding tie) and rounded once to fp32.  For fp64 fields the
    plain expression is used, patched with extended precision where 0.25*s is subnormal.
    """
    dt = C.dtype.type
    s = E + W
    s = s + N
    s = s + S
    s = s + dt(-4.0) * C
    if C.dtype == np.float32:
        return (s.astype(np.float64) * 0.25 + C.astype(np.float64)).astype(np.float32)
    out = dt(0.25) * s + C
    tiny = np.abs(s) < np.finfo(C.dtype).tiny * 8
    if np.any(tiny & (s != 0)):
        ls, lc = s[tiny].astype(np.longdouble), C[tiny].astype(np.longdouble)
        out[tiny] = (ls * np.longdouble(0.25) + lc).astype(C.dtype)
    return out


def mdf_step(u: np.ndarray) -> np.ndarray:
    """One Jacobi generation; the boundary ring is held fixed (Dirichlet)."""
    out = u.copy()
    if u.shape[0] < 3 or u.shape[1] < 3:
        return out
    C = u[1:-1, 1:-1]
    out[1:-1, 1:-1] = mdf_update_cells(C, u[:-2, 1:-1], u[2:, 1:-1], u[1:-1, :-2], u[1:-1, 2:])
    return out


def mdf_run(u: np.ndarray, generations: int) -> np.ndarray:
    for _ in range(generations):
        u = mdf_step(u)
    return u


def mdf_update_cells_ref_fp64(C, N, S, W, E):
    """Literal mixed-precision form of MDF_kernel.cu:20 (fp32 adds, fp64 final fma)."""
    assert C.dtype == np.float32
    s = E + W
    s = s + N
    s = s + S
    s = s + np.float32(-4.0) * C
    return (s.astype(np.float64) * 0.25 + C.astype(np.float64)).astype(np.float32)


def mdf_update_cells_simplified(C, N, S, W, E):
    """The algebraically simplified ``0.25*(E+W+N+S)`` -- NOT bit-equal; kept to test that claim."""
    dt = C.dtype.type
    return dt(0.25) * (((E + W) + N) + S)


# --------------------------------------------------------------------------- residuals


def residual(u_new: np.ndarray, u_old: np.ndarray, kind: str = "linf") -> float:
    """Global residual of one generation: the difference is formed in the field's own dtype (what every kernel
    does), its magnitude / square are reduced in fp64."""
    d = (u_new - u_old).astype(np.float64)
    if kind == "linf":
        return float(np.max(np.abs(d))) if d.size else 0.0
    if kind == "l2":
        return float(np.sqrt(np.sum(d * d)))
    raise ValueError(kind)


# --------------------------------------------------------------------------- Game of Life

_RAND_MAX = 2147483647


class GlibcRand:
    """glibc 2.x ``rand()`` (TYPE_3 additive feedback generator), default seed 1.

    kernel.cu:141 calls ``rand()`` without ever seeding it, so every run (and every rank)
    sees this exact stream; first value 1804289383.
    """

    def __init__(self, seed: int = 1):
        seed = seed or 1
        r = [0] * 34
        r[0] = seed
        for i in range(1, 31):
            hi, lo = divmod(r[i - 1], 127773)
            word = 16807 * lo - 2836 * hi
            if word < 0:
                word += 2147483647
            r[i] = word
        for i in range(31, 34):
            r[i] = r[i - 31]
        self._r = [x & 0xFFFFFFFF for x in r]
        for _ in range(34, 344):
            self._advance()

    def _advance(self) -> int:
        r = self._r
        v = (r[-31] + r[-3]) & 0xFFFFFFFF
        r.append(v)
        if len(r) > 64:
            del r[: len(r) - 34]
        return v

    def next(self) -> int:
        return self._advance() >> 1

    def take(self, n: int) -> np.ndarray:
        return np.fromiter((self.next() for _ in range(n)), dtype=np.int64, count=n)


def life_init_reference(h: int, w: int, prob: float = 0.15, seed: int = 1) -> np.ndarray:
    """``create_universe`` of kernel.cu:131-146 with the unseeded glibc stream.

    One ``rand()`` per interior cell in row-major order; alive iff
    ``(float)rand()/RAND_MAX <= prob`` evaluated in fp32.
    """
    u = np.zeros((h, w), dtype=np.int32)
    if h < 3 or w < 3:
        return u
    n = (h - 2) * (w - 2)
    r = GlibcRand(seed).take(n)
    rp = r.astype(np.float32) / np.float32(_RAND_MAX)
    alive = ~(rp > np.float32(prob))
    u[1:-1, 1:-1] = alive.reshape(h - 2, w - 2).astype(np.int32)
    return u


def hash_u32(x: np.ndarray, seed: int) -> np.ndarray:
    """Stateless 32-bit mixer used by the device init kernels for seeded synthetic fields.

    (lowbias32-style avalanche over ``index ^ seed*golden``; restated here so device init
    can be checked exactly.)
    """
    x = (x.astype(np.uint64) ^ np.uint64((seed * 0x9E3779B9) & 0xFFFFFFFF)).astype(np.uint32)
    x = x ^ (x >> np.uint32(16))
    x = (x.astype(np.uint64) * np.uint64(0x7FEB352D)).astype(np.uint32)
    x = x ^ (x >> np.uint32(15))
    x = (x.astype(np.uint64) * np.uint64(0x846CA68B)).astype(np.uint32)
    x = x ^ (x >> np.uint32(16))
    return x


def life_init_hash(h: int, w: int, prob: float = 0.15, seed: int = 0) -> np.ndarray:
    """Seeded synthetic init: ring dead, interior alive iff hash(i*w+j) * 2^-32 < prob."""
    idx = (np.arange(h, dtype=np.int64)[:, None] * w + np.arange(w, dtype=np.int64)[None, :])
    hsh = hash_u32(idx.astype(np.uint32), seed)
    thr = np.uint32(min(int(prob * 4294967296.0), 0xFFFFFFFF))
    u = (hsh < thr).astype(np.int32)
    u[0, :] = 0
    u[-1, :] = 0
    u[:, 0] = 0
    u[:, -1] = 0
    return u


def life_step(u: np.ndarray, bc: str = "dead-outside") -> np.ndarray:
    """One generation of B3/S23 (kernel.cu:61,66); cells outside the grid are dead.

    bc="dead-outside": every cell evolves.  bc="fixed-ring": the outer ring keeps its value.
    """
    h, w = u.shape
    p = np.zeros((h + 2, w + 2), dtype=np.int32)
    p[1:-1, 1:-1] = u
    n = (p[:-2, :-2] + p[:-2, 1:-1] + p[:-2, 2:] + p[1:-1, :-2] + p[1:-1, 2:]
         + p[2:, :-2] + p[2:, 1:-1] + p[2:, 2:])
    out = ((n == 3) | ((n == 2) & (u != 0))).astype(np.int32)
    if bc == "fixed-ring":
        out[0, :] = u[0, :]
        out[-1, :] = u[-1, :]
        out[:, 0] = u[:, 0]
        out[:, -1] = u[:, -1]
    elif bc != "dead-outside":
        raise ValueError(bc)
    return out


def life_run(u: np.ndarray, generations: int, bc: str = "dead-outside") -> np.ndarray:
    for _ in range(generations):
        u = life_step(u, bc)
    return u


# --------------------------------------------------------------------------- heat 3-D (extension)

HEAT3D_COEFF = 1.0 / 6.0


def heat3d_init(d: int, h: int, w: int, ring: float = 100.0, interior: float = 0.0, dtype=np.float32) -> np.ndarray:
    """3-D analogue of MDF_kernel.cu:88-99: boundary shell = ring, interior = interior."""
    u = np.full((d, h, w), interior, dtype=dtype)
    u[0], u[-1] = ring, ring
    u[:, 0], u[:, -1] = ring, ring
    u[:, :, 0], u[:, :, -1] = ring, ring
    return u


def heat3d_update_cells(C, N, S, W, E, U, D, coeff=HEAT3D_COEFF):
    """7-point FTCS step, every operation individually rounded (no fused multiply-add):

    ``s = ((((E+W)+N)+S)+U)+D ; d = s - 6*C ; out = C + coeff*d``

    (x: W/E, y: N/S, z: U/D).  6*C and coeff*d are not exact, so the device kernel uses
    __fmul_rn/__fadd_rn to forbid contraction and match this bit for bit.
    """
    dt = C.dtype.type
    s = E + W
    s = s + N
    s = s + S
    s = s + U
    s = s + D
    d = s - dt(6.0) * C
    return C + dt(coeff) * d


def heat3d_step(u: np.ndarray, coeff=HEAT3D_COEFF) -> np.ndarray:
    out = u.copy()
    if min(u.shape) < 3:
        return out
    C = u[1:-1, 1:-1, 1:-1]
    out[1:-1, 1:-1, 1:-1] = heat3d_update_cells(
        C,
        u[1:-1, :-2, 1:-1], u[1:-1, 2:, 1:-1],
        u[1:-1, 1:-1, :-2], u[1:-1, 1:-1, 2:],
        u[:-2, 1:-1, 1:-1], u[2:, 1:-1, 1:-1], coeff)
    return out


def heat3d_run(u: np.ndarray, generations: int, coeff=HEAT3D_COEFF) -> np.ndarray:
    for _ in range(generations):
        u = heat3d_step(u, coeff)
    return u


# --------------------------------------------------------------------------- dispatch


def init_field(cfg) -> np.ndarray:
    """Global initial field for a RunConfig (single-domain)."""
    if cfg.workload == "mdf":
        dt = np.float32 if cfg.dtype == "fp32" else np.float64
        return mdf_init_shape(cfg.height, cfg.width, cfg.init, cfg.ring, cfg.interior, cfg.seed or 0, dt)
    if cfg.workload == "life":
        if cfg.seed is None:
            return life_init_reference(cfg.height, cfg.width, cfg.prob)
        return life_init_hash(cfg.height, cfg.width, cfg.prob, cfg.seed)
    if cfg.workload == "heat3d":
        dt = np.float32 if cfg.dtype == "fp32" else np.float64
        return heat3d_init(cfg.depth, cfg.height, cfg.width, cfg.ring, cfg.interior, dt)
    raise ValueError(cfg.workload)


def step_field(cfg, u: np.ndarray) -> np.ndarray:
    if cfg.workload == "mdf":
        return mdf_step(u)
    if cfg.workload == "life":
        return life_step(u, cfg.bc)
    return heat3d_step(u)


def run_field(cfg, u: np.ndarray, generations: int) -> np.ndarray:
    for _ in range(generations):
        u = step_field(cfg, u)
    return u


def checksum_rows(rows: np.ndarray, grow0: int, w: int) -> int:
    """NumPy restatement of ``mcp_checksum2d`` (ops/csrc/checksum.cu): wrapping sum over cells of
    mix64(bit pattern ^ golden * global cell index)."""
    a = np.ascontiguousarray(rows)
    bits = a.view(np.uint64 if a.dtype.itemsize == 8 else np.uint32).astype(np.uint64)
    h = a.shape[0]
    with np.errstate(over="ignore"):
        idx = (np.arange(h, dtype=np.uint64)[:, None] + np.uint64(grow0)) * np.uint64(w) + np.arange(w, dtype=np.uint64)[None, :]
        z = bits ^ (idx * np.uint64(0x9E3779B97F4A7C15))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
        return int(z.sum(dtype=np.uint64))


def checksum_field(field: np.ndarray) -> int:
    return checksum_rows(field, 0, field.shape[1])
