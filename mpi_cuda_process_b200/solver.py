"""Solver: the generation loop, for every backend.

Reference counterpart: the two copies of the ``while (iter_count > 0)`` loop in ``main()`` -- rank 0 at
MDF_kernel.cu:155-188, "everyone else" at MDF_kernel.cu:189-222 (kernel.cu:202-269) -- which per generation
upload the whole grid, launch middle_kernel, receive the neighbour's border row element by element, upload
the whole grid again, launch border_kernel, synchronise, download the whole grid and send the own border
row element by element.  As written that loop never swaps its buffers (MDF_kernel.cu:164), copies the
input back (MDF_kernel.cu:177) and deadlocks for g >= 2 (MDF_kernel.cu:168 vs 215); the Solver implements
the intended semantics with one code path for every rank.

Backends
  fused   persistent sm_100a kernel: TMA-fed, temporally blocked, halo rows stored into the neighbour
          GPUs over NVLink inside the kernel, device-side flags and grid barrier (the product)
  nccl    plain kernel per generation + grouped NCCL send/recv of ghost rows (baseline)
  staged  the reference's host-staged pattern: D2H, host messages, H2D every generation (baseline)
  cpu     torch CPU tensors + gloo (plumbing / correctness without a GPU)
"""
from __future__ import annotations

import ctypes as C
import time
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import numpy as np
import torch

from . import models, oracle
from .config import RunConfig
from .parallel.context import DistContext, init_context
from .parallel.decomp import Decomposition
from .parallel import host_p2p

_TORCH_DTYPE = {"fp32": torch.float32, "fp64": torch.float64, "int32": torch.int32, "bits": torch.int32}
_NP_DTYPE = {"fp32": np.float32, "fp64": np.float64, "int32": np.int32, "bits": np.int32}
_MCP_DTYPE = {"fp32": 0, "fp64": 1, "int32": 2, "bits": 3}


@dataclass
class RunStats:
    generations: int = 0
    device_ms: float = 0.0          # max over ranks, CUDA events (wall clock on the CPU backend)
    cells: int = 0
    launches: int = 0
    halo_messages: int = 0
    residual_history: List[Tuple[int, float]] = field(default_factory=list)
    stop_reason: str = ""           # run_to_tolerance: "converged" | "max_generations" | "slot_limit"

    @property
    def glups(self) -> float:
        return (self.cells * self.generations) / (self.device_ms * 1e-3) / 1e9 if self.device_ms > 0 else 0.0


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class PendingResult:
    """Result of a solve that has been enqueued but not waited for (:meth:`Solver.run_async`): the kernel status and
    the global residual of the last generation travel back in one asynchronous 32-byte copy; :meth:`wait` blocks on
    its event only.  Lets a host loop enqueue step i+1 before it looks at step i, so that the ranks' launches do not
    drift apart by a host round trip per step."""

    def __init__(self, solver, slot, event, stats, value=None):
        self._solver, self._slot, self._event, self.stats, self._value = solver, slot, event, stats, value

    def wait(self) -> Optional[float]:
        if self._event is not None:
            self._event.synchronize()
            self._event = None
            self._value = self._solver._decode_result(self._slot, self.stats)
        return self._value


class _ResultRing:
    """Bookkeeping of the pinned result slots: a fixed ring, so a slot may still belong to a handle nobody has waited
    for when its turn comes again -- that handle is settled (its value read out) before the slot is handed on."""

    def __init__(self, slots: int):
        self.slots, self.next, self.owner = slots, 0, {}

    def claim(self) -> int:
        idx = self.next % self.slots
        self.next += 1
        older = self.owner.pop(idx, None)
        if older is not None:
            older.wait()
        return idx

    def bind(self, idx: int, handle):
        self.owner[idx] = handle
        return handle


class Solver:
    """Distributed stencil solver.  Use as a context manager or call :meth:`close`."""

    def __init__(self, cfg: RunConfig, ctx: Optional[DistContext] = None, device: Optional[str] = None):
        self.cfg = cfg
        self._own_ctx = ctx is None
        self.ctx = ctx if ctx is not None else init_context(device="cpu" if cfg.backend == "cpu" else device)
        self.backend = self._pick_backend()
        self.model = models.get(cfg.workload)    # rule, operator code, time-block limits, boundary rule
        if cfg.dtype not in self.model.dtypes:
            raise ValueError(f"workload {cfg.workload!r} stores {self.model.dtypes}, not {cfg.dtype!r}")
        self.bits = cfg.dtype == "bits"          # bit-packed life: 32 cells per stored 32-bit word
        if self.bits and self.backend != "fused":
            raise ValueError("dtype='bits' (bit-packed life) runs on the fused GPU backend only")
        self.ndim = cfg.ndim
        self.tdtype = _TORCH_DTYPE[cfg.dtype]
        self.npdtype = _NP_DTYPE[cfg.dtype]
        self.esize = 8 if cfg.dtype == "fp64" else 4
        # ghost depth: the fused kernel exchanges `time_block` rows every `time_block` generations
        if self.backend == "fused" and self.ndim == 2:
            tbmax = self.model.time_block_limit(cfg.dtype)
            tb = 1
            while tb * 2 <= min(cfg.time_block, tbmax):
                tb *= 2
            self.time_block = tb
            self.ghost = tb
        else:
            self.time_block = 1
            self.ghost = 1
        self.decomp = self._make_decomp()
        self.block = self.decomp.block(self.ctx.rank)
        self.generation = 0
        self.total_stats = RunStats(cells=cfg.cells)
        self._arena = None
        self._plan = None
        self._lib = None
        self._closed = False
        self._ghosts_valid = True
        self._last_residual: Optional[Tuple[float, float]] = None
        self.last_gather_stats: Optional[dict] = None
        self._scratch = None
        self._alloc()

    # ------------------------------------------------------------------ setup
    def _pick_backend(self) -> str:
        b = self.cfg.backend
        if b == "auto":
            if not self.ctx.is_cuda:
                return "cpu"
            return "fused"
        if b != "cpu" and not self.ctx.is_cuda:
            raise RuntimeError(f"backend {b!r} needs a GPU")
        return b

    def _make_decomp(self) -> Decomposition:
        cfg, world = self.cfg, self.ctx.world
        if cfg.ndim == 3 and cfg.decomposition == "pencil":
            return Decomposition.pencil(cfg.shape, world, cfg.pencil, ghost=self.ghost)
        return Decomposition.slab(cfg.shape, world, ghost=self.ghost)

    @property
    def _fix_bounds(self):
        """Inclusive global box of cells that evolve: (ylo, yhi, xlo, xhi)."""
        return self.model.evolving_box(self.cfg)

    def _alloc(self) -> None:
        cfg, G = self.cfg, self.ghost
        own = self.block.shape
        self.own_shape = own
        # ghost layers on every non-contiguous axis (unused at the physical boundary: stay zero)
        self.alloc_shape = tuple(n + 2 * G for n in own[:-1]) + (own[-1],)
        if self.backend == "fused" and self.ndim == 3:
            # the 3-D kernel stores into neighbours' buffers: every rank uses the shape of the largest block so
            # that plane and row strides agree across ranks
            mx = [max(b.shape[a] for b in self.decomp.blocks()) for a in range(self.ndim)]
            self.alloc_shape = tuple(n + 2 * G for n in mx[:-1]) + (own[-1],)
        if self.ctx.is_cuda:
            from .ops import _lib
            self._lib = _lib.cuda_lib()
            self.pitch = _round_up((cfg.width + 31) // 32, 32) if self.bits else _round_up(cfg.width, 32)
        else:
            self.pitch = cfg.width
        store_shape = self.alloc_shape[:-1] + (self.pitch,)
        self.store_shape = store_shape
        if self.ctx.is_cuda:
            from .parallel.arena import SymmetricArena
            # identical layout on every rank: size buffers for the largest block
            max_own = [max(b.shape[a] for b in self.decomp.blocks()) for a in range(self.ndim)]
            max_alloc = 1
            for n in max_own[:-1]:
                max_alloc *= n + 2 * G
            self._buf_bytes = _round_up(max_alloc * self.pitch * self.esize, 4096)
            # signalling pad in front of the buffers: halo counters @0, residual arrival counter @256,
            # residual inbox ring (64 slots x 16 ranks x 16 B) @4096
            self._flags_bytes = 65536        # + per-strip halo flags (2 x 2048 x 8 B) @32768
            need_peers = self.backend == "fused" and self.ctx.distributed
            # halo pushes need the two neighbours; the fused residual all-reduce stores to every rank
            peers = list(range(self.ctx.world)) if need_peers else []
            if need_peers and self.ctx.world > 16:
                raise ValueError("the fused backend supports up to 16 ranks (MCP_MAX_RANKS)")
            total = self._flags_bytes + 2 * self._buf_bytes
            if need_peers:
                self._arena = SymmetricArena(self.ctx, total, peers=peers)
            else:
                self._arena = SymmetricArena(_SoloCtx(self.ctx), total)
            self.bufs = [self._arena.view(self._flags_bytes + i * self._buf_bytes, store_shape, self.tdtype)
                         for i in range(2)]
            self._scratch = torch.empty(max(1, self._lib.mcp_residual2d_scratch_bytes() // 8) + 2,
                                        dtype=torch.float64, device=self.ctx.device)
        else:
            self.bufs = [torch.zeros(store_shape, dtype=self.tdtype) for _ in range(2)]
        self.cur = 0

    def _buf_ptr(self, i: int, rank: Optional[int] = None) -> int:
        off = self._flags_bytes + i * self._buf_bytes
        return self._arena.peer_ptr(self.ctx.rank if rank is None else rank, off)

    def _stream(self) -> C.c_void_p:
        return C.c_void_p(torch.cuda.current_stream(self.ctx.device).cuda_stream)

    def _ensure_plan3d(self):
        from .ops import _lib
        cfg = self.cfg
        d = _lib.Fused3dDesc()
        d.dtype = _MCP_DTYPE[cfg.dtype]
        d.d, d.h, d.w = cfg.depth, cfg.height, cfg.width
        d.z_start, d.nz = self.block.start[0], self.own_shape[0]
        d.y_start, d.ny = self.block.start[1], self.own_shape[1]
        d.pitch = self.pitch
        d.pz_alloc, d.py_alloc = self.store_shape[0], self.store_shape[1]
        zchunk = int(cfg.extra.get("zchunk", 0))
        d.zchunk = zchunk
        nbrs = [self.block.neighbors[0][0], self.block.neighbors[0][1],
                self.block.neighbors[1][0], self.block.neighbors[1][1]]
        opposite = [1, 0, 3, 2]
        counts = (C.c_int * 4)()
        for i in range(2):
            d.buf[i] = self._buf_ptr(i)
        for f, r in enumerate(nbrs):
            if r is None:
                continue
            nb = self.decomp.block(r)
            for i in range(2):
                d.nbr_buf[f][i] = self._buf_ptr(i, r)
            d.nbr_n[f] = nb.shape[0] if f < 2 else nb.shape[1]
            d.nbr_flag[f] = self._arena.peer_ptr(r, 512 + 8 * opposite[f])
            _lib.check(self._lib.mcp_fused3d_face_tiles(nb.shape[0], nb.shape[1], cfg.width, d.dtype, zchunk, counts))
            d.expect[f] = counts[opposite[f]]
        d.flags = self._arena.peer_ptr(self.ctx.rank, 512)
        d.coeff = oracle.HEAT3D_COEFF
        d.world, d.rank = self.ctx.world, self.ctx.rank
        self._res_in_kernel = False
        if self.ctx.distributed and not cfg.extra.get("host_residual", False):
            # same pad layout as the 2-D kernel: residual arrival counter @256, inbox ring @4096
            d.res_inbox = self._arena.peer_ptr(self.ctx.rank, 4096)
            d.res_flag = self._arena.peer_ptr(self.ctx.rank, 256)
            for r in range(self.ctx.world):
                d.peer_res_inbox[r] = self._arena.peer_ptr(r, 4096)
                d.peer_res_flag[r] = self._arena.peer_ptr(r, 256)
            self._res_in_kernel = True
        plan = C.c_void_p()
        _lib.check(self._lib.mcp_fused3d_create(C.byref(d), C.byref(plan)), "mcp_fused3d_create")
        self._plan = plan
        self._desc = d
        if cfg.extra.get("timeout_ms") is not None:
            _lib.check(self._lib.mcp_fused3d_set_timeout_ms(plan, int(cfg.extra["timeout_ms"])))
        self._result_pool = torch.zeros((8, 4), dtype=torch.float64).pin_memory()

    def _ensure_plan(self):
        if self._plan is not None:
            return
        if self.ndim == 3:
            return self._ensure_plan3d()
        from .ops import _lib
        cfg, G = self.cfg, self.ghost
        d = _lib.FusedDesc()
        d.op, d.dtype = self.model.op_code, _MCP_DTYPE[cfg.dtype]
        d.h, d.w = cfg.height, cfg.width
        d.row_start, d.rows_own, d.ghost = self.block.start[0], self.own_shape[0], G
        d.pitch = self.pitch
        up, dn = self.block.neighbors[0]
        for i in range(2):
            d.buf[i] = self._buf_ptr(i)
            d.up_buf[i] = self._buf_ptr(i, up) if up is not None else None
            d.dn_buf[i] = self._buf_ptr(i, dn) if dn is not None else None
        d.up_rows_own = self.decomp.block(up).shape[0] if up is not None else 0
        d.flags = self._arena.peer_ptr(self.ctx.rank, 0)
        d.up_flag = self._arena.peer_ptr(up, 8) if up is not None else None      # up's flags[1]
        d.dn_flag = self._arena.peer_ptr(dn, 0) if dn is not None else None      # down's flags[0]
        if self.ctx.distributed and not cfg.extra.get("coarse_halo_flags", False):
            d.strip_flags = self._arena.peer_ptr(self.ctx.rank, 32768)
            d.up_strip_flag = self._arena.peer_ptr(up, 32768 + 2048 * 8) if up is not None else None
            d.dn_strip_flag = self._arena.peer_ptr(dn, 32768) if dn is not None else None
        d.ylo, d.yhi, d.xlo, d.xhi = self._fix_bounds
        d.chunk_rows = int(cfg.extra.get("chunk_rows", 0))
        d.max_ctas = int(cfg.extra.get("max_ctas", 0))
        d.world, d.rank = self.ctx.world, self.ctx.rank
        self._res_in_kernel = False
        if self.ctx.distributed and cfg.extra.get("residual_allreduce") == "nvls":
            # NVLS variant: a multicast-mapped pad from torch's symmetric memory (bootstrap only; the reduction
            # itself is multimem.* instructions inside the sweep kernel)
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm
            self._nvls_pad = symm.empty(4096, dtype=torch.uint8, device=self.ctx.device)
            self._nvls_pad.zero_()
            hdl = symm.rendezvous(self._nvls_pad, dist.group.WORLD.group_name)
            if not hdl.multicast_ptr:
                raise RuntimeError("this system exposes no multicast (NVLS) mapping")
            torch.cuda.synchronize(self.ctx.device)
            self.ctx.barrier()
            d.nvls_local = self._nvls_pad.data_ptr()
            d.nvls_mc = hdl.multicast_ptr
            self._nvls_handle = hdl
            self._res_in_kernel = True
        elif self.ctx.distributed and not cfg.extra.get("host_residual", False):
            d.res_inbox = self._arena.peer_ptr(self.ctx.rank, 4096)
            d.res_flag = self._arena.peer_ptr(self.ctx.rank, 256)
            for r in range(self.ctx.world):
                d.peer_res_inbox[r] = self._arena.peer_ptr(r, 4096)
                d.peer_res_flag[r] = self._arena.peer_ptr(r, 256)
            self._res_in_kernel = True
        plan = C.c_void_p()
        _lib.check(self._lib.mcp_fused2d_create(C.byref(d), C.byref(plan)), "mcp_fused2d_create")
        self._plan = plan
        self._desc = d
        # plan options: explicit (cfg.extra), no environment lookups inside the native library
        for key, opt in (("timeout_ms", 1), ("debug_flags", 2), ("sync_mode", 3), ("prefetch", 4)):
            if cfg.extra.get(key) is not None:
                _lib.check(self._lib.mcp_fused2d_set_option(plan, opt, int(cfg.extra[key])), "mcp_fused2d_set_option")
        # one pinned 32-byte block for (status, stop word, sweeps done, last residual): read back with one async copy
        self._result_pool = torch.zeros((8, 4), dtype=torch.float64).pin_memory()

    # ------------------------------------------------------------------ initial condition
    def initialize(self, field: Optional[np.ndarray] = None) -> None:
        """Set generation 0.  ``field`` (global array, same on every rank) overrides the config's rule."""
        cfg = self.cfg
        if field is None and self.ctx.is_cuda and (cfg.workload != "life" or cfg.seed is not None):
            self._device_init()
        else:
            if field is None:
                field = self._host_init_global()
            self.load_global(field)
        self.generation = 0
        self._ghosts_valid = True
        self._reset_plan_parity()
        self.ctx.barrier()

    def _host_init_global(self) -> np.ndarray:
        cfg = self.cfg
        if cfg.workload == "life" and cfg.seed is None and cfg.height * cfg.width > 1 << 16:
            from .ops import _lib
            u = np.zeros((cfg.height, cfg.width), dtype=np.int32)
            _lib.host_lib().mcp_host_life_init_reference(u.ctypes.data, cfg.height, cfg.width, C.c_float(cfg.prob))
            return u
        return oracle.init_field(cfg)

    def _device_init(self) -> None:
        from .ops import _lib
        cfg, G = self.cfg, self.ghost
        s = self._stream()
        for b in self.bufs:
            if self.ndim == 2:
                row0 = self.block.start[0] - G
                if self.bits:
                    _lib.check(self._lib.mcp_init_life_hash_bits(b.data_ptr(), self.pitch, self.store_shape[0], row0,
                                                                 cfg.height, cfg.width, cfg.prob, cfg.seed, s))
                elif cfg.workload == "life":
                    _lib.check(self._lib.mcp_init_life_hash(b.data_ptr(), self.pitch, self.store_shape[0], row0,
                                                            cfg.height, cfg.width, cfg.prob, cfg.seed, s))
                elif cfg.workload == "mdf" and cfg.init != "ring":
                    _lib.check(self._lib.mcp_init_shape2d(_MCP_DTYPE[cfg.dtype], 1 if cfg.init == "hash" else 2,
                                                          b.data_ptr(), self.pitch, self.store_shape[0], row0,
                                                          cfg.height, cfg.width, cfg.ring, cfg.interior,
                                                          cfg.seed or 0, s))
                else:
                    _lib.check(self._lib.mcp_init_ring2d(_MCP_DTYPE[cfg.dtype], b.data_ptr(), self.pitch,
                                                         self.store_shape[0], row0, cfg.height, cfg.width,
                                                         cfg.ring, cfg.interior, s))
            else:
                z0, y0 = self.block.start[0] - G, self.block.start[1] - G
                _lib.check(self._lib.mcp_init_shell3d(_MCP_DTYPE[cfg.dtype], b.data_ptr(), self.pitch,
                                                      self.store_shape[1] * self.pitch, self.store_shape[0],
                                                      self.store_shape[1], z0, y0, cfg.depth, cfg.height, cfg.width,
                                                      cfg.ring, cfg.interior, s))
        self.cur = 0
        torch.cuda.synchronize(self.ctx.device)

    def _local_from_global(self, field: np.ndarray) -> np.ndarray:
        """This rank's allocation (owned block + ghost layers, zero outside the domain) cut from a global array."""
        G = self.ghost
        cells_shape = self.store_shape[:-1] + (self.cfg.width,) if self.bits else self.store_shape
        loc = np.zeros(cells_shape, dtype=self.npdtype)
        src_sl, dst_sl = [], []
        for a in range(self.ndim - 1):
            s, e = self.block.start[a] - G, self.block.stop[a] + G
            cs, ce = max(s, 0), min(e, self.cfg.shape[a])
            src_sl.append(slice(cs, ce))
            dst_sl.append(slice(cs - s, ce - s))
        src_sl.append(slice(0, self.cfg.width))
        dst_sl.append(slice(0, self.cfg.width))
        loc[tuple(dst_sl)] = field[tuple(src_sl)]
        return self._pack_bits(loc) if self.bits else loc

    def _pack_bits(self, cells: np.ndarray) -> np.ndarray:
        """(rows, w) cells in {0,1} -> (rows, pitch) int32 words, cell x = bit x%32 of word x//32."""
        b = np.packbits((cells != 0).astype(np.uint8), axis=-1, bitorder="little")
        out = np.zeros(cells.shape[:-1] + (self.pitch * 4,), dtype=np.uint8)
        out[..., :b.shape[-1]] = b
        return out.view("<u4").astype(np.uint32).view(np.int32).reshape(cells.shape[:-1] + (self.pitch,))

    def _unpack_bits(self, words: np.ndarray) -> np.ndarray:
        b = np.ascontiguousarray(words).view(np.uint8)
        return np.unpackbits(b, axis=-1, bitorder="little")[..., :self.cfg.width].astype(np.int32)

    def load_global(self, field: np.ndarray) -> None:
        if tuple(field.shape) != tuple(self.cfg.shape):
            raise ValueError(f"field shape {field.shape} != configured {self.cfg.shape}")
        loc = torch.from_numpy(self._local_from_global(np.asarray(field, dtype=self.npdtype)))
        for b in self.bufs:
            b.copy_(loc)
        self.cur = 0
        if self.ctx.is_cuda:
            torch.cuda.synchronize(self.ctx.device)

    def load_local(self, local: torch.Tensor, non_blocking: bool = True) -> None:
        """Upload this rank's slab (owned rows + ghost rows, shape ``store_shape``) from (pinned) host memory
        into the current buffer.  This is the public per-step input path timed by ``bench.py``'s e2e number."""
        if tuple(local.shape) != tuple(self.store_shape):
            raise ValueError(f"local shape {tuple(local.shape)} != {self.store_shape}")
        if self.backend == "fused" and self.ctx.distributed and self.ndim == 3:
            # a neighbour's last halo push of the previous run may still be in flight towards our ghost layers
            # (the 2-D kernel drains its incoming pushes before it ends, so stream order is enough there)
            self.ctx.barrier()
        self.bufs[self.cur].copy_(local, non_blocking=non_blocking)
        self._ghosts_valid = True

    def stage_input(self, local: torch.Tensor) -> None:
        """Start an asynchronous upload of the NEXT input slab (pinned host memory, shape ``store_shape``) on a
        side stream into a device staging buffer; it overlaps whatever the solver is computing.  Pair with
        :meth:`commit_input`."""
        if tuple(local.shape) != tuple(self.store_shape):
            raise ValueError(f"local shape {tuple(local.shape)} != {self.store_shape}")
        if not self.ctx.is_cuda:
            self._staged = local.clone()
            return
        if getattr(self, "_stage_buf", None) is None:
            self._stage_buf = torch.empty(self.store_shape, dtype=self.tdtype, device=self.ctx.device)
            self._copy_stream = torch.cuda.Stream(device=self.ctx.device)
            self._stage_done = torch.cuda.Event(enable_timing=True)
            self._stage_begin = torch.cuda.Event(enable_timing=True)
            self._stage_free = torch.cuda.Event()
            self._stage_free.record()
        self._copy_stream.wait_event(self._stage_free)          # previous commit has consumed the staging buffer
        with torch.cuda.stream(self._copy_stream):
            self._stage_begin.record(self._copy_stream)
            self._stage_buf.copy_(local, non_blocking=True)
            self._stage_done.record(self._copy_stream)

    def last_upload_ms(self) -> Optional[float]:
        """Device time of the last completed :meth:`stage_input` copy (host -> device), or None."""
        if getattr(self, "_stage_buf", None) is None:
            return None
        self._stage_done.synchronize()
        return float(self._stage_begin.elapsed_time(self._stage_done))

    def commit_input(self) -> None:
        """Make the staged slab the current state (device-to-device copy once the upload has landed)."""
        if not self.ctx.is_cuda:
            self.bufs[self.cur].copy_(self._staged)
            self._ghosts_valid = True
            return
        if self.backend == "fused" and self.ctx.distributed and self.ndim == 3:
            self.ctx.barrier()      # neighbours' last halo pushes of the previous run must have landed first
        # 2-D fused: no host barrier -- before a sweep kernel ends it waits for the flags of every halo push the
        # neighbours address to this rank (drain_incoming), so the copy below is ordered behind them by the stream
        cur = torch.cuda.current_stream(self.ctx.device)
        cur.wait_event(self._stage_done)
        self.bufs[self.cur].copy_(self._stage_buf, non_blocking=True)
        self._stage_free.record(cur)
        self._ghosts_valid = True

    def _reset_plan_parity(self):
        if self._plan is not None:
            from .ops import _lib
            setp = self._lib.mcp_fused3d_set_parity if self.ndim == 3 else self._lib.mcp_fused2d_set_parity
            _lib.check(setp(self._plan, self.cur))

    # ------------------------------------------------------------------ stepping
    def run(self, generations: Optional[int] = None, residual=False, timed: bool = True) -> RunStats:
        """Advance ``generations`` generations (default: the configured count).

        ``residual``: False = none; True / "history" = global residual of the last generation of every time block
        (every generation on the unfused backends); "final" = only the residual of the very last generation
        (the sweep kernel runs its cheaper residual-free instantiation for all but the last time block)."""
        g = self.cfg.generations if generations is None else int(generations)
        self._res_final = False
        if residual == "final" and self.backend == "fused":
            self._res_final = True      # one launch: only the last time block runs the residual row loop
            residual = True
        if residual == "final":
            tail = self.time_block if self.backend == "fused" else 1
            if g > tail:
                st0 = self.run(g - tail, residual=False, timed=timed)
                st1 = self.run(tail, residual=True, timed=timed)
                st1.generations += st0.generations
                st1.device_ms += st0.device_ms
                st1.launches += st0.launches
                st1.halo_messages += st0.halo_messages
                return st1
            residual = True
        residual = bool(residual)
        st = RunStats(generations=g, cells=self.cfg.cells)
        if g <= 0:
            return st
        if self.ctx.is_cuda:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if timed:
                ev0.record()
        else:
            t0 = time.perf_counter()
        if self.ctx.is_cuda:
            torch.cuda.nvtx.range_push(f"mcp.run[{self.backend}] {self.cfg.workload} g={g}")
        if self.backend == "fused":
            self._run_fused(g, residual, st)
        elif self.backend == "nccl":
            self._run_nccl(g, residual, st)
        elif self.backend == "staged":
            self._run_staged(g, residual, st)
        else:
            self._run_cpu(g, residual, st)
        if self.ctx.is_cuda:
            torch.cuda.nvtx.range_pop()
            if timed:
                ev1.record()
                ev1.synchronize()
                st.device_ms = self.ctx.all_reduce_max(ev0.elapsed_time(ev1)) if self.ctx.distributed else ev0.elapsed_time(ev1)
        else:
            st.device_ms = self.ctx.all_reduce_max((time.perf_counter() - t0) * 1e3)
        self.generation += g
        self.total_stats.generations += g
        self.total_stats.device_ms += st.device_ms
        self.total_stats.launches += st.launches
        self.total_stats.halo_messages += st.halo_messages
        return st

    def step(self, residual=False) -> RunStats:
        """One generation -- one trip through the reference's loop body (MDF_kernel.cu:157-187).  On the fused 2-D
        backend a launch advances whole time blocks, so single steps run the TB = 1 instantiation; prefer
        :meth:`run` with the full count."""
        return self.run(1, residual=residual)

    # -- fused ---------------------------------------------------------------------------------
    def _run_fused3d(self, g: int, residual: bool, st: RunStats) -> None:
        """One launch; the residual (every generation, or the last one for residual="final") is accumulated by the
        sweep kernel itself and all-reduced over NVLink -- no second pass over the field, no host-side reduction."""
        from .ops import _lib
        self._ensure_plan()
        before = self._lib.mcp_fused3d_launches(self._plan)
        in_kernel = residual and (self._res_in_kernel or not self.ctx.distributed)
        if in_kernel:
            final = getattr(self, "_res_final", False)
            done = 0
            hist = []
            while done < g:                                    # the device history holds 16384 generations per call
                n = g - done if final else min(g - done, 16384)
                _lib.check(self._lib.mcp_fused3d_run_residual(self._plan, n, 2 if final else 1, self._stream()),
                           "mcp_fused3d_run_residual")
                buf = (C.c_double * (2 * 16384))()
                k = _lib.check(self._lib.mcp_fused3d_residuals(self._plan, buf, 16384), "mcp_fused3d_residuals")
                hist += [(buf[2 * i], buf[2 * i + 1]) for i in range(k)]
                done += n
            _lib.check(self._lib.mcp_fused3d_status(self._plan), "fused kernel status")
            kind = self.cfg.residual if self.cfg.residual != "none" else "linf"
            if hist:
                self._last_residual = hist[-1]
            base = self.generation + (g - len(hist))
            st.residual_history = [(base + i + 1, float(np.sqrt(ss)) if kind == "l2" else mx)
                                   for i, (ss, mx) in enumerate(hist)]
        else:
            n = g - 1 if residual else g          # host residual (extra={"host_residual": True}): last step apart
            if n > 0:
                _lib.check(self._lib.mcp_fused3d_run(self._plan, n, self._stream()), "mcp_fused3d_run")
            if residual:
                prev = self._lib.mcp_fused3d_parity(self._plan)
                _lib.check(self._lib.mcp_fused3d_run(self._plan, 1, self._stream()), "mcp_fused3d_run")
                cur = self._lib.mcp_fused3d_parity(self._plan)
                self.cur = cur
                self._record_residual(st, self.bufs[prev], self.bufs[cur])
        st.launches = int(self._lib.mcp_fused3d_launches(self._plan) - before)
        self.cur = self._lib.mcp_fused3d_parity(self._plan)

    def _run_fused(self, g: int, residual: bool, st: RunStats) -> None:
        if self.ndim == 3:
            return self._run_fused3d(g, residual, st)
        from .ops import _lib
        self._ensure_plan()
        before = self._lib.mcp_fused2d_launches(self._plan)
        final = residual and getattr(self, "_res_final", False)
        mode = 2 if final else (1 if residual else 0)
        _lib.check(self._lib.mcp_fused2d_run(self._plan, g, self.time_block, mode, self._stream()), "mcp_fused2d_run")
        st.launches = int(self._lib.mcp_fused2d_launches(self._plan) - before)
        self.cur = self._lib.mcp_fused2d_parity(self._plan)
        if final and (self._res_in_kernel or not self.ctx.distributed):
            self._collect_final_residual(st)
        elif residual:
            self._collect_fused_residual(st)

    def _enqueue_result(self, st: RunStats) -> PendingResult:
        """One asynchronous 32-byte copy of the plan's result block into a pinned slot + one event."""
        from .ops import _lib
        ring = self.__dict__.get("_result_ring")
        if ring is None:
            ring = self._result_ring = _ResultRing(self._result_pool.shape[0])
        idx = ring.claim()      # settles the handle that still owns the slot, if any
        slot = self._result_pool[idx]
        _lib.check(self._lib.mcp_fused2d_result_async(self._plan, slot.data_ptr(), self._stream()),
                   "mcp_fused2d_result_async")
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.ctx.device))
        return ring.bind(idx, PendingResult(self, slot, ev, st))

    def _decode_result(self, slot: torch.Tensor, st: RunStats) -> float:
        from .ops import _lib
        status = int(slot.view(torch.int32)[0])
        if status != 0:
            raise _lib.NativeError(f"fused kernel recorded status {status}"
                                   + (" (a flag / barrier / TMA wait timed out)" if status == -4 else ""))
        ss, mx = float(slot[2]), float(slot[3])
        self._last_residual = (ss, mx)
        kind = self.cfg.residual if self.cfg.residual != "none" else "linf"
        r = float(np.sqrt(ss)) if kind == "l2" else mx
        st.residual_history = [(-1, r)]
        return r

    def _collect_final_residual(self, st: RunStats) -> None:
        """Status + global residual of the run's last generation: one asynchronous 32-byte copy into pinned memory
        and one event, no blocking library call."""
        self._enqueue_result(st).wait()

    def run_async(self, generations: Optional[int] = None) -> PendingResult:
        """Enqueue a solve of ``generations`` generations (one launch, fused residual of the last generation) and
        return without waiting; ``.wait()`` on the returned handle yields the global residual and raises if the kernel
        recorded an error.  Eight result slots: enqueueing a ninth solve first waits for the oldest.  Backends without an in-kernel residual run synchronously."""
        g = self.cfg.generations if generations is None else int(generations)
        fused2d = self.backend == "fused" and self.ndim == 2
        if fused2d:
            self._ensure_plan()
        if not fused2d or (self.ctx.distributed and not self._res_in_kernel) or g <= 0:
            st = self.run(g, residual="final", timed=False)
            return PendingResult(self, None, None, st, st.residual_history[-1][1] if st.residual_history else None)
        from .ops import _lib
        st = RunStats(generations=g, cells=self.cfg.cells)
        before = self._lib.mcp_fused2d_launches(self._plan)
        _lib.check(self._lib.mcp_fused2d_run(self._plan, g, self.time_block, 2, self._stream()), "mcp_fused2d_run")
        st.launches = int(self._lib.mcp_fused2d_launches(self._plan) - before)
        self.cur = self._lib.mcp_fused2d_parity(self._plan)
        self.generation += g
        self.total_stats.generations += g
        self.total_stats.launches += st.launches
        return self._enqueue_result(st)

    def _collect_fused_residual(self, st: RunStats) -> None:
        buf = (C.c_double * (2 * 16384))()
        n = self._lib.mcp_fused2d_residuals(self._plan, buf, 16384)
        from .ops import _lib
        _lib.check(n, "mcp_fused2d_residuals")
        status = self._lib.mcp_fused2d_status(self._plan)
        _lib.check(status, "fused kernel status")
        local = [(buf[2 * i], buf[2 * i + 1]) for i in range(n)]
        if getattr(self, "_res_in_kernel", False) or not self.ctx.distributed:
            hist = local                        # already global: all-reduced inside the kernel over NVLink
        else:
            allr = self.ctx.all_gather_object(local)
            hist = []
            for i in range(n):
                ss = 0.0
                mx = 0.0
                for r in range(self.ctx.world):     # fixed rank order -> same bits everywhere
                    ss += allr[r][i][0]
                    mx = max(mx, allr[r][i][1])
                hist.append((ss, mx))
        if hist:
            self._last_residual = hist[-1]
        kind = self.cfg.residual if self.cfg.residual != "none" else "linf"
        st.residual_history = [(-1, (np.sqrt(ss) if kind == "l2" else mx)) for ss, mx in hist]

    def check_status(self) -> None:
        """Raise if a fused kernel recorded an error (e.g. a halo-flag wait timed out)."""
        if self._plan is not None:
            from .ops import _lib
            stat = self._lib.mcp_fused3d_status if self.ndim == 3 else self._lib.mcp_fused2d_status
            _lib.check(stat(self._plan), "fused kernel status")

    def verify_protocol(self) -> dict:
        """Host-side check of the in-kernel flag protocol (2-D fused backend): after all ranks have finished their
        runs, every halo counter must equal the number of arrivals the plan has accounted for -- a lost or duplicated
        signal (or ranks that issued different run sequences) shows up here instead of as a silent hang later."""
        if self.backend != "fused" or self.ndim != 2 or self._plan is None:
            return {}
        from .ops import _lib
        torch.cuda.synchronize(self.ctx.device)
        self.ctx.barrier()                      # neighbours' last pushes have been signalled
        exp = (C.c_uint64 * 2)()
        _lib.check(self._lib.mcp_fused2d_expected_flags(self._plan, exp))
        got = self._arena.view(0, (2,), torch.int64).cpu().tolist()
        up, dn = self.block.neighbors[0]
        out = {"up": (got[0], int(exp[0])) if up is not None else None,
               "down": (got[1], int(exp[1])) if dn is not None else None}
        for k, v in out.items():
            if v is not None and v[0] != v[1]:
                raise RuntimeError(f"halo flag protocol violated on rank {self.ctx.rank} ({k}): counter {v[0]}, expected {v[1]}")
        return out

    # -- plain kernel + exchange helpers ----------------------------------------------------------
    def _launch_plain(self, src: torch.Tensor, dst: torch.Tensor) -> None:
        from .ops import _lib
        cfg, G = self.cfg, self.ghost
        s = self._stream()
        if self.ndim == 2:
            ylo, yhi, xlo, xhi = self._fix_bounds
            op = self.model.op_code | (0x100 if cfg.strict_fp64_accum and cfg.dtype == "fp32" else 0)
            _lib.check(self._lib.mcp_step2d(op, _MCP_DTYPE[cfg.dtype], src.data_ptr(), dst.data_ptr(), self.pitch,
                                            self.store_shape[0], self.block.start[0] - G, cfg.height, cfg.width,
                                            G, G + self.own_shape[0], ylo, yhi, xlo, xhi, s), "mcp_step2d")
        else:
            step3d = self._lib.mcp_step3d_cluster if cfg.extra.get("cluster3d") else self._lib.mcp_step3d
            _lib.check(step3d(_MCP_DTYPE[cfg.dtype], src.data_ptr(), dst.data_ptr(), self.pitch,
                                            self.store_shape[1] * self.pitch, self.store_shape[0],
                                            self.store_shape[1], self.block.start[0] - G, self.block.start[1] - G,
                                            cfg.depth, cfg.height, cfg.width, G, G + self.own_shape[0], G,
                                            G + self.own_shape[1], oracle.HEAT3D_COEFF, s), "mcp_step3d")

    def _residual_plain(self, a: torch.Tensor, b: torch.Tensor) -> Tuple[float, float]:
        """(sum of squares, max abs) of a-b over this rank's owned cells."""
        G = self.ghost
        if self.ctx.is_cuda and self.ndim == 2:
            from .ops import _lib
            out = self._scratch[-2:]
            _lib.check(self._lib.mcp_residual2d(_MCP_DTYPE[self.cfg.dtype], a.data_ptr(), b.data_ptr(), self.pitch,
                                                self.cfg.width, G, G + self.own_shape[0],
                                                self._scratch.data_ptr(), out.data_ptr(), self._stream()))
            v = out.cpu()
            return float(v[0]), float(v[1])
        sl = tuple(slice(G, G + n) for n in self.own_shape[:-1]) + (slice(0, self.cfg.width),)
        d = (a[sl] - b[sl]).to(torch.float64)          # difference in the field's own dtype, like the kernels
        return float((d * d).sum()), float(d.abs().max()) if d.numel() else 0.0

    def _global_residual(self, local: Tuple[float, float]) -> Tuple[float, float]:
        allr = self.ctx.all_gather_object(local)
        ss = 0.0
        mx = 0.0
        for s_, m_ in allr:
            ss += s_
            mx = max(mx, m_)
        return ss, mx

    def _record_residual(self, st: RunStats, src, dst) -> None:
        ss, mx = self._global_residual(self._residual_plain(dst, src))
        self._last_residual = (ss, mx)
        kind = self.cfg.residual if self.cfg.residual != "none" else "linf"
        st.residual_history.append((self.generation + len(st.residual_history) + 1,
                                    float(np.sqrt(ss)) if kind == "l2" else mx))

    # -- nccl baseline --------------------------------------------------------------------------
    def _run_nccl(self, g: int, residual: bool, st: RunStats) -> None:
        from .parallel import nccl_halo
        G = self.ghost
        for _ in range(g):
            src, dst = self.bufs[self.cur], self.bufs[self.cur ^ 1]
            if not self._ghosts_valid and self.ctx.distributed:
                for a in self.decomp.split_axes():
                    lo, hi = self.block.neighbors[a]
                    st.halo_messages += nccl_halo.exchange_axis(src, a, G, self.own_shape[a], lo, hi)
            self._launch_plain(src, dst)
            st.launches += 1
            if residual:
                self._record_residual(st, src, dst)
            self.cur ^= 1
            self._ghosts_valid = False

    # -- host-staged baseline (the reference's pattern) -------------------------------------------
    def _run_staged(self, g: int, residual: bool, st: RunStats) -> None:
        """Per generation: kernel, full-slab D2H, host messages, full-slab H2D -- MDF_kernel.cu:161-183 with the
        buffer swap, mirrored peers and Dirichlet mask put right."""
        G = self.ghost
        if not hasattr(self, "_host_stage"):
            self._host_stage = torch.empty(self.store_shape, dtype=self.tdtype).pin_memory()
        for _ in range(g):
            src, dst = self.bufs[self.cur], self.bufs[self.cur ^ 1]
            if not self._ghosts_valid and self.ctx.distributed:
                self._host_stage.copy_(src)                       # D2H full slab (MDF_kernel.cu:177)
                torch.cuda.synchronize(self.ctx.device)
                for a in self.decomp.split_axes():
                    lo, hi = self.block.neighbors[a]
                    if a == 0 and self.ndim == 2:
                        st.halo_messages += host_p2p.exchange_rows(self._host_stage[:, :self.cfg.width],
                                                                   G, self.own_shape[0], lo, hi, self.ctx.cpu_group(),
                                                                   mode=self.cfg.halo_message)
                    else:
                        st.halo_messages += host_p2p.exchange_axis(self._host_stage, a, G, self.own_shape[a], lo, hi,
                                                                   self.ctx.cpu_group())
                src.copy_(self._host_stage, non_blocking=True)    # H2D full slab (MDF_kernel.cu:161,171)
            self._launch_plain(src, dst)
            st.launches += 1
            if residual:
                self._record_residual(st, src, dst)
            self.cur ^= 1
            self._ghosts_valid = False

    # -- cpu --------------------------------------------------------------------------------------
    def _cpu_step(self, src: torch.Tensor, dst: torch.Tensor) -> None:
        cfg, G = self.cfg, self.ghost
        dst.copy_(src)
        if self.ndim == 2:
            ylo, yhi, xlo, xhi = self._fix_bounds
            r0 = self.block.start[0]
            lo = max(r0, ylo) - r0 + G
            hi = min(self.block.stop[0] - 1, yhi) - r0 + G + 1
            if hi <= lo or xhi < xlo:
                return
            if cfg.workload == "mdf":
                C_ = src[lo:hi, xlo:xhi + 1]
                N_, S_ = src[lo - 1:hi - 1, xlo:xhi + 1], src[lo + 1:hi + 1, xlo:xhi + 1]
                W_, E_ = src[lo:hi, xlo - 1:xhi], src[lo:hi, xlo + 1:xhi + 2]
                s = E_ + W_
                s = s + N_
                s = s + S_
                s = s + (-4.0) * C_
                if self.tdtype == torch.float32:
                    # true fma(s, 0.25, C): exact in fp64, one rounding (0.25*s is inexact for subnormal s)
                    dst[lo:hi, xlo:xhi + 1] = (s.double() * 0.25 + C_.double()).float()
                else:
                    dst[lo:hi, xlo:xhi + 1] = 0.25 * s + C_
            else:
                p = torch.nn.functional.pad(src, (1, 1))
                n = torch.zeros((hi - lo, cfg.width), dtype=torch.int32)
                for dy in (-1, 0, 1):
                    for dx in (0, 1, 2):
                        if dy == 0 and dx == 1:
                            continue
                        n += p[lo + dy:hi + dy, dx:dx + cfg.width]
                cur = src[lo:hi, :cfg.width]
                new = ((n == 3) | ((n == 2) & (cur != 0))).to(torch.int32)
                dst[lo:hi, xlo:xhi + 1] = new[:, xlo:xhi + 1]
        else:
            z0, y0 = self.block.start[0], self.block.start[1]
            zl = max(z0, 1) - z0 + G
            zh = min(self.block.stop[0] - 1, cfg.depth - 2) - z0 + G + 1
            yl = max(y0, 1) - y0 + G
            yh = min(self.block.stop[1] - 1, cfg.height - 2) - y0 + G + 1
            if zh <= zl or yh <= yl or cfg.width < 3:
                return
            xs = slice(1, cfg.width - 1)
            C_ = src[zl:zh, yl:yh, xs]
            s = src[zl:zh, yl:yh, 2:cfg.width] + src[zl:zh, yl:yh, 0:cfg.width - 2]
            s = s + src[zl:zh, yl - 1:yh - 1, xs]
            s = s + src[zl:zh, yl + 1:yh + 1, xs]
            s = s + src[zl - 1:zh - 1, yl:yh, xs]
            s = s + src[zl + 1:zh + 1, yl:yh, xs]
            d = s - 6.0 * C_
            coeff = torch.tensor(oracle.HEAT3D_COEFF, dtype=self.tdtype)
            dst[zl:zh, yl:yh, xs] = C_ + coeff * d

    def _run_cpu(self, g: int, residual: bool, st: RunStats) -> None:
        G = self.ghost
        for _ in range(g):
            src, dst = self.bufs[self.cur], self.bufs[self.cur ^ 1]
            if not self._ghosts_valid and self.ctx.distributed:
                for a in self.decomp.split_axes():
                    lo, hi = self.block.neighbors[a]
                    if a == 0 and self.ndim == 2:
                        st.halo_messages += host_p2p.exchange_rows(src, G, self.own_shape[0], lo, hi,
                                                                   self.ctx.cpu_group(), mode=self.cfg.halo_message)
                    else:
                        st.halo_messages += host_p2p.exchange_axis(src, a, G, self.own_shape[a], lo, hi,
                                                                   self.ctx.cpu_group())
            self._cpu_step(src, dst)
            if residual:
                self._record_residual(st, src, dst)
            self.cur ^= 1
            self._ghosts_valid = False

    # ------------------------------------------------------------------ convergence
    def run_to_tolerance(self, tol: Optional[float] = None, max_generations: Optional[int] = None,
                         check_every: Optional[int] = None) -> RunStats:
        """Iterate until the global residual of a generation drops to ``tol`` (extension; the reference has a
        fixed generation count only, MDF_kernel.cu:153,157,186)."""
        tol = self.cfg.tol if tol is None else tol
        if tol is None:
            raise ValueError("no tolerance given")
        max_g = (max_generations or self.cfg.max_generations
                 or (self.cfg.generations if self.cfg.generations > 1 else 100000))   # 1 = "not given"
        every = check_every or self.cfg.check_every or max(self.time_block, 1) * 25
        if self.backend == "fused":
            every = max(self.time_block, every // self.time_block * self.time_block)
        total = RunStats(cells=self.cfg.cells)
        kind = self.cfg.residual if self.cfg.residual != "none" else "linf"
        if self.backend == "fused" and self.ndim == 3 and not self.cfg.extra.get("host_tolerance_loop", False):
            self._ensure_plan()
            if not self.ctx.distributed or self._res_in_kernel:
                return self._run_to_tolerance_fused3d(tol, max_g, kind)
        if self.backend == "fused" and self.ndim == 2 and not self.cfg.extra.get("host_tolerance_loop", False):
            self._ensure_plan()
            if self.ctx.distributed and self.cfg.extra.get("residual_allreduce") == "nvls" and kind == "l2":
                raise ValueError("residual_allreduce='nvls' cannot drive an L2 stop rule: the switch's summation order "
                                 "may round differently per rank and the ranks would stop in different sweeps; use "
                                 "residual='linf' with NVLS or the default peer-to-peer exchange")
            # the single-launch path needs a stop decision that is identical on every rank: only with the in-kernel
            # all-reduce; otherwise fall through to the host loop (global residual through the control plane)
            if not self.ctx.distributed or self._res_in_kernel:
                return self._run_to_tolerance_fused(tol, max_g, kind)
        while total.generations < max_g:
            n = min(every, max_g - total.generations)
            if self.backend == "fused":
                # residual of the last generation of the last time block of this batch
                if n > self.time_block:
                    s0 = self.run(n - self.time_block, residual=False)
                    total.device_ms += s0.device_ms
                    total.launches += s0.launches
                    s1 = self.run(self.time_block, residual=True)
                else:
                    s1 = self.run(n, residual=True)
            else:
                if n > 1:
                    s0 = self.run(n - 1, residual=False)
                    total.device_ms += s0.device_ms
                    total.launches += s0.launches
                    total.halo_messages += s0.halo_messages
                s1 = self.run(1, residual=True)
            total.device_ms += s1.device_ms
            total.launches += s1.launches
            total.halo_messages += s1.halo_messages
            total.generations += n
            ss, mx = self._last_residual
            r = float(np.sqrt(ss)) if kind == "l2" else mx
            total.residual_history.append((self.generation, r))
            if r <= tol:
                total.stop_reason = "converged"
                break
        if not total.stop_reason:
            total.stop_reason = "max_generations"
        return total

    def _run_to_tolerance_fused(self, tol: float, max_g: int, kind: str) -> RunStats:
        """The kernel all-reduces the residual of every time block over NVLink and stops itself: one launch per 16384
        time blocks (the size of the device-side residual history), normally one launch in all."""
        from .ops import _lib
        self._ensure_plan()
        tb = self.time_block
        max_g = max(tb, max_g // tb * tb)
        st = RunStats(cells=self.cfg.cells)
        while True:
            done = C.c_int(0)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            before = self._lib.mcp_fused2d_launches(self._plan)
            ev0.record()
            _lib.check(self._lib.mcp_fused2d_run_to_tolerance(self._plan, max_g - st.generations, tb,
                                                              2 if kind == "l2" else 1, float(tol), C.byref(done),
                                                              self._stream()), "mcp_fused2d_run_to_tolerance")
            ev1.record()
            ev1.synchronize()
            st.device_ms += self.ctx.all_reduce_max(ev0.elapsed_time(ev1))
            st.launches += int(self._lib.mcp_fused2d_launches(self._plan) - before)
            self.cur = self._lib.mcp_fused2d_parity(self._plan)
            part = RunStats()
            self._collect_fused_residual(part)
            base = self.generation + st.generations
            st.residual_history += [(base + (i + 1) * tb, r) for i, (_, r) in enumerate(part.residual_history)]
            st.generations += int(done.value)
            st.stop_reason = {1: "converged", 2: "max_generations", 3: "slot_limit"}.get(
                int(self._lib.mcp_fused2d_stop_reason(self._plan)), "")
            # "slot_limit": the history table filled up before the tolerance was met -- not converged, go on
            if st.stop_reason != "slot_limit" or st.generations >= max_g or done.value <= 0:
                break
        self.generation += st.generations
        self.total_stats.generations += st.generations
        self.total_stats.device_ms += st.device_ms
        self.total_stats.launches += st.launches
        return st

    def _run_to_tolerance_fused3d(self, tol: float, max_g: int, kind: str) -> RunStats:
        """3-D: the sweep kernel all-reduces the residual of every generation and stops itself (one launch per 16384
        generations)."""
        from .ops import _lib
        st = RunStats(cells=self.cfg.cells)
        while True:
            done = C.c_int(0)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            before = self._lib.mcp_fused3d_launches(self._plan)
            ev0.record()
            _lib.check(self._lib.mcp_fused3d_run_to_tolerance(self._plan, max_g - st.generations,
                                                              2 if kind == "l2" else 1, float(tol), C.byref(done),
                                                              self._stream()), "mcp_fused3d_run_to_tolerance")
            ev1.record()
            ev1.synchronize()
            st.device_ms += self.ctx.all_reduce_max(ev0.elapsed_time(ev1))
            st.launches += int(self._lib.mcp_fused3d_launches(self._plan) - before)
            self.cur = self._lib.mcp_fused3d_parity(self._plan)
            _lib.check(self._lib.mcp_fused3d_status(self._plan), "fused kernel status")
            buf = (C.c_double * (2 * 16384))()
            k = _lib.check(self._lib.mcp_fused3d_residuals(self._plan, buf, 16384), "mcp_fused3d_residuals")
            base = self.generation + st.generations
            st.residual_history += [(base + i + 1, float(np.sqrt(buf[2 * i])) if kind == "l2" else buf[2 * i + 1])
                                    for i in range(k)]
            if k:
                self._last_residual = (buf[2 * (k - 1)], buf[2 * (k - 1) + 1])
            st.generations += int(done.value)
            st.stop_reason = {1: "converged", 2: "max_generations", 3: "slot_limit"}.get(
                int(self._lib.mcp_fused3d_stop_reason(self._plan)), "")
            if st.stop_reason != "slot_limit" or st.generations >= max_g or done.value <= 0:
                break
        self.generation += st.generations
        self.total_stats.generations += st.generations
        self.total_stats.device_ms += st.device_ms
        self.total_stats.launches += st.launches
        return st

    def residual(self) -> Optional[float]:
        """Last computed global residual (according to ``cfg.residual``), or None."""
        if self._last_residual is None:
            return None
        ss, mx = self._last_residual
        return float(np.sqrt(ss)) if self.cfg.residual == "l2" else mx

    # ------------------------------------------------------------------ output
    def local_owned(self) -> torch.Tensor:
        """View of this rank's owned cells in the current buffer (no ghosts, no pad)."""
        G = self.ghost
        if self.bits:
            torch.cuda.synchronize(self.ctx.device)
            words = self.bufs[self.cur][G:G + self.own_shape[0]].cpu().numpy()
            return torch.from_numpy(self._unpack_bits(words))
        sl = tuple(slice(G, G + n) for n in self.own_shape[:-1]) + (slice(0, self.cfg.width),)
        return self.bufs[self.cur][sl]

    def gather(self, as_fp32: bool = False, copy: bool = True) -> Optional[np.ndarray]:
        """Assemble the global field on rank 0 (None elsewhere).

        On GPUs this is the fused gather + pack: every rank's pack kernel strips ghosts and pad (optionally casts
        fp64 -> fp32, or unpacks bit-packed life cells) while storing its block straight into rank 0's output buffer
        over NVLink in 16-byte vectors; no intermediate copies, no NCCL.  Rank 0 then copies the dense field into a
        pinned host buffer.  ``copy=False`` returns a view of that pinned buffer (valid until the next gather) instead
        of a private array.  The reference has no gather: every rank holds (a stale copy of) the whole grid
        (MDF_kernel.cu:137-144)."""
        if self.ctx.is_cuda:
            return self._gather_p2p(as_fp32, copy)
        arr = self.local_owned().detach().cpu().contiguous().numpy()
        if not self.ctx.distributed:
            return arr.copy()
        parts = self.ctx.all_gather_object((self.block.start, arr))
        if self.ctx.rank != 0:
            return None
        out = np.zeros(self.cfg.shape, dtype=self.npdtype)
        for start, a in parts:
            sl = tuple(slice(s, s + n) for s, n in zip(start, a.shape))
            out[sl] = a
        return out

    def _gather_p2p(self, as_fp32: bool, copy: bool = True) -> Optional[np.ndarray]:
        from .ops import _lib
        from .parallel.arena import SharedBuffer
        cfg, G = self.cfg, self.ghost
        cast = as_fp32 and cfg.dtype == "fp64"
        out_t = torch.float32 if cast else self.tdtype
        out_np = np.float32 if cast else self.npdtype
        esz = 4 if (cast or cfg.dtype != "fp64") else 8
        key = (esz,)
        if getattr(self, "_gather_buf", None) is None or self._gather_key != key:
            if getattr(self, "_gather_buf", None) is not None:
                self._gather_buf.close()
            self._gather_buf = SharedBuffer(self.ctx, cfg.cells * esz, owner=0)
            self._gather_key = key
            self._gather_host = None
            self._gather_ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True),
                               torch.cuda.Event(enable_timing=True))
        shape3 = cfg.shape if self.ndim == 3 else (1,) + tuple(cfg.shape)
        if self.ndim == 3:
            z_lo, nz, y_lo, ny = G, self.own_shape[0], G, self.own_shape[1]
            plane_pitch = self.store_shape[1] * self.pitch
            dz0, dy0 = self.block.start[0], self.block.start[1]
        else:
            z_lo, nz, y_lo, ny = 0, 1, G, self.own_shape[0]
            plane_pitch = self.store_shape[0] * self.pitch
            dz0, dy0 = 0, self.block.start[0]
        self.ctx.barrier()          # rank 0's previous reader is done with the buffer
        ev0, ev1, ev2 = self._gather_ev
        ev0.record()
        if self.bits:
            _lib.check(self._lib.mcp_pack_rows(3, 3, self.bufs[self.cur].data_ptr(), self.pitch, cfg.width, G,
                                               G + self.own_shape[0], self._gather_buf.ptr, self.block.start[0],
                                               self._stream()), "mcp_pack_rows")
        else:
            _lib.check(self._lib.mcp_pack_block(_MCP_DTYPE[cfg.dtype], 1 if cast else 0, self.bufs[self.cur].data_ptr(),
                                                self.pitch, plane_pitch, z_lo, nz, y_lo, ny, cfg.width,
                                                self._gather_buf.ptr, shape3[1], dz0, dy0, self._stream()),
                       "mcp_pack_block")
        ev1.record()
        torch.cuda.synchronize(self.ctx.device)
        pack_ms = self.ctx.all_reduce_max(ev0.elapsed_time(ev1))     # also: every rank's stores have landed
        own_bytes = int(np.prod(self.own_shape[:-1])) * cfg.width * esz
        self.last_gather_stats = {"pack_ms": round(pack_ms, 4), "bytes_per_rank": own_bytes,
                                  "bytes_into_rank0": cfg.cells * esz,
                                  "rank0_ingress_gbs": round(cfg.cells * esz / (pack_ms * 1e-3) / 1e9, 2) if pack_ms > 0 else None,
                                  "nvlink_peak_gbs": 770.0}
        if self.ctx.rank != 0:
            return None
        dev = self._gather_buf.view(cfg.shape, out_t)
        if self._gather_host is None:
            self._gather_host = torch.empty(cfg.shape, dtype=out_t).pin_memory()
        self._gather_host.copy_(dev, non_blocking=True)
        ev2.record()
        ev2.synchronize()
        self.last_gather_stats["d2h_ms"] = round(ev1.elapsed_time(ev2), 4)
        arr = self._gather_host.numpy()
        return arr.astype(out_np, copy=True) if copy else arr

    def checksum(self) -> int:
        """64-bit position-dependent checksum of the GLOBAL field (same value on every rank), computed on the devices:
        per-rank sums over owned cells (``mcp_checksum2d``) added mod 2^64.  Equal checksums <=> equal fields (up to
        hash collisions), for any rank count and backend -- the benchmark-size correctness check."""
        if self.ndim != 2 or self.bits:
            raise ValueError("checksum() covers 2-D fp32/fp64/int32 fields")
        G = self.ghost
        if self.ctx.is_cuda:
            from .ops import _lib
            out = torch.zeros(1, dtype=torch.int64, device=self.ctx.device)
            _lib.check(self._lib.mcp_checksum2d(_MCP_DTYPE[self.cfg.dtype], self.bufs[self.cur].data_ptr(), self.pitch,
                                                self.cfg.width, G, G + self.own_shape[0], self.block.start[0],
                                                out.data_ptr(), self._stream()), "mcp_checksum2d")
            local = int(out.cpu()[0]) & 0xFFFFFFFFFFFFFFFF
        else:
            local = oracle.checksum_rows(self.bufs[self.cur][G:G + self.own_shape[0], :self.cfg.width].numpy(),
                                         self.block.start[0], self.cfg.width)
        return sum(self.ctx.all_gather_object(local)) & 0xFFFFFFFFFFFFFFFF

    def gather_ascii(self) -> Optional[bytes]:
        """Life grids: the bytes the reference's ``print_array`` would print (kernel.cu:115-129), produced on the
        GPUs -- every rank packs its rows to '0' / ' ' / newline bytes straight into rank 0's buffer."""
        if self.cfg.workload != "life":
            raise ValueError("print_array takes int grids only (it cannot print the float field, SURVEY D16)")
        if not self.ctx.is_cuda or self.bits:
            from . import io as mio
            f = self.gather()
            return mio.format_array(f) if f is not None else None
        from .ops import _lib
        from .parallel.arena import SharedBuffer
        h, w, G = self.cfg.height, self.cfg.width, self.ghost
        buf = SharedBuffer(self.ctx, h * (w + 1), owner=0)
        try:
            _lib.check(self._lib.mcp_pack_rows(2, 2, self.bufs[self.cur].data_ptr(), self.pitch, w, G,
                                               G + self.own_shape[0], buf.ptr, self.block.start[0], self._stream()),
                       "mcp_pack_rows")
            torch.cuda.synchronize(self.ctx.device)
            self.ctx.barrier()
            out = None
            if self.ctx.rank == 0:
                out = b"\n" + buf.view((h * (w + 1),), torch.uint8).cpu().numpy().tobytes() + b"\n"
        finally:
            buf.close()
        return out

    # ------------------------------------------------------------------ teardown
    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        if self.ctx.is_cuda:
            torch.cuda.synchronize(self.ctx.device)
        if self._plan is not None:
            (self._lib.mcp_fused3d_destroy if self.ndim == 3 else self._lib.mcp_fused2d_destroy)(self._plan)
            self._plan = None
        self.bufs = []
        if getattr(self, "_gather_buf", None) is not None:
            self._gather_buf.close()
            self._gather_buf = None
        if self._arena is not None:
            self._arena.close()
            self._arena = None
        if self._own_ctx:
            self.ctx.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


class _SoloCtx:
    """Arena context for allocations that are not shared with peers."""

    def __init__(self, ctx):
        self.rank, self.world, self.device, self.distributed = ctx.rank, 1, ctx.device, False

    def barrier(self):
        pass

    def all_gather_object(self, o):
        return [o]
