"""Executable model of the fused 2-D kernel's synchronisation protocol (host-side protocol checker).

The kernel (ops/csrc/stencil2d_fused.cu) orders its memory traffic with nothing but flags: per-tile completion words
inside a GPU, per-strip sequence flags or one arrival counter per neighbour between GPUs, two field buffers per rank.
This module restates those rules on a small abstract machine -- ranks x chunks x strips tiles, warps that draw tickets in
order and advance one phase at a time (wait, read, write, push, signal, publish) -- and lets an adversarial random
scheduler interleave the phases of all warps of all ranks.  Every cell region carries the number of the sweep that last
wrote it, so a read of data that is too old (a missed read-after-write dependency) or too new (a write-after-read
hazard: the region was overwritten before this reader got to it) is detected the moment it happens.

It is a model, not the kernel: it checks the protocol's rules, and it documents which rule breaks which guarantee --
``flags="coarse"`` with ``sync="dataflow"`` is the combination the injected-delay stress test caught on the GPU (one
arrival counter cannot tell a barrier-free neighbour's sweeps apart); the model finds the same violation in
milliseconds.  ``launches`` > 1 models back-to-back kernels with no host barrier between the ranks; with ``upload``
the host overwrites the whole slab between launches (Solver.load_local / commit_input), which is only safe because a
launch drains its incoming halo pushes before it ends (``drain=False`` shows the race; without uploads the sequence flags
alone are enough).  Warps draw their next ticket when they start a tile (``lookahead``) and read that tile's input early
when a non-blocking probe of its nine completion words succeeds (``prefetch``), as the kernel does; ``prefetch_ghost``
is the variant the kernel avoids (prefetching a tile whose condition includes the neighbour GPU's flags).  The reference has no protocol to check: its only synchronisation is cudaDeviceSynchronize and blocking
one-element MPI calls (MDF_kernel.cu:166-183), and its D2H -> send ordering holds by accident.
"""
from __future__ import annotations

import random
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple


class ProtocolViolation(AssertionError):
    pass


@dataclass
class _Warp:
    rank: int
    ticket: Optional[Tuple[int, int, int]] = None   # (sweep, chunk, strip)
    phase: int = 0                                   # 0 wait, 1 read, 2 write, 3 push, 4 signal, 5 publish
    nxt: Optional[Tuple[int, int, int]] = None       # look-ahead ticket, drawn when the current tile starts
    looked: bool = False                             # the look-ahead draw of the current tile has been attempted
    pre: bool = False                                # the look-ahead tile's input has been read already (prefetch)
    early_step: int = 0                              # early_up_signal: 0 = push pending, 1 = signal pending
    early_done: bool = False                         # early_up_signal: this tile has already pushed + signalled upwards


@dataclass
class Model:
    ranks: int = 2
    chunks: int = 3
    strips: int = 4
    sweeps: int = 5
    warps: int = 3
    flags: str = "strip"          # "strip": per-strip sequence flags; "coarse": one arrival counter per neighbour
    sync: str = "dataflow"        # "dataflow": per-tile completion words; "barrier": grid barrier between sweeps
    launches: int = 1             # kernel launches per rank, `sweeps` sweeps each, NO host barrier between the ranks
    upload: bool = True           # between launches the host overwrites the whole slab (ghost rows included)
    drain: bool = True            # a launch ends only when every halo push addressed to this rank has landed
    lookahead: bool = True        # a warp draws its next ticket when it starts a tile (grab() behind the current tile)
    prefetch: bool = True         # ... and reads that tile's input early if a non-blocking probe finds it settled
    prefetch_ghost: bool = False  # UNSOUND variant: prefetch also tiles that read ghost rows, probing tile_done only
    early_up_signal: bool = False  # PROPOSED variant (not in the kernel): a top boundary tile pushes and signals its
    #                                first rows upwards as soon as they exist, before it computes the rest of the tile
    seed: int = 0
    stall_rate: float = 0.02      # probability per scheduler round that the adversary freezes a warp or a rank
    max_stall: int = 400          # longest freeze, in scheduler rounds
    log: List[str] = field(default_factory=list)

    def __post_init__(self):
        R, C, S = self.ranks, self.chunks, self.strips
        self.rng = random.Random(self.seed)
        # version[r][buf][c][s]: sweep count of the data in that tile region (buffer 0 starts as generation 0)
        self.version = [[[[0 if b == 0 else -1 for _ in range(S)] for _ in range(C)] for b in range(2)] for _ in range(R)]
        # ghost[r][buf][dir][s]: version of the neighbour's boundary rows as seen in rank r's ghost rows (dir 0 = up)
        self.ghost = [[[[0 if b == 0 else -1 for _ in range(S)] for _ in range(2)] for b in range(2)] for _ in range(R)]
        self.tile_done = [[[0] * S for _ in range(C)] for _ in range(R)]          # last sweep + 1 each tile finished
        self.strip_flag = [[[0] * S for _ in range(2)] for _ in range(R)]         # [r][dir][s] = sweeps done by nbr
        self.counter = [[0, 0] for _ in range(R)]                                 # arrivals from up / down
        self.next_ticket = [0] * R
        self.finished = [[0] * (self.sweeps * self.launches) for _ in range(R)]   # tiles finished per global sweep
        self.launched = [0] * R                                                   # launch each rank is in
        self.pool = [_Warp(r) for r in range(R) for _ in range(self.warps)]
        self.order = self._tile_order()

    # outside-in chunk order, all strips of a chunk together (decode_tile in the kernel)
    def _tile_order(self) -> List[Tuple[int, int]]:
        C, S = self.chunks, self.strips
        out = []
        for ci in range(C):
            c = C - 1 - (ci >> 1) if (ci & 1) else (ci >> 1)
            out += [(c, s) for s in range(S)]
        return out

    def _has(self, r: int, d: int) -> bool:
        return (r > 0) if d == 0 else (r < self.ranks - 1)

    # ------------------------------------------------------------------ one phase of one warp
    _EPOCH = 10_000               # uploaded data of launch L carries version L*_EPOCH + (global sweep)

    def _tag(self, launch: int) -> int:
        return launch * self._EPOCH if self.upload else 0

    def _drained(self, r: int, g: int) -> bool:
        """All pushes the neighbours address to rank r during their first g sweeps have been signalled."""
        for d in (0, 1):
            if not self._has(r, d):
                continue
            if self.flags == "strip":
                if any(f < g for f in self.strip_flag[r][d]):
                    return False
            elif self.counter[r][d] < g * self.strips:
                return False
        return True

    def _draw_ticket(self, r: int, transition: bool = True) -> Optional[Tuple[int, int, int]]:
        """Next ticket of rank r, or None if none can be handed out now.  A launch boundary is crossed only with
        ``transition`` (the look-ahead draw never crosses one: the kernel's ticket range ends with the launch)."""
        n_tiles = self.chunks * self.strips
        t = self.next_ticket[r]
        if t >= self.launches * self.sweeps * n_tiles:
            return None
        g = t // n_tiles                                  # global sweep: seq_base + sw in the kernel
        launch = g // self.sweeps
        if launch > self.launched[r]:
            if not transition:
                return None
            # the previous kernel of this rank has to end first: all its tiles, then the drain; then the host's copy
            if any(self.finished[r][g2] < n_tiles for g2 in range(g)):
                return None
            if self.drain and not self._drained(r, g):
                return None
            if self.upload:
                b, v = g & 1, self._tag(launch) + g
                for c in range(self.chunks):
                    for s in range(self.strips):
                        self.version[r][b][c][s] = v
                for d in (0, 1):
                    for s in range(self.strips):
                        self.ghost[r][b][d][s] = v
            self.launched[r] = launch
        if self.sync == "barrier" and g % self.sweeps > 0 and self.finished[r][g - 1] < n_tiles:
            return None                                   # grid barrier: a sweep starts when the previous one is complete
        self.next_ticket[r] = t + 1
        c, s = self.order[t % n_tiles]
        return (g, c, s)

    def _reads_ghost(self, r: int, ticket) -> bool:
        _, c, _ = ticket
        return (c == 0 and self._has(r, 0)) or (c == self.chunks - 1 and self._has(r, 1))

    def _deps_ready(self, r: int, ticket, local_only: bool = False) -> bool:
        k, c, s = ticket
        C, S = self.chunks, self.strips
        if self.sync == "dataflow" and k % self.sweeps > 0:
            for c2 in (c - 1, c, c + 1):
                for s2 in (s - 1, s, s + 1):
                    if 0 <= c2 < C and 0 <= s2 < S and self.tile_done[r][c2][s2] < k:
                        return False
        if local_only:
            return True
        for d, edge in ((0, c == 0), (1, c == C - 1)):
            if edge and self._has(r, d) and k > 0:
                if self.flags == "strip":
                    if any(0 <= s2 < S and self.strip_flag[r][d][s2] < k for s2 in (s - 1, s, s + 1)):
                        return False
                elif self.counter[r][d] < k * S:
                    return False
        return True

    def _read(self, r: int, ticket) -> None:
        k, c, s = ticket
        C, S = self.chunks, self.strips
        b = k & 1
        for c2 in (c - 1, c, c + 1):
            for s2 in (s - 1, s, s + 1):
                if not 0 <= s2 < S:
                    continue
                if 0 <= c2 < C:
                    got, what = self.version[r][b][c2][s2], f"tile ({c2},{s2})"
                elif self._has(r, 0 if c2 < 0 else 1):
                    got, what = self.ghost[r][b][0 if c2 < 0 else 1][s2], f"ghost {'up' if c2 < 0 else 'down'} strip {s2}"
                else:
                    continue                                  # physical boundary: fixed cells
                want = self._tag(k // self.sweeps) + k
                if got != want:
                    kind = "stale data (missed read-after-write)" if got < want else "overwritten data (write-after-read)"
                    raise ProtocolViolation(f"rank {r} sweep {k} tile ({c},{s}) reads {what} at version {got}, "
                                            f"expected {want}: {kind}; flags={self.flags} sync={self.sync} "
                                            f"drain={self.drain} seed={self.seed}")

    def _step(self, w: _Warp) -> bool:
        """Advance warp ``w`` by one phase if it can; returns whether anything happened."""
        r = w.rank
        if w.ticket is None:
            if w.nxt is not None:                              # the look-ahead ticket becomes the current tile
                w.ticket, w.nxt, w.phase, w.looked = w.nxt, None, (2 if w.pre else 0), False
                w.pre = False
                return True
            t = self._draw_ticket(r)
            if t is None:
                return False
            w.ticket, w.phase, w.looked = t, 0, False
            return True
        k, c, s = w.ticket
        C = self.chunks
        tag = self._tag(k // self.sweeps)
        if self.lookahead and not w.looked:                    # grab() of the next ticket, before anything else
            w.looked = True
            w.nxt = self._draw_ticket(r, transition=False)
            return True
        if (self.prefetch and self.sync == "dataflow" and w.nxt is not None and not w.pre and 2 <= w.phase <= 4
                and self.rng.random() < 0.5 and (self.prefetch_ghost or not self._reads_ghost(r, w.nxt))
                and self._deps_ready(r, w.nxt, local_only=True)):
            # tail of the row loop: the probe found the next tile's nine completion words set -> its loads are issued
            self._read(r, w.nxt)
            w.pre = True
            return True
        if w.phase == 0:
            if not self._deps_ready(r, w.ticket):
                return False
            w.phase = 1
        elif w.phase == 1:
            self._read(r, w.ticket)
            w.phase = 2
        elif w.phase == 2:
            early = self.early_up_signal and c == 0 and self._has(r, 0) and not w.early_done
            if early:
                # the rows streamed first are the upper ghost rows (all read by now) and the tile's top rows: push them
                # and raise the upper neighbour's flag, then go on with the rest of the tile (two scheduler steps)
                if w.early_step == 0:
                    self.ghost[r - 1][(k + 1) & 1][1][s] = tag + k + 1
                    w.early_step = 1
                else:
                    self.strip_flag[r - 1][1][s] = k + 1
                    self.counter[r - 1][1] += 1
                    w.early_step, w.early_done = 0, True
                return True
            self.version[r][(k + 1) & 1][c][s] = tag + k + 1
            w.phase = 3
        elif w.phase == 3:                                     # halo push into the neighbours' ghost rows
            if c == 0 and self._has(r, 0) and not w.early_done:
                self.ghost[r - 1][(k + 1) & 1][1][s] = tag + k + 1
            if c == C - 1 and self._has(r, 1):
                self.ghost[r + 1][(k + 1) & 1][0][s] = tag + k + 1
            w.phase = 4
        elif w.phase == 4:                                     # signals behind the fence
            if c == 0 and self._has(r, 0) and not w.early_done:
                self.strip_flag[r - 1][1][s] = k + 1
                self.counter[r - 1][1] += 1
            if c == C - 1 and self._has(r, 1):
                self.strip_flag[r + 1][0][s] = k + 1
                self.counter[r + 1][0] += 1
            w.early_done = False
            w.phase = 5
        else:
            self.tile_done[r][c][s] = k + 1
            self.finished[r][k] += 1
            w.ticket = None
        return True

    def run(self, max_steps: int = 1_000_000) -> int:
        """Run to completion under a random schedule; returns the number of steps.  Raises ProtocolViolation on a
        hazard and RuntimeError on a deadlock."""
        steps = 0
        total = self.ranks * self.launches * self.sweeps * self.chunks * self.strips
        stalled: Dict[int, int] = {}                       # warp index -> scheduler rounds it stays frozen
        while sum(sum(f) for f in self.finished) < total:
            # adversary: now and then freeze a warp (or every warp of a rank) for many rounds, the way a delayed tile
            # or a descheduled CTA would be -- short stalls never separate two sweeps of one neighbour
            if self.rng.random() < self.stall_rate:
                length = self.rng.randrange(20, self.max_stall)
                if self.rng.random() < 0.25:
                    r = self.rng.randrange(self.ranks)
                    for i, w in enumerate(self.pool):
                        if w.rank == r:
                            stalled[i] = length
                else:
                    stalled[self.rng.randrange(len(self.pool))] = length
            order = list(range(len(self.pool)))
            self.rng.shuffle(order)
            progressed = skipped = False
            for i in order:
                if stalled.get(i, 0) > 0:
                    skipped = True
                    continue
                if self._step(self.pool[i]):
                    progressed = True
                    steps += 1
                    if self.rng.random() < 0.7:
                        break                              # mostly one phase at a time: fine-grained interleaving
            for i in list(stalled):
                stalled[i] -= 1
                if stalled[i] <= 0:
                    del stalled[i]
            if not progressed:
                if skipped:
                    stalled.clear()                        # everything runnable is frozen: thaw and try again
                    continue
                raise RuntimeError(f"deadlock: flags={self.flags} sync={self.sync} seed={self.seed}")
            if steps > max_steps:
                raise RuntimeError("model did not finish")
        return steps


def explore(seeds: int = 200, **kw) -> Dict[str, int]:
    """Run ``seeds`` random schedules; returns {"ok": n, "violations": m} (first violation message under "first")."""
    out: Dict[str, object] = {"ok": 0, "violations": 0}
    for seed in range(seeds):
        try:
            Model(seed=seed, **kw).run()
            out["ok"] += 1
        except ProtocolViolation as e:
            out["violations"] += 1
            out.setdefault("first", str(e))
    return out


if __name__ == "__main__":                                 # python -m mpi_cuda_process_b200.utils.protocol_model
    import argparse
    import json

    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--seeds", type=int, default=200)
    ap.add_argument("--ranks", type=int, default=3)
    ap.add_argument("--chunks", type=int, default=3)
    ap.add_argument("--strips", type=int, default=4)
    ap.add_argument("--sweeps", type=int, default=5)
    ap.add_argument("--warps", type=int, default=3)
    ap.add_argument("--launches", type=int, default=1)
    ap.add_argument("--flags", choices=["strip", "coarse"], default="strip")
    ap.add_argument("--sync", choices=["dataflow", "barrier"], default="dataflow")
    ap.add_argument("--no-drain", action="store_true")
    ap.add_argument("--no-upload", action="store_true")
    ap.add_argument("--no-prefetch", action="store_true")
    ap.add_argument("--prefetch-ghost", action="store_true", help="the unsound variant the kernel avoids")
    a = ap.parse_args()
    print(json.dumps(explore(a.seeds, ranks=a.ranks, chunks=a.chunks, strips=a.strips, sweeps=a.sweeps, warps=a.warps,
                             launches=a.launches, flags=a.flags, sync=a.sync, drain=not a.no_drain,
                             upload=not a.no_upload, prefetch=not a.no_prefetch, prefetch_ghost=a.prefetch_ghost)))
