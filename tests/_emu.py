"""Several fused-2D "ranks" on ONE GPU, wired to each other exactly as ranks on different GPUs are (peer buffers, halo
counters, per-strip flags, residual inboxes) -- only the pointer values differ.  They are advanced together by
``mcp_fused2d_run_group`` (one cooperative launch, CTA b works for rank b % n)."""
import ctypes as C

import numpy as np
import torch

from mpi_cuda_process_b200.ops import _lib as lib_mod

MAX_STRIPS = 2048
RES_RING, MAX_RANKS = 64, 16


class EmulatedRanks:
    def __init__(self, field: np.ndarray, n: int, tb: int, strip_flags: bool = True, residual: bool = True,
                 chunk_rows: int = 0, options=None):
        self.lib = lib_mod.cuda_lib()
        h, w = field.shape
        self.h, self.w, self.n, self.tb, self.G = h, w, n, tb, tb
        self.pitch = (w + 31) // 32 * 32
        self.starts = [h * r // n for r in range(n)]
        self.rows = [h * (r + 1) // n - h * r // n for r in range(n)]
        G = self.G
        self.bufs, self.flags, self.sflags, self.inbox, self.resflag = [], [], [], [], []
        for r in range(n):
            loc = np.zeros((self.rows[r] + 2 * G, self.pitch), dtype=np.float32)
            lo, hi = self.starts[r] - G, self.starts[r] + self.rows[r] + G
            cs, ce = max(lo, 0), min(hi, h)
            loc[cs - lo:ce - lo, :w] = field[cs:ce]
            t = torch.from_numpy(loc).cuda()
            self.bufs.append([t.clone(), t.clone()])
            self.flags.append(torch.zeros(8, dtype=torch.int64, device="cuda"))
            self.sflags.append(torch.zeros(2 * MAX_STRIPS, dtype=torch.int64, device="cuda"))
            self.inbox.append(torch.zeros(RES_RING * MAX_RANKS * 2, dtype=torch.float64, device="cuda"))
            self.resflag.append(torch.zeros(8, dtype=torch.int64, device="cuda"))
        self.plans = []
        for r in range(n):
            d = lib_mod.FusedDesc()
            d.op, d.dtype, d.h, d.w = 0, 0, h, w
            d.row_start, d.rows_own, d.ghost, d.pitch = self.starts[r], self.rows[r], G, self.pitch
            up, dn = (r - 1 if r > 0 else None), (r + 1 if r < n - 1 else None)
            for i in range(2):
                d.buf[i] = self.bufs[r][i].data_ptr()
                d.up_buf[i] = self.bufs[up][i].data_ptr() if up is not None else None
                d.dn_buf[i] = self.bufs[dn][i].data_ptr() if dn is not None else None
            d.up_rows_own = self.rows[up] if up is not None else 0
            d.flags = self.flags[r].data_ptr()
            d.up_flag = self.flags[up].data_ptr() + 8 if up is not None else None
            d.dn_flag = self.flags[dn].data_ptr() if dn is not None else None
            if strip_flags:
                d.strip_flags = self.sflags[r].data_ptr()
                d.up_strip_flag = self.sflags[up].data_ptr() + MAX_STRIPS * 8 if up is not None else None
                d.dn_strip_flag = self.sflags[dn].data_ptr() if dn is not None else None
            d.ylo, d.yhi, d.xlo, d.xhi = 1, h - 2, 1, w - 2
            d.chunk_rows = chunk_rows
            d.world, d.rank = n, r
            if residual and n > 1:
                d.res_inbox = self.inbox[r].data_ptr()
                d.res_flag = self.resflag[r].data_ptr()
                for q in range(n):
                    d.peer_res_inbox[q] = self.inbox[q].data_ptr()
                    d.peer_res_flag[q] = self.resflag[q].data_ptr()
            p = C.c_void_p()
            lib_mod.check(self.lib.mcp_fused2d_create(C.byref(d), C.byref(p)), "create")
            for opt, val in (options or {}).items():
                lib_mod.check(self.lib.mcp_fused2d_set_option(p, opt, val), "set_option")
            self.plans.append(p)
        self._arr = (C.c_void_p * n)(*[p.value for p in self.plans])

    def run(self, generations: int, residual_mode: int = 0) -> None:
        lib_mod.check(self.lib.mcp_fused2d_run_group(self._arr, self.n, generations, self.tb, residual_mode, None),
                      "run_group")
        torch.cuda.synchronize()

    def status(self):
        return [self.lib.mcp_fused2d_status(p) for p in self.plans]

    def field(self) -> np.ndarray:
        out = np.zeros((self.h, self.w), dtype=np.float32)
        for r in range(self.n):
            cur = self.lib.mcp_fused2d_parity(self.plans[r])
            out[self.starts[r]:self.starts[r] + self.rows[r]] = \
                self.bufs[r][cur][self.G:self.G + self.rows[r], :self.w].cpu().numpy()
        return out

    def result(self, r: int):
        host = torch.zeros(4, dtype=torch.float64).pin_memory()
        lib_mod.check(self.lib.mcp_fused2d_result_async(self.plans[r], host.data_ptr(), None))
        torch.cuda.synchronize()
        w = host.view(torch.int32)
        return {"status": int(w[0]), "sweeps_done": int(w[2]), "sumsq": float(host[2]), "max": float(host[3])}

    def close(self):
        for p in self.plans:
            self.lib.mcp_fused2d_destroy(p)
        self.plans = []
