#!/usr/bin/env python
"""Device times of the auxiliary kernels (init, boundary re-imposition, residual, checksum, gather/pack, bit unpack)
on one GPU: CUDA events, L2 flushed between iterations, median of 10, reported as GB/s of the bytes the kernel must
move against the measured HBM copy bandwidth (MEASURED_PEAKS.json).  One JSON line per kernel (also appended to
``--log``).  Reference counterparts: create_universe (MDF_kernel.cu:88-99), the "else -> 100" branches (:35,43,59,67),
print_array (kernel.cu:115-129); the reference times nothing.

    python benchmarks/microbench.py [--n 16384] [--log gpurun_out/microbench.jsonl]
"""
import argparse
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_cuda_process_b200.ops import _lib  # noqa: E402
from mpi_cuda_process_b200.utils.log import JsonlLogger  # noqa: E402
from mpi_cuda_process_b200.utils.metrics import measured_peaks  # noqa: E402
from mpi_cuda_process_b200.utils.timing import L2Flusher, time_cuda  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=16384)
ap.add_argument("--log", default=None)
a = ap.parse_args()
lib = _lib.cuda_lib()
dev = torch.device("cuda", 0)
n, pitch, G = a.n, a.n, 8
rows = n // 8 + 2 * G                         # one rank's slab of an 8-way decomposition
f32 = [torch.zeros((rows, pitch), dtype=torch.float32, device=dev) for _ in range(2)]
f64 = torch.zeros((rows, pitch), dtype=torch.float64, device=dev)
i32 = torch.zeros((rows, pitch), dtype=torch.int32, device=dev)
words = torch.randint(0, 2**31 - 1, (rows, (n // 32 + 31) // 32 * 32), dtype=torch.int32, device=dev)
dense = torch.empty((rows - 2 * G) * n, dtype=torch.float64, device=dev)
scratch = torch.empty(max(1, lib.mcp_residual2d_scratch_bytes() // 8) + 2, dtype=torch.float64, device=dev)
out64 = torch.zeros(1, dtype=torch.int64, device=dev)
peak = measured_peaks()["hbm_gbs"]
flush = L2Flusher(dev)
log = JsonlLogger(a.log)
own = rows - 2 * G
cases = [
    ("mcp_init_ring2d f32", rows * pitch * 4,
     lambda: lib.mcp_init_ring2d(0, f32[0].data_ptr(), pitch, rows, 100, n, n, 100.0, 0.0, None)),
    ("mcp_init_shape2d hash f32", rows * pitch * 4,
     lambda: lib.mcp_init_shape2d(0, 1, f32[1].data_ptr(), pitch, rows, 100, n, n, 100.0, 0.0, 7, None)),
    ("mcp_init_life_hash i32", rows * pitch * 4,
     lambda: lib.mcp_init_life_hash(i32.data_ptr(), pitch, rows, 100, n, n, 0.15, 7, None)),
    ("mcp_apply_ring2d f32", 2 * n * 4 + 2 * own * 4,
     lambda: lib.mcp_apply_ring2d(0, f32[0].data_ptr(), pitch, rows, 0, n, n, 100.0, None)),
    ("mcp_residual2d f32", 2 * own * n * 4,
     lambda: lib.mcp_residual2d(0, f32[0].data_ptr(), f32[1].data_ptr(), pitch, n, G, G + own, scratch.data_ptr(),
                                scratch[-2:].data_ptr(), None)),
    ("mcp_checksum2d f32", own * n * 4,
     lambda: lib.mcp_checksum2d(0, f32[0].data_ptr(), pitch, n, G, G + own, 0, out64.data_ptr(), None)),
    ("mcp_pack_block f32 (gather, local destination)", 2 * own * n * 4,
     lambda: lib.mcp_pack_block(0, 0, f32[0].data_ptr(), pitch, rows * pitch, 0, 1, G, own, n, dense.data_ptr(), n, 0, 0, None)),
    ("mcp_pack_block f64->f32 (gather + cast)", own * n * 12,
     lambda: lib.mcp_pack_block(1, 1, f64.data_ptr(), pitch, rows * pitch, 0, 1, G, own, n, dense.data_ptr(), n, 0, 0, None)),
    ("mcp_pack_rows life -> print_array bytes", own * n * 4 + own * (n + 1),
     lambda: lib.mcp_pack_rows(2, 2, i32.data_ptr(), pitch, n, G, G + own, dense.data_ptr(), 0, None)),
    ("mcp_pack_rows bit-packed life -> int32 cells", own * n // 8 + own * n * 4,
     lambda: lib.mcp_pack_rows(3, 3, words.data_ptr(), words.shape[1], n, G, G + own, dense.data_ptr(), 0, None)),
]
for name, nbytes, fn in cases:
    def call(fn=fn):
        _lib.check(fn(), name)
    ms = statistics.median(time_cuda(call, iters=10, warmup=3, flush=flush))
    gbs = nbytes / (ms * 1e-3) / 1e9
    log.log(kernel=name, ms=round(ms, 4), bytes=int(nbytes), gbs=round(gbs, 1),
            fraction_of_measured_hbm=round(gbs / peak, 3), grid=[own, n], l2="flushed between iterations")
    print(f"{name:52s} {ms:8.4f} ms {gbs:8.1f} GB/s  {gbs / peak:5.2f} of measured HBM copy bandwidth")
log.close()
