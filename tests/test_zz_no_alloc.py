"""The allocation-free residual entry point, called through the C ABI exactly as a foreign host would.

Some users of this library are not allowed to allocate device memory themselves: the GPU's memory may be
fully claimed by another allocator, the process may run under a fixed memory budget, or the call may sit in
a loop where the device-wide stall of a cudaMalloc/cudaFree is unacceptable.  ``mcp_residual2d_noalloc``
serves them with a scratch array that is part of the library image.  This test checks that it returns what the
regular entry point (``mcp_residual2d`` with caller-provided scratch) returns -- on two calls in a row, because
the built-in array is shared and keeps the previous call's contents -- and that it refuses a row range larger
than the array can hold.

The file name sorts last on purpose: if this test fails under ``-x`` every other test has already run.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(lib_mod, lib, dtype_id, a, b, pitch, w, lo, hi, scratch):
    ref = torch.zeros(2, dtype=torch.float64, device="cuda")
    got = torch.zeros(2, dtype=torch.float64, device="cuda")
    lib_mod.check(lib.mcp_residual2d(dtype_id, a.data_ptr(), b.data_ptr(), pitch, w, lo, hi, scratch.data_ptr(),
                                     ref.data_ptr(), None), "mcp_residual2d")
    lib_mod.check(lib.mcp_residual2d_noalloc(dtype_id, a.data_ptr(), b.data_ptr(), pitch, w, lo, hi,
                                             got.data_ptr(), None), "mcp_residual2d_noalloc")
    torch.cuda.synchronize()
    return ref.cpu().numpy(), got.cpu().numpy()


def test_noalloc_matches_regular_entry_point_twice():
    from mpi_cuda_process_b200.ops import _lib as lib_mod
    lib = lib_mod.cuda_lib()
    rng = np.random.default_rng(11)
    scratch = torch.empty(lib.mcp_residual2d_scratch_bytes() // 8, dtype=torch.float64, device="cuda")

    # call 1: fp32, many rows
    h, w, pitch = 700, 333, 352
    a = torch.from_numpy(rng.normal(size=(h, pitch)).astype(np.float32)).cuda()
    b = torch.from_numpy(rng.normal(size=(h, pitch)).astype(np.float32)).cuda()
    ref1, got1 = _both(lib_mod, lib, 0, a, b, pitch, w, 5, 690, scratch)
    # call 2, straight after: fp64, fewer rows and different data -- stale entries of call 1 must not leak in
    h2, w2, pitch2 = 90, 61, 64
    a2 = torch.from_numpy(rng.normal(size=(h2, pitch2))).cuda()
    b2 = torch.from_numpy(rng.normal(size=(h2, pitch2))).cuda()
    ref2, got2 = _both(lib_mod, lib, 1, a2, b2, pitch2, w2, 0, 90, scratch)

    for ref, got in ((ref1, got1), (ref2, got2)):
        assert got[1] == ref[1]                                   # max |a-b| is order independent: exact
        assert abs(got[0] - ref[0]) <= 1e-12 * abs(ref[0])        # fp64 sums in a different fixed order
    # and both agree with a plain PyTorch fp64 evaluation of the same quantity
    d = (a[5:690, :w] - b[5:690, :w]).double()      # the library forms a-b in the field's own precision
    assert abs(got1[0] - float((d * d).sum())) <= 1e-12 * got1[0] and got1[1] == float(d.abs().max())
    d2 = a2[:, :w2] - b2[:, :w2]
    assert abs(got2[0] - float((d2 * d2).sum())) <= 1e-12 * got2[0] and got2[1] == float(d2.abs().max())


def test_noalloc_rejects_more_rows_than_its_scratch_holds():
    from mpi_cuda_process_b200.ops import _lib as lib_mod
    lib = lib_mod.cuda_lib()
    cap = lib.mcp_residual2d_noalloc_max_rows()
    assert cap == 65536
    tiny = torch.zeros(64, dtype=torch.float32, device="cuda")
    out = torch.zeros(2, dtype=torch.float64, device="cuda")
    # refused before any memory is touched, so the small buffers are never read out of bounds
    rc = lib.mcp_residual2d_noalloc(0, tiny.data_ptr(), tiny.data_ptr(), 32, 32, 0, cap + 1, out.data_ptr(), None)
    assert rc < 0
    msg = lib.mcp_last_error().decode()
    assert "exceed" in msg and str(cap) in msg
    with pytest.raises(lib_mod.NativeError):
        lib_mod.check(rc, "mcp_residual2d_noalloc")
