"""C-ABI argument validation: every bad call returns a negative code and a message (the reference checks nothing)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _desc(lib_mod, buf0, buf1, **kw):
    d = lib_mod.FusedDesc()
    d.op, d.dtype, d.h, d.w = 0, 0, 64, 64
    d.row_start, d.rows_own, d.ghost, d.pitch = 0, 64, 4, 64
    d.buf[0], d.buf[1] = buf0.data_ptr(), buf1.data_ptr()
    d.ylo, d.yhi, d.xlo, d.xhi = 1, 62, 1, 62
    for k, v in kw.items():
        setattr(d, k, v)
    return d


def test_fused2d_create_rejects_bad_descriptors():
    from mpi_cuda_process_b200.ops import _lib as lib_mod
    lib = lib_mod.cuda_lib()
    a = torch.zeros((72, 64), dtype=torch.float32, device="cuda")
    b = torch.zeros_like(a)
    plan = C.c_void_p()
    for kw, needle in [(dict(pitch=60), "pitch"), (dict(ghost=0), "ghost"), (dict(ghost=9), "ghost"),
                       (dict(op=1), "dtype"), (dict(dtype=2), "dtype"), (dict(rows_own=0), "empty")]:
        rc = lib.mcp_fused2d_create(C.byref(_desc(lib_mod, a, b, **kw)), C.byref(plan))
        assert rc < 0, kw
        assert needle in lib.mcp_last_error().decode(), (kw, lib.mcp_last_error())
    # a neighbour without flags, and a slab thinner than the ghost depth
    d = _desc(lib_mod, a, b, rows_own=2)
    d.up_buf[0], d.up_buf[1] = a.data_ptr(), b.data_ptr()
    assert lib.mcp_fused2d_create(C.byref(d), C.byref(plan)) < 0
    # a good one still works afterwards
    d = _desc(lib_mod, a, b)
    lib_mod.check(lib.mcp_fused2d_create(C.byref(d), C.byref(plan)))
    assert lib.mcp_fused2d_run(plan, -1, 4, 0, None) < 0
    assert lib.mcp_fused2d_run(plan, 4, 0, 0, None) < 0
    lib_mod.check(lib.mcp_fused2d_run(plan, 4, 4, 0, None))
    torch.cuda.synchronize()
    assert lib.mcp_fused2d_status(plan) == 0
    lib.mcp_fused2d_destroy(plan)


def test_plain_and_pack_entry_points_reject_bad_arguments():
    from mpi_cuda_process_b200.ops import _lib as lib_mod
    lib = lib_mod.cuda_lib()
    a = torch.zeros((16, 64), dtype=torch.float32, device="cuda")
    assert lib.mcp_step2d(0, 0, a.data_ptr(), a.data_ptr(), 64, 16, 0, 16, 64, 1, 15, 1, 14, 1, 62, None) < 0   # aliased
    b = torch.zeros_like(a)
    assert lib.mcp_step2d(0, 0, a.data_ptr(), b.data_ptr(), 62, 16, 0, 16, 64, 1, 15, 1, 14, 1, 62, None) < 0   # pitch
    assert lib.mcp_step2d(7, 0, a.data_ptr(), b.data_ptr(), 64, 16, 0, 16, 64, 1, 15, 1, 14, 1, 62, None) < 0   # op
    assert lib.mcp_pack_rows(0, 2, a.data_ptr(), 64, 64, 0, 16, b.data_ptr(), 0, None) < 0                      # ascii of floats
    assert lib.mcp_init_shape2d(0, 5, a.data_ptr(), 64, 16, 0, 16, 64, 100.0, 0.0, 0, None) < 0                 # kind
    with pytest.raises(lib_mod.NativeError):
        lib_mod.check(lib.mcp_residual2d(9, a.data_ptr(), b.data_ptr(), 64, 64, 0, 16, a.data_ptr(), a.data_ptr(), None))


def test_apply_ring_kernel():
    from mpi_cuda_process_b200.ops import _lib as lib_mod
    lib = lib_mod.cuda_lib()
    h, w, pitch, row0 = 20, 50, 64, -2          # local rows 0..11 hold global rows -2..9
    u = torch.full((12, pitch), 7.0, dtype=torch.float32, device="cuda")
    lib_mod.check(lib.mcp_apply_ring2d(0, u.data_ptr(), pitch, 12, row0, h, w, 100.0, None))
    v = u.cpu().numpy()
    assert (v[2, :w] == 100).all()                       # global row 0
    assert (v[3:, 0] == 100).all() and (v[3:, w - 1] == 100).all()
    assert (v[3:, 1:w - 1] == 7).all() and (v[:2] == 7).all() and (v[:, w:] == 7).all()
