"""ctypes loader for the native layer (``include/mcp_b200.h``).

The CUDA library is mandatory whenever a GPU is present: if it cannot be loaded on a GPU box the
import fails loudly instead of falling back to PyTorch ops.  On a CPU-only box only the host subset
(``libmcp_host.so``) is loaded and every GPU entry point raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

from . import build as _build

_cuda_lib: Optional[C.CDLL] = None
_host_lib: Optional[C.CDLL] = None

c_void_pp = C.POINTER(C.c_void_p)


class NativeError(RuntimeError):
    pass


class FusedDesc(C.Structure):
    """mcp_fused2d_desc (include/mcp_b200.h)."""

    _fields_ = [
        ("op", C.c_int), ("dtype", C.c_int), ("h", C.c_int), ("w", C.c_int),
        ("row_start", C.c_int), ("rows_own", C.c_int), ("ghost", C.c_int),
        ("pitch", C.c_int64),
        ("buf", C.c_void_p * 2), ("up_buf", C.c_void_p * 2), ("dn_buf", C.c_void_p * 2),
        ("up_rows_own", C.c_int),
        ("flags", C.c_void_p), ("up_flag", C.c_void_p), ("dn_flag", C.c_void_p),
        ("ylo", C.c_int), ("yhi", C.c_int), ("xlo", C.c_int), ("xhi", C.c_int),
        ("chunk_rows", C.c_int), ("max_ctas", C.c_int),
        ("world", C.c_int), ("rank", C.c_int),
        ("res_inbox", C.c_void_p), ("res_flag", C.c_void_p),
        ("peer_res_inbox", C.c_void_p * 16), ("peer_res_flag", C.c_void_p * 16),
        ("nvls_local", C.c_void_p), ("nvls_mc", C.c_void_p),
        ("strip_flags", C.c_void_p), ("up_strip_flag", C.c_void_p), ("dn_strip_flag", C.c_void_p),
    ]


class Fused3dDesc(C.Structure):
    """mcp_fused3d_desc (include/mcp_b200.h)."""

    _fields_ = [
        ("dtype", C.c_int), ("d", C.c_int), ("h", C.c_int), ("w", C.c_int),
        ("z_start", C.c_int), ("nz", C.c_int), ("y_start", C.c_int), ("ny", C.c_int),
        ("pitch", C.c_int64), ("py_alloc", C.c_int), ("pz_alloc", C.c_int),
        ("buf", C.c_void_p * 2), ("nbr_buf", (C.c_void_p * 2) * 4), ("nbr_n", C.c_int * 4),
        ("flags", C.c_void_p), ("nbr_flag", C.c_void_p * 4), ("expect", C.c_int * 4),
        ("coeff", C.c_double), ("zchunk", C.c_int),
        ("world", C.c_int), ("rank", C.c_int),
        ("res_inbox", C.c_void_p), ("res_flag", C.c_void_p),
        ("peer_res_inbox", C.c_void_p * 16), ("peer_res_flag", C.c_void_p * 16),
    ]


def _declare_host(lib: C.CDLL) -> None:
    lib.mcp_host_life_init_reference.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float]
    lib.mcp_host_life_init_reference.restype = C.c_int
    lib.mcp_host_mdf_step_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
    lib.mcp_host_mdf_step_f32.restype = C.c_int
    lib.mcp_host_mdf_step_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    lib.mcp_host_mdf_step_f64.restype = C.c_int
    lib.mcp_host_life_step.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
    lib.mcp_host_life_step.restype = C.c_int
    lib.mcp_host_print_array.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_int64]
    lib.mcp_host_print_array.restype = C.c_int64


def _declare_cuda(lib: C.CDLL) -> None:
    _declare_host(lib)
    i, i64, vp, dbl, u32 = C.c_int, C.c_int64, C.c_void_p, C.c_double, C.c_uint32
    lib.mcp_last_error.restype = C.c_char_p
    lib.mcp_version.restype = i
    lib.mcp_device_info.argtypes = [C.POINTER(i), C.POINTER(i), C.POINTER(i), C.POINTER(C.c_size_t)]
    lib.mcp_arena_alloc.argtypes = [C.c_size_t, c_void_pp]
    lib.mcp_arena_free.argtypes = [vp]
    lib.mcp_ipc_handle_size.restype = i
    lib.mcp_ipc_get_handle.argtypes = [vp, vp]
    lib.mcp_ipc_open_handle.argtypes = [vp, c_void_pp]
    lib.mcp_ipc_close_handle.argtypes = [vp]
    lib.mcp_enable_peer_access.argtypes = [i]
    lib.mcp_init_ring2d.argtypes = [i, vp, i64, i, i, i, i, dbl, dbl, vp]
    lib.mcp_init_shape2d.argtypes = [i, i, vp, i64, i, i, i, i, dbl, dbl, u32, vp]
    lib.mcp_init_life_hash.argtypes = [vp, i64, i, i, i, i, dbl, u32, vp]
    lib.mcp_init_life_hash_bits.argtypes = [vp, i64, i, i, i, i, dbl, u32, vp]
    lib.mcp_apply_ring2d.argtypes = [i, vp, i64, i, i, i, i, dbl, vp]
    lib.mcp_init_shell3d.argtypes = [i, vp, i64, i64, i, i, i, i, i, i, i, dbl, dbl, vp]
    lib.mcp_step2d.argtypes = [i, i, vp, vp, i64, i, i, i, i, i, i, i, i, i, i, vp]
    lib.mcp_step3d.argtypes = [i, vp, vp, i64, i64, i, i, i, i, i, i, i, i, i, i, i, dbl, vp]
    lib.mcp_step3d_cluster.argtypes = [i, vp, vp, i64, i64, i, i, i, i, i, i, i, i, i, i, i, dbl, vp]
    lib.mcp_fused2d_create.argtypes = [C.POINTER(FusedDesc), c_void_pp]
    lib.mcp_fused2d_run.argtypes = [vp, i, i, i, vp]
    lib.mcp_fused2d_run_to_tolerance.argtypes = [vp, i, i, i, dbl, C.POINTER(i), vp]
    lib.mcp_fused2d_stop_reason.argtypes = [vp]
    lib.mcp_fused2d_set_option.argtypes = [vp, i, C.c_longlong]
    lib.mcp_fused2d_run_group.argtypes = [c_void_pp, i, i, i, i, vp]
    lib.mcp_fused2d_plan_schedule2.argtypes = [i, i, i, i, i, i, i, i, i, vp, i, C.POINTER(i)]
    lib.mcp_fused2d_result_async.argtypes = [vp, vp, vp]
    lib.mcp_fused2d_clear_status.argtypes = [vp]
    lib.mcp_fused2d_tile_at.argtypes = [i, i, i, C.POINTER(i), C.POINTER(i)]
    lib.mcp_fused2d_plan_schedule.argtypes = [i, i, i, i, i, i, i, i, vp, i, C.POINTER(i)]
    lib.mcp_fused2d_tile_trace.argtypes = [vp, i, vp, vp, vp, i]
    lib.mcp_fused2d_tile_trace2.argtypes = [vp, i, i, vp, i]
    lib.mcp_fused2d_expected_flags.argtypes = [vp, vp]
    lib.mcp_fused2d_parity.argtypes = [vp]
    lib.mcp_fused2d_set_parity.argtypes = [vp, i]
    lib.mcp_fused2d_residuals.argtypes = [vp, vp, i]
    lib.mcp_fused2d_status.argtypes = [vp]
    lib.mcp_fused2d_launches.argtypes = [vp]
    lib.mcp_fused2d_launches.restype = i64
    lib.mcp_fused2d_geometry.argtypes = [vp, i, C.POINTER(i), C.POINTER(i), C.POINTER(i), C.POINTER(i)]
    lib.mcp_fused2d_chunks.argtypes = [vp, i, vp, i]
    lib.mcp_fused2d_destroy.argtypes = [vp]
    lib.mcp_fused3d_create.argtypes = [C.POINTER(Fused3dDesc), c_void_pp]
    lib.mcp_fused3d_face_tiles.argtypes = [i, i, i, i, i, C.POINTER(i)]
    lib.mcp_fused3d_run.argtypes = [vp, i, vp]
    lib.mcp_fused3d_run_residual.argtypes = [vp, i, i, vp]
    lib.mcp_fused3d_run_to_tolerance.argtypes = [vp, i, i, dbl, C.POINTER(i), vp]
    lib.mcp_fused3d_stop_reason.argtypes = [vp]
    lib.mcp_fused3d_residuals.argtypes = [vp, vp, i]
    lib.mcp_fused3d_result_async.argtypes = [vp, vp, vp]
    lib.mcp_fused3d_set_timeout_ms.argtypes = [vp, C.c_longlong]
    lib.mcp_fused3d_parity.argtypes = [vp]
    lib.mcp_fused3d_set_parity.argtypes = [vp, i]
    lib.mcp_fused3d_status.argtypes = [vp]
    lib.mcp_fused3d_launches.argtypes = [vp]
    lib.mcp_fused3d_launches.restype = i64
    lib.mcp_fused3d_destroy.argtypes = [vp]
    lib.mcp_residual2d_scratch_bytes.restype = C.c_size_t
    lib.mcp_residual2d.argtypes = [i, vp, vp, i64, i, i, i, vp, vp, vp]
    lib.mcp_residual2d_noalloc_max_rows.restype = i
    lib.mcp_residual2d_noalloc.argtypes = [i, vp, vp, i64, i, i, i, vp, vp]
    lib.mcp_checksum2d.argtypes = [i, vp, i64, i, i, i, i64, vp, vp]
    lib.mcp_pack_rows.argtypes = [i, i, vp, i64, i, i, i, vp, i64, vp]
    lib.mcp_pack_block.argtypes = [i, i, vp, i64, i64, i, i, i, i, i, vp, i64, i64, i64, vp]


def host_lib() -> C.CDLL:
    """CPU reference subset; built on demand with g++ (no CUDA needed)."""
    global _host_lib
    if _host_lib is None:
        path = _build.build_host()          # returns at once when the source hash matches the built library
        _host_lib = C.CDLL(str(path))
        _declare_host(_host_lib)
    return _host_lib


def cuda_lib() -> C.CDLL:
    """The sm_100a library.  Raises NativeError if it is missing and cannot be built."""
    global _cuda_lib
    if _cuda_lib is None:
        # always go through the build: a library left over from older sources (e.g. a changed descriptor layout)
        # must never be loaded; the call returns at once when the source hash matches
        import os
        override = os.environ.get("MCP_B200_LIB")      # development only: load an experimental build of the library
        try:
            path = override or _build.build_cuda()
        except Exception as e:  # pragma: no cover - depends on toolchain
            raise NativeError(f"libmcp_b200.so is stale or missing and could not be built: {e}") from e
        try:
            _cuda_lib = C.CDLL(str(path))
        except OSError as e:
            raise NativeError(f"cannot load {path}: {e}") from e
        _declare_cuda(_cuda_lib)
    return _cuda_lib


def cuda_lib_loaded() -> bool:
    return _cuda_lib is not None


def check(rc: int, what: str = "") -> int:
    """Raise NativeError with the library's message if ``rc`` is a negative MCP_ERR_* code."""
    if rc < 0:
        msg = cuda_lib().mcp_last_error().decode(errors="replace") if _cuda_lib is not None else ""
        raise NativeError(f"{what or 'native call'} failed (code {rc}): {msg}")
    return rc


def library_paths():
    return {"cuda": str(_build.lib_path()), "host": str(_build.host_lib_path())}
