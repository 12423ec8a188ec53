"""Output: the reference's ``print_array`` format and field dumps.

``print_array`` (kernel.cu:115-129, MDF_kernel.cu:72-86) is the only output format the reference defines:
a leading newline, then per cell ``"0"`` if the cell equals 1 else ``" "``, a newline after every ``w``
cells, and a trailing newline.  It takes ``int[]`` in both programs (so it cannot print the float field
of the MDF program) and every call site is commented out (kernel.cu:232,265, MDF_kernel.cu:185,218).
Here it is byte-compatible for life grids; float fields get raw / .npy / text dumps (no reference format
exists to match).
"""
from __future__ import annotations

import io as _io
import sys
from typing import Optional

import numpy as np


def format_array(u: np.ndarray) -> bytes:
    """Bytes that the reference's ``print_array(arr, w, size)`` would write for an int grid."""
    u = np.asarray(u)
    if u.ndim != 2:
        raise ValueError("print_array is defined for 2-D grids")
    h, w = u.shape
    body = np.full((h, w + 1), ord(" "), dtype=np.uint8)
    body[:, :w][u == 1] = ord("0")
    body[:, w] = ord("\n")
    return b"\n" + body.tobytes() + b"\n"


def print_array(u: np.ndarray, file=None) -> None:
    out = file if file is not None else sys.stdout
    data = format_array(u)
    if hasattr(out, "buffer"):
        out.flush()
        out.buffer.write(data)
        out.buffer.flush()
    elif isinstance(out, (_io.RawIOBase, _io.BufferedIOBase)):
        out.write(data)
    else:
        out.write(data.decode("ascii"))


def parse_array(data: bytes, w: int) -> np.ndarray:
    """Inverse of :func:`format_array`."""
    if not (data.startswith(b"\n") and data.endswith(b"\n")):
        raise ValueError("not print_array output")
    body = np.frombuffer(data[1:-1], dtype=np.uint8)
    if body.size % (w + 1):
        raise ValueError("size mismatch")
    body = body.reshape(-1, w + 1)
    return (body[:, :w] == ord("0")).astype(np.int32)


def dump_field(u: np.ndarray, path: str, fmt: str = "npy") -> None:
    """Write a gathered field.  fmt: ascii (life only) | npy | raw | text."""
    if fmt == "ascii":
        if not np.issubdtype(u.dtype, np.integer):
            raise ValueError("the ascii format is print_array's and is defined for int grids only; "
                             "use npy/raw/text for float fields")
        with open(path, "wb") as f:
            f.write(format_array(u))
    elif fmt == "npy":
        np.save(path, u)
    elif fmt == "raw":
        np.ascontiguousarray(u).tofile(path)
    elif fmt == "text":
        np.savetxt(path, u.reshape(-1, u.shape[-1]), fmt="%.9g" if u.dtype != np.int32 else "%d")
    else:
        raise ValueError(f"unknown output format {fmt!r}")


def load_field(path: str, fmt: str = "npy", shape=None, dtype=None, w: Optional[int] = None) -> np.ndarray:
    if fmt == "npy":
        return np.load(path)
    if fmt == "raw":
        return np.fromfile(path, dtype=dtype).reshape(shape)
    if fmt == "ascii":
        with open(path, "rb") as f:
            return parse_array(f.read(), w)
    if fmt == "text":
        return np.loadtxt(path, dtype=dtype).reshape(shape)
    raise ValueError(fmt)
