"""Development helper: build an experimental copy of libmcp_b200.so with extra -D flags for stencil2d_fused.cu /
heat3d_fused.cu (everything else is taken from the regular build), e.g.

    python scripts/build_variant.py v_scalar12 -DMCP_F32_SCALAR -DMCP_WARPS_F32_TB8=12
    MCP_B200_LIB=$PWD/mpi_cuda_process_b200/ops/_lib/variants/v_scalar12.so python bench.py ...
"""
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_cuda_process_b200.ops import build as b  # noqa: E402

name, flags = sys.argv[1], sys.argv[2:]
files = [f[5:] for f in flags if f.startswith("FILE=")] or ["stencil2d_fused.cu"]
flags = [f for f in flags if not f.startswith("FILE=")]
b.build_cuda()
vdir = b.LIBDIR / "variants"
vdir.mkdir(exist_ok=True)
objs = []
for src in b.CUDA_SOURCES:
    if src in files:
        obj = vdir / f"{name}.{src}.o"
        cmd = [b._nvcc(), *b.NVCC_FLAGS, *flags, "-Xptxas", "-v", "-I", str(b.INCLUDE), "-c", str(b.CSRC / src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            sys.exit(r.stderr)
        (vdir / f"{name}.{src}.ptxas.txt").write_text(r.stderr)
        objs.append(obj)
    else:
        objs.append(b.OBJDIR / (src + ".o"))
out = vdir / f"{name}.so"
subprocess.run([b._nvcc(), "-shared", "-o", str(out), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a",
                "-Xcompiler", "-fPIC"], check=True)
print(out)
