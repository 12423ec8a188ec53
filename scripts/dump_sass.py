"""SASS evidence for profiles/sass/: full `cuobjdump -sass` listings of the named kernels (gzip) and, uncompressed, the
parts that matter -- the lean row loop of the fused 2-D kernel, and every line with a TMA, mbarrier, cluster, NVLink
signalling, multimem or packed-fp32 instruction.  Runs on the build box (no GPU needed).

    python scripts/dump_sass.py
"""
import gzip
import io
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "mpi_cuda_process_b200", "ops", "_build")
OUT = os.path.join(ROOT, "profiles", "sass")
os.makedirs(OUT, exist_ok=True)

KERNELS = [  # (object, mangled-name regex, file stem)
    ("stencil2d_fused.cu.o", r"fused2d_kernelINS_12MdfF32PackedELi8ELb0ELi8", "fused2d_mdf_f32_tb8"),
    ("stencil2d_fused.cu.o", r"fused2d_kernelINS_12MdfF32PackedELi8ELb1ELi8", "fused2d_mdf_f32_tb8_residual"),
    ("stencil2d_fused.cu.o", r"fused2d_kernelINS_12MdfF32ScalarELi8ELb0ELi8", "fused2d_mdf_f32_scalar_tb8"),
    ("stencil2d_fused.cu.o", r"fused2d_kernelINS_6MdfF64ELi4ELb1ELi16", "fused2d_mdf_f64_tb4_residual"),
    ("stencil2d_fused.cu.o", r"fused2d_kernelINS_6LifeB1ELi2ELb0ELi12", "fused2d_life_bits_tb2"),
    ("stencil2d_fused.cu.o", r"fused2d_group_kernelINS_12MdfF32PackedELi8ELb0", "fused2d_group_mdf_f32_tb8"),
    ("heat3d_fused.cu.o", r"heat3d_fused_kernelIfLb0", "heat3d_fused_f32"),
    ("heat3d_fused.cu.o", r"heat3d_fused_kernelIdLb1", "heat3d_fused_f64_residual"),
    ("heat3d_cluster.cu.o", r"step3d_cluster_kernelIf", "heat3d_cluster_f32"),
    ("gather_pack.cu.o", r"pack_block_kernelIffLb1", "pack_block_f32_vector"),
    ("gather_pack.cu.o", r"unpack_bits_kernel", "unpack_bits"),
    ("init_bc.cu.o", r"fill_rows_kernelIfNS_8RingFillIfEELb1", "init_ring2d_f32_vector"),
    ("checksum.cu.o", r"checksum2d_kernelIj", "checksum2d_4byte"),
    ("residual.cu.o", r"residual_pass1If", "residual_pass1_f32"),
]
KEY = re.compile(r"UTMALDG|UTMACCTL|SYNCS|UCGABAR|CGAERRBAR|MEMBAR|FENCE|ERRBAR|\.SYS|MULTIMEM|LDGMC|REDGMC|REDG|RED\.|"
                 r"ATOMG|LD\.E\.128|CCTL|S2UR")
COUNT_ONLY = re.compile(r"FFMA2|FADD2")   # thousands of lines: counted in the README, listed in the row loop


def functions(obj):
    txt = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, obj)], capture_output=True, text=True).stdout
    parts = re.split(r"(?m)^\s*Function : ", txt)
    return {p.split("\n", 1)[0].strip(): p for p in parts[1:]}


cache = {}
index = ["# profiles/sass -- SASS of the sm_100a kernels (cuobjdump -sass, CUDA 12.9)", "",
         "Full listings are `<kernel>.sass.gz`; `<kernel>.key.txt` holds every TMA / mbarrier / cluster / NVLink-signal / "
         "multimem / packed-fp32 line with its address, and `fused2d_mdf_f32_tb8.rowloop.sass` the complete lean row loop "
         "(6 input rows x 8 time levels, no branches).", "", "| kernel | instructions | notable mnemonics (count) |", "|---|---|---|"]
for obj, pat, stem in KERNELS:
    fns = cache.setdefault(obj, functions(obj))
    name = next((n for n in fns if re.search(pat, n)), None)
    if name is None:
        print("not found:", pat, file=sys.stderr)
        continue
    body = fns[name]
    with open(os.path.join(OUT, stem + ".sass.gz"), "wb") as raw, \
            gzip.GzipFile(filename="", fileobj=raw, mode="wb", mtime=0) as gz, \
            io.TextIOWrapper(gz) as f:          # mtime=0: identical input -> identical file
        f.write("Function : " + body)
    ins = [l for l in body.split("\n") if re.match(r"\s+/\*[0-9a-f]{4,6}\*/", l)]
    keys = [re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", l).rstrip() for l in ins if KEY.search(l)]
    open(os.path.join(OUT, stem + ".key.txt"), "w").write(name + "\n" + "\n".join(keys) + "\n")
    count = {}
    for l in [x for x in ins if COUNT_ONLY.search(x)] + keys:
        m = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
        if m:
            count[m.group(1)] = count.get(m.group(1), 0) + 1
    top = ", ".join(f"`{k}` {v}" for k, v in sorted(count.items(), key=lambda x: -x[1])[:10])
    index.append(f"| `{stem}` | {len(ins)} | {top} |")
    if stem == "fused2d_mdf_f32_tb8":
        # the lean row loop = the longest branch-free run with 5+ STG and no FSEL
        blocks, cur = [], []
        for l in ins:
            cur.append(l)
            if re.search(r"\*/\s+(@!?U?P\d+\s+)?(BRA|EXIT|RET|BRX|CALL|BSYNC|WARPSYNC)", l):
                blocks.append(cur)
                cur = []
        best = max((b for b in blocks if sum("STG" in x for x in b) >= 5 and not any("FSEL" in x for x in b)
                    and not any("WARPSYNC" in x for x in b)), key=len, default=None)
        if best:
            clean = [re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", l).rstrip() for l in best]
            open(os.path.join(OUT, stem + ".rowloop.sass"), "w").write(
                f"// lean row loop of {name}: {len(best)} instructions for 6 rows x 8 levels x 4 cells per lane = 192 "
                f"cell updates per lane ({len(best) / 192:.2f} instructions per update)\n" + "\n".join(clean) + "\n")


def resource_usage():
    """Registers / stack / static shared / local (spill) bytes of every kernel in the built library."""
    lib = os.path.join(ROOT, "mpi_cuda_process_b200", "ops", "_lib", "libmcp_b200.so")
    out = subprocess.run(["cuobjdump", "--dump-resource-usage", lib], capture_output=True, text=True).stdout.splitlines()
    rows = []
    for i, l in enumerate(out):
        m = re.match(r"\s*Function (\S+):", l)
        if not m:
            continue
        name = subprocess.run(["cu++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        get = lambda k: int((re.search(k + r":(\d+)", out[i + 1]) or [None, "0"])[1])
        rows.append((name, get("REG"), get("STACK"), get("SHARED"), get("LOCAL"), get(r"CONSTANT\[0\]")))
    rows.sort()
    with open(os.path.join(OUT, "resource_usage.txt"), "w") as f:
        f.write("# cuobjdump --dump-resource-usage of libmcp_b200.so (sm_100a, -lineinfo).  STACK is the call frame of "
                "the\n# __noinline__ wait helpers (8 bytes) and/or register spills (ptxas -v figures: "
                "ptxas_spills.txt).\n# Dynamic shared memory (TMA rings) is set at launch and not listed.\n")
        f.write(f"{'REG':>4} {'STACK':>6} {'SMEM':>6} {'LOCAL':>6} {'CONST0':>7}  kernel\n")
        for n, r, st, sh, lo, c in rows:
            f.write(f"{r:>4} {st:>6} {sh:>6} {lo:>6} {c:>7}  {n}\n")
    return rows


rows = resource_usage()
index.append("")
index.append(f"`resource_usage.txt`: registers, stack, static shared and local memory of all {len(rows)} kernels in the "
             f"library (`cuobjdump --dump-resource-usage`); `ptxas_spills.txt`: the `ptxas -v` figures of the "
             f"sweep kernels.  The 8-warp time-block-8 kernels (up to {max(r[1] for r in rows)} registers) do not spill; "
             f"{sum(1 for r in rows if r[2] > 8)} of the 16-warp instantiations at the 128-register cap spill 16-228 "
             f"bytes (time block 4: fp64, packed fp32, the emulated-ranks kernel; life int32 TB = 2 with residual), and "
             f"so does the fp64 residual instantiation of the 3-D kernel (16 bytes).")
open(os.path.join(OUT, "README.md"), "w").write("\n".join(index) + "\n")
print("\n".join(index[4:]))
