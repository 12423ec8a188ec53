"""Speed-of-light of the fused 2-D sweep kernel as built: how fast could the row loop go if every SM sub-partition
issued one instruction of it per cycle (and, for the packed operator, if the FMA pipe never waited)?

The kernel is not HBM-bound at TB = 8 (DRAM 39 % in ncu), so the HBM roofline says little about it; the binding
resources are instruction issue (scalar operator) and the FMA pipe (packed operator: an f32x2 instruction holds the pipe
for two cycles).  Both bounds follow from the SASS of the lean row loop, which this script reads from the built object
(no GPU needed), and from three schedule factors:

    ceiling = SMs x 4 sub-partitions x clock x (cell updates per stage and warp) / (cycles per stage)
              x (lanes that produce output) / 32 x (output rows of a tile) / (rows a tile processes)

    python scripts/issue_ceiling.py [--mhz 1965] [--sms 148] [--rows 236] [--measured 'fp32 scalar TB=8=3279']
"""
import argparse
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "mpi_cuda_process_b200", "ops", "_build", "stencil2d_fused.cu.o")

# (label, mangled-name regex, TB, cells per lane, margin lanes per side)
VARIANTS = [
    ("fp32 scalar TB=8", r"fused2d_kernelINS_12MdfF32ScalarELi8ELb0ELi8", 8, 4, 2),
    ("fp32 packed TB=8", r"fused2d_kernelINS_12MdfF32PackedELi8ELb0ELi8", 8, 4, 2),
    ("fp32 packed TB=4", r"fused2d_kernelINS_12MdfF32PackedELi4ELb0ELi16", 4, 4, 1),
    ("fp64 TB=8", r"fused2d_kernelINS_6MdfF64ELi8ELb0ELi8", 8, 2, 4),
    ("fp64 TB=4", r"fused2d_kernelINS_6MdfF64ELi4ELb0ELi16", 4, 2, 2),
]
ROWS_PER_STAGE = 6
FMA_PIPE_2CYC = re.compile(r"^(FADD2|FFMA2|FMUL2)")
FMA_PIPE_1CYC = re.compile(r"^(FADD|FFMA|FMUL|IMAD|FSEL|FMNMX)(\.|$)")
FP64 = re.compile(r"^(DADD|DFMA|DMUL)")


def lean_loop(pattern):
    txt = subprocess.run(["cuobjdump", "-sass", OBJ], capture_output=True, text=True, check=True).stdout
    for part in txt.split("Function : ")[1:]:
        if not re.search(pattern, part.split("\n", 1)[0]):
            continue
        ins = [l for l in part.split("\n") if re.match(r"\s+/\*[0-9a-f]{4,6}\*/", l)]
        blocks, cur = [], []
        for l in ins:
            cur.append(l)
            if re.search(r"\*/\s+(@!?U?P\d+\s+)?(BRA|EXIT|RET|BRX|CALL|BSYNC|WARPSYNC)", l):
                blocks.append(cur)
                cur = []
        # the lean stage body: the longest branch-free run that stores a whole stage and selects nothing
        cand = [b for b in blocks if sum("STG" in x for x in b) >= 5 and not any("FSEL" in x for x in b)]
        return max(cand, key=len) if cand else []
    return []


def opcode(line):
    m = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    return m.group(1) if m else ""


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mhz", type=float, default=1965.0, help="SM clock under load (bench.py clocks.sm_mhz)")
    ap.add_argument("--sms", type=int, default=148)
    ap.add_argument("--rows", type=int, default=236, help="output rows per tile (chunk height)")
    ap.add_argument("--measured", default="", help="measured GLUPS per GPU as label=value pairs, e.g. "
                    "'fp32 scalar TB=8=3279,fp32 packed TB=8=3105'")
    a = ap.parse_args()
    measured = {k.strip(): float(v) for k, _, v in (item.rpartition("=") for item in a.measured.split(",") if item)}
    slots = a.sms * 4 * a.mhz * 1e6                      # warp-instructions per second the GPU can issue
    print(f"{a.sms} SMs x 4 sub-partitions x {a.mhz:.0f} MHz = {slots / 1e9:.0f} G warp-instructions/s")
    print(f"{'kernel':18} {'instr':>6} {'fma-pipe cyc':>12} {'fp64':>5} {'bound cyc':>9} {'upd/stage':>9} "
          f"{'lanes':>6} {'rows':>6} {'ceiling GLUPS':>13}")
    for label, pat, tb, cpl, margin in VARIANTS:
        body = lean_loop(pat)
        if not body:
            print(f"{label:18} (not found in {os.path.basename(OBJ)})")
            continue
        ops = [opcode(l) for l in body]
        n = len(ops)
        fma = sum(2 for o in ops if FMA_PIPE_2CYC.match(o)) + sum(1 for o in ops if FMA_PIPE_1CYC.match(o))
        f64 = sum(1 for o in ops if FP64.match(o))
        # fp64: B200's FP64 pipe issues one warp instruction per 2 cycles per sub-partition (64 lanes/SM/clk, ncu's
        # sm__pipe_fp64 denominator); everything else one per cycle
        cycles = max(n, fma, 2 * f64)
        updates = ROWS_PER_STAGE * tb * cpl * 32
        lane_eff = (32 - 2 * margin) / 32
        row_eff = a.rows / (a.rows + 2 * tb)
        ceiling = slots * updates / cycles * lane_eff * row_eff / 1e9
        which = "issue" if cycles == n else ("fma pipe" if cycles == fma else "fp64 pipe")
        print(f"{label:18} {n:6d} {fma:12d} {f64:5d} {cycles:9d} {updates:9d} {lane_eff:6.3f} {row_eff:6.3f} "
              f"{ceiling:13.0f}  ({which})"
              + (f"  measured {measured[label]:.0f} = {measured[label] / ceiling:.2f}" if label in measured else ""))


if __name__ == "__main__":
    main()
