"""bench.py's contract pieces that do not need a GPU: the reference arm prints ONE JSON line saying why the unmodified
reference cannot run (two bare .cu programs, no MPI on the box, deadlock for g >= 2: MDF_kernel.cu:168 vs :215) and exits
0; the ref-fixed arm degrades to an 'unavailable' line where no GPU is visible; the flags the driver passes parse."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _one_json_line(args, env=None):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True,
                       timeout=300, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stderr
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    return json.loads(lines[0])


def test_reference_arm_prints_unavailable_and_exits_zero():
    for n in ("1", "8"):
        rec = _one_json_line(["--impl", "reference", "--gpus", n, "--steps", "3", "--warmup", "3"])
        assert rec["impl"] == "reference" and "unavailable" in rec and "\n" not in rec["unavailable"]
        assert rec["n_gpus"] == int(n) and "GLUPS" in rec["metric"]


def test_reference_arm_under_a_launcher_prints_once():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, timeout=300, cwd=ROOT, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""          # only rank 0 prints


def test_ref_fixed_arm_degrades_without_a_gpu():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    rec = _one_json_line(["--impl", "ref-fixed", "--gpus", "1", "--height", "64", "--width", "64"], env=env)
    assert rec["impl"] == "ref-fixed" and ("unavailable" in rec or rec["value"] > 0)


def test_driver_flags_parse():
    sys.path.insert(0, ROOT)
    import bench
    a = bench.parse_args(["--gpus", "8", "--steps", "20", "--warmup", "5"])
    assert (a.gpus, a.steps, a.warmup, a.impl, a.height, a.width, a.gens_per_step) == (8, 20, 5, "fused", 16384, 16384, 1000)
    d = bench.parse_args([])
    assert d.gpus == 1 and d.warmup >= 3 and not d.no_numa_bind
