"""The Python examples run (CPU backend here), and the multi-rank run of the glider prints the same frames as the
single-process run -- the glider crosses the slab cuts.  The reference has no examples and no output at all (its
print_array calls are commented out, kernel.cu:232,265)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = dict(os.environ, CUDA_VISIBLE_DEVICES="")          # examples pick the CPU backend when no GPU is visible


def _run(cmd, **kw):
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=ENV, **kw)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


def test_heat_plate_example_converges():
    out = _run([sys.executable, "examples/heat_plate.py", "--size", "40", "--tol", "0.05"])
    assert "40 x 40 fp32 on 1 rank(s)" in out and "generations" in out
    last = float(out.split("last max|du| = ")[1].split(",")[0])
    assert last <= 0.05


def test_life_glider_example_same_frames_on_three_ranks():
    args = ["examples/life_glider.py", "--height", "18", "--width", "20", "--generations", "24", "--every", "8"]
    one = _run([sys.executable] + args)
    three = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=3",
                  "--master-addr", "127.0.0.1", "--master-port", "29621"] + args)
    frames = one[one.index("generation 8"):]
    assert frames.count("generation") == 3 and "5 live cells" in frames
    assert frames in three                               # torchrun adds banner lines around it
