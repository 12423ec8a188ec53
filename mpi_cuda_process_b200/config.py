"""Run configuration.

The reference has no configuration system at all: three ``scanf("%d")`` values
(generations, height, width; reference MDF_kernel.cu:103-112, kernel.cu:150-159), a
compile-time ``NUM_THREADS 512`` (MDF_kernel.cu:6) and literals scattered through the
code (``0.15`` at kernel.cu:193, ``100`` at MDF_kernel.cu:35,93, ``0.25`` at
MDF_kernel.cu:20, world size 2 at MDF_kernel.cu:29,37).  ``RunConfig`` names every one
of those knobs and adds the extension knobs (3-D, fp64, tolerance, decomposition).
"""
from __future__ import annotations

import dataclasses
import json
from dataclasses import dataclass, field
from typing import Optional, Tuple

WORKLOADS = ("mdf", "life", "heat3d")
BACKENDS = ("auto", "fused", "nccl", "staged", "cpu")
DTYPES = ("fp32", "fp64", "int32", "bits")
BC_MODES = ("fixed-ring", "dead-outside")


@dataclass
class RunConfig:
    """Everything one run needs.

    workload      "mdf" (2-D 5-point heat/Jacobi, MDF_kernel.cu:10-22), "life"
                  (Game of Life B3/S23, kernel.cu:10-68) or "heat3d" (3-D 7-point,
                  extension named by BASELINE.json config 3).
    generations   number of sweeps ``g`` (MDF_kernel.cu:106); ignored when ``tol`` is set
                  except as an upper bound (``max_generations``).
    height/width  grid ``h`` x ``w`` (MDF_kernel.cu:109,112); ``depth`` adds the third axis.
    dtype         storage/compute type.  The reference is fp32 for MDF (MDF_kernel.cu:137)
                  and int32 for life (kernel.cu:184); fp64 is an extension.
    bc            life only: "dead-outside" = what kernel.cu:23-64 tries to say (cells
                  outside the grid are dead, edge cells evolve); "fixed-ring" = what
                  create_universe + the halo exchange say (ring held dead,
                  kernel.cu:137-139,214,228).  MDF is always fixed-ring (Dirichlet).
    ring/interior MDF initial values (MDF_kernel.cu:92-96).
    prob          life initial alive probability (kernel.cu:193).
    seed          None = reproduce the reference's unseeded glibc ``rand()`` stream
                  (kernel.cu:141); an int = device Philox-style hash init (extension).
    backend       "fused" = persistent sm_100a kernel with in-kernel NVLink halo pushes;
                  "nccl" = plain kernel + NCCL send/recv halo (baseline); "staged" =
                  the reference's host-staged pattern (baseline); "cpu" = torch CPU + gloo.
    time_block    temporal blocking depth T of the fused kernel (generations advanced per
                  HBM pass); also the ghost-layer depth exchanged every T generations.
    tol           run-to-tolerance on the global residual (extension); None = fixed g.
    residual      "none" | "l2" | "linf": fused global residual of u' - u.
    """

    workload: str = "mdf"
    generations: int = 1
    height: int = 256
    width: int = 256
    depth: int = 0
    dtype: str = "fp32"
    bc: Optional[str] = None             # None = workload default (life: dead-outside, others: fixed-ring)
    ring: float = 100.0
    interior: float = 0.0
    prob: float = 0.15
    init: str = "ring"                   # mdf: ring (reference) | hash (random interior) | box (hot box)
    seed: Optional[int] = None
    backend: str = "auto"
    decomposition: str = "slab"          # "slab" (1-D rows) or "pencil" (2-D over a 3-D grid)
    pencil: Tuple[int, int] = (0, 0)     # (Py, Pz) for pencil; 0 = choose
    time_block: int = 8
    tol: Optional[float] = None
    max_generations: int = 0
    residual: str = "none"
    check_every: int = 0                 # residual check period (0 = every time block)
    halo_message: str = "row"            # staged baseline: "element" (reference) or "row"
    output: Optional[str] = None
    output_format: str = "ascii"         # ascii | npy | raw | text
    strict_fp64_accum: bool = False      # MDF fp32: do the final fma in fp64 like the reference's DFMA
    checkpoint_every: int = 0
    checkpoint_dir: Optional[str] = None
    extra: dict = field(default_factory=dict)

    def __post_init__(self):
        self.validate()

    def validate(self) -> None:
        if self.workload not in WORKLOADS:
            raise ValueError(f"workload must be one of {WORKLOADS}, got {self.workload!r}")
        if self.backend not in BACKENDS:
            raise ValueError(f"backend must be one of {BACKENDS}, got {self.backend!r}")
        if self.dtype not in DTYPES:
            raise ValueError(f"dtype must be one of {DTYPES}, got {self.dtype!r}")
        if self.bc is None:
            self.bc = "dead-outside" if self.workload == "life" else "fixed-ring"
        if self.bc not in BC_MODES:
            raise ValueError(f"bc must be one of {BC_MODES}, got {self.bc!r}")
        if self.workload == "life":
            if self.dtype != "bits":        # "bits" = one cell per bit (fused GPU backend only); default int32 cells
                self.dtype = "int32"
        elif self.dtype in ("int32", "bits"):
            raise ValueError(f"{self.dtype} is only valid for the life workload")
        if self.workload == "heat3d" and self.depth <= 0:
            raise ValueError("heat3d needs depth > 0")
        if self.workload != "heat3d" and self.depth not in (0, 1):
            raise ValueError("depth is only meaningful for heat3d")
        if self.generations < 0:
            raise ValueError("generations must be >= 0")
        if self.height < 1 or self.width < 1:
            raise ValueError("height and width must be >= 1")
        if self.time_block < 1:
            raise ValueError("time_block must be >= 1")
        if self.residual not in ("none", "l2", "linf"):
            raise ValueError("residual must be none|l2|linf")
        if self.tol is not None and self.residual == "none":
            self.residual = "linf"
        if self.init not in ("ring", "hash", "box"):
            raise ValueError("init must be ring|hash|box")
        if self.halo_message not in ("row", "element"):
            raise ValueError("halo_message must be row|element")

    @property
    def ndim(self) -> int:
        return 3 if self.workload == "heat3d" else 2

    @property
    def shape(self) -> Tuple[int, ...]:
        if self.ndim == 3:
            return (self.depth, self.height, self.width)
        return (self.height, self.width)

    @property
    def cells(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n

    def to_json(self) -> str:
        return json.dumps(dataclasses.asdict(self), sort_keys=True)

    @classmethod
    def from_json(cls, s: str) -> "RunConfig":
        d = json.loads(s)
        if "pencil" in d:
            d["pencil"] = tuple(d["pencil"])
        return cls(**d)
