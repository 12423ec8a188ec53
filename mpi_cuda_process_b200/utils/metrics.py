"""Throughput and roofline arithmetic (GLUPS, achieved bandwidth, fraction of the measured peaks)."""
from __future__ import annotations

import json
from pathlib import Path
from typing import Dict, Optional

FALLBACK_HBM_GBS = 6650.0
NVLINK_GBS_PER_DIR = 770.0      # measured peer copy, B200_PROFILING.md


def measured_peaks(repo_root: Optional[str] = None) -> Dict:
    """MEASURED_PEAKS.json (driver-written) or the documented fallback."""
    root = Path(repo_root) if repo_root else Path(__file__).resolve().parents[2]
    p = root / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            d = json.loads(p.read_text())
            d["source"] = "measured"
            return d
        except (OSError, ValueError):
            pass
    return {"hbm_gbs": FALLBACK_HBM_GBS, "source": "fallback"}


def glups(cells: int, generations: int, ms: float) -> float:
    return cells * generations / (ms * 1e-3) / 1e9 if ms > 0 else 0.0


def roofline(cells_per_gpu: int, generations: int, ms: float, bytes_per_update: int, halo_bytes_per_gen: float = 0.0,
             peaks: Optional[Dict] = None, time_block: int = 1) -> Dict:
    """Achieved fraction of the per-GPU roofline: the slower of HBM stencil time and NVLink halo time.

    ``fraction_of_roofline`` uses the one-pass denominator (one read + one write per update per generation); a
    temporally blocked kernel advances ``time_block`` generations per pass over the field, so that fraction can
    exceed 1.  ``fraction_of_time_block_roofline`` divides by the kernel's own bound instead: one read + one write
    per ``time_block`` generations (what its DRAM traffic actually is, see the ncu summaries in profiles/); it is
    the number to read as "how close to the memory system's limit" -- below 1 means something else (instruction
    issue, FMA pipe) is the binding resource.
    """
    peaks = peaks or measured_peaks()
    hbm = peaks["hbm_gbs"] * 1e9
    t_hbm = cells_per_gpu * bytes_per_update / hbm
    t_halo = halo_bytes_per_gen / (NVLINK_GBS_PER_DIR * 1e9)
    t_roof = max(t_hbm, t_halo)
    t_gen = ms * 1e-3 / max(generations, 1)
    return {
        "t_gen_us": t_gen * 1e6, "t_roofline_us": t_roof * 1e6,
        "bound": "hbm" if t_hbm >= t_halo else "nvlink",
        "fraction_of_roofline": t_roof / t_gen if t_gen > 0 else 0.0,
        "time_block": time_block,
        "t_time_block_roofline_us": max(t_hbm / max(time_block, 1), t_halo / max(time_block, 1)) * 1e6,
        "fraction_of_time_block_roofline": max(t_hbm, t_halo) / max(time_block, 1) / t_gen if t_gen > 0 else 0.0,
        "one_pass_equivalent_gbs": cells_per_gpu * bytes_per_update / t_gen / 1e9 if t_gen > 0 else 0.0,
        "dram_gbs_estimate": cells_per_gpu * bytes_per_update / max(time_block, 1) / t_gen / 1e9 if t_gen > 0 else 0.0,
        "peaks_source": peaks.get("source", "measured") + (" (of measured)" if peaks.get("source") == "measured" else " (of fallback)"),
    }
