"""Clock / throttle sampling during a timed region (nvidia-smi in the background, or NVML)."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
import time
from typing import Dict, List, Optional

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class NvmlClockSampler:
    """Samples SM clock, power and throttle reasons through NVML from a background thread (few-ms period),
    so that even a 100 ms timed region gets tens of samples.  Falls back to :class:`ClockSampler`."""

    def __init__(self, gpu_index: int = 0, period_ms: float = 5.0):
        self.gpu_index, self.period = gpu_index, period_ms / 1e3
        self.samples: List[Dict] = []
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self._h = None
        self.available = False
        try:
            import pynvml
            self._nv = pynvml
            pynvml.nvmlInit()
            self._h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self._max = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self.available = True
        except Exception:
            self.available = False

    def _loop(self):
        nv = self._nv
        R = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
        }
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while not self._stop.is_set():
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                pw = nv.nvmlDeviceGetPowerUsage(self._h) / 1e3
                mask = int(get_reasons(self._h))
                s = {"sm_mhz": mhz, "sm_max_mhz": self._max, "power_w": pw, "t": time.time()}
                for k, bit in R.items():
                    s[k] = bool(mask & bit)
                self.samples.append(s)
            except Exception:
                pass
            time.sleep(self.period)

    def start(self):
        if self.available:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def stop(self) -> Dict:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2)
        return self.summary()

    def summary(self) -> Dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        reasons = [k for k in ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
                   if any(s[k] for s in self.samples)]
        pmax = max(s["power_w"] for s in self.samples)
        loaded = [s for s in self.samples if s["power_w"] >= 0.6 * pmax] or self.samples
        return {"sm_mhz": statistics.median(s["sm_mhz"] for s in loaded), "sm_max_mhz": self._max, "reasons": reasons,
                "samples": len(self.samples), "power_w_max": round(pmax, 1), "source": "nvml"}


def make_sampler(gpu_index: int = 0):
    s = NvmlClockSampler(gpu_index)
    return s if s.available else ClockSampler(gpu_index, period_ms=50)


class ClockSampler:
    """Samples SM clocks and throttle reasons of one GPU every ``period_ms`` while active."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 200):
        self.gpu_index = gpu_index
        self.period_ms = period_ms
        self.samples: List[Dict] = []
        self._proc: Optional[subprocess.Popen] = None
        self._thread: Optional[threading.Thread] = None
        self._stop = threading.Event()
        self.available = shutil.which("nvidia-smi") is not None

    def _reader(self):
        assert self._proc is not None and self._proc.stdout is not None
        for line in self._proc.stdout:
            if self._stop.is_set():
                break
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9 or not parts[0].isdigit():
                continue
            try:
                self.samples.append({
                    "sm_mhz": float(parts[1].split()[0]), "sm_max_mhz": float(parts[2].split()[0]),
                    "power_w": float(parts[3].split()[0]) if parts[3][0].isdigit() else 0.0,
                    "hw_slowdown": parts[5].lower().startswith("active"),
                    "hw_thermal_slowdown": parts[6].lower().startswith("active"),
                    "sw_thermal_slowdown": parts[7].lower().startswith("active"),
                    "sw_power_cap": parts[8].lower().startswith("active"),
                    "t": time.time(),
                })
            except (ValueError, IndexError):
                continue

    def start(self):
        if not self.available:
            return self
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={_QUERY}", "--format=csv,noheader",
                 "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self._thread = threading.Thread(target=self._reader, daemon=True)
            self._thread.start()
        except OSError:
            self.available = False
        return self

    def stop(self) -> Dict:
        self._stop.set()
        if self._proc is not None:
            self._proc.terminate()       # exact PID we started
            try:
                self._proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self._proc.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)
        return self.summary()

    def summary(self) -> Dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        mx = max(s["sm_max_mhz"] for s in self.samples)
        loaded = [s for s in self.samples if s["power_w"] > 250.0] or self.samples
        reasons = [k for k in ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
                   if any(s[k] for s in self.samples)]
        return {"sm_mhz": statistics.median(s["sm_mhz"] for s in loaded), "sm_max_mhz": mx, "reasons": reasons,
                "samples": len(self.samples), "power_w_max": max(s["power_w"] for s in self.samples)}

    def __enter__(self):
        return self.start()

    def __exit__(self, *a):
        self.stop()
        return False
