"""Device timing + clock sampling helpers (B200_PROFILING.md timing hygiene)."""
from __future__ import annotations

import statistics
import subprocess
import threading
import time


class ClockSampler:
    """Samples ``nvidia-smi`` SM clocks and throttle reasons in a background thread."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_s: float = 0.2):
        self.gpu_index = gpu_index
        self.period_s = period_s
        self.samples: list[dict] = []
        self._stop = threading.Event()
        self._thr = None

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(
                    ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                     "-i", str(self.gpu_index)], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 8:
                    self.samples.append({"sm": float(f[1]), "sm_max": float(f[2]), "power": float(f[3]),
                                         "hw_slowdown": f[4], "hw_thermal": f[5], "sw_thermal": f[6],
                                         "sw_power_cap": f[7]})
            except Exception:
                pass
            self._stop.wait(self.period_s)

    def start(self):
        self._thr = threading.Thread(target=self._loop, daemon=True)
        self._thr.start()
        return self

    def stop(self) -> dict:
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        reasons = []
        for key, name in (("hw_slowdown", "hw_slowdown"), ("hw_thermal", "hw_thermal_slowdown"),
                          ("sw_thermal", "sw_thermal_slowdown"), ("sw_power_cap", "sw_power_cap")):
            if any(s[key].lower().startswith("active") for s in self.samples):
                reasons.append(name)
        return {"sm_mhz": statistics.median(s["sm"] for s in self.samples),
                "sm_max_mhz": max(s["sm_max"] for s in self.samples),
                "power_w_max": max(s["power"] for s in self.samples),
                "samples": len(self.samples), "reasons": reasons}


def wall() -> float:
    return time.perf_counter()
