"""nvidia-smi clock / throttle sampler for the timed region of a benchmark (B200_PROFILING.md)."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
from typing import Dict, List, Optional

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    def __init__(self, gpu_index: int = 0, period_ms: int = 100):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self._proc: Optional[subprocess.Popen] = None
        self._lines: List[str] = []
        self._thr: Optional[threading.Thread] = None

    def start(self):
        exe = shutil.which("nvidia-smi")
        if not exe:
            return self
        self._proc = subprocess.Popen(
            [exe, f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i", str(self.gpu_index),
             "-lms", str(self.period_ms)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)

        def pump():
            for line in self._proc.stdout:
                self._lines.append(line.strip())
        self._thr = threading.Thread(target=pump, daemon=True)
        self._thr.start()
        return self

    def wait_first_sample(self, timeout_s: float = 2.0) -> bool:
        import time
        t0 = time.time()
        while not self._lines and time.time() - t0 < timeout_s:
            time.sleep(0.01)
        return bool(self._lines)

    def mark(self) -> None:
        """Samples before this point (warm-up) are dropped from the summary."""
        self._skip = len(self._lines)

    def stop(self) -> Dict:
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self._proc.kill()
            if self._thr:
                self._thr.join(timeout=2)
        return self.summary()

    def summary(self) -> Dict:
        sm, mx, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        lines = self._lines[getattr(self, "_skip", 0):] or self._lines[-1:]
        for line in lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); mx.append(float(parts[2])); power.append(float(parts[3]))
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(power),
                "reasons": sorted(reasons), "samples": len(sm)}
