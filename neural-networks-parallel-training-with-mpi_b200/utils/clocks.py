"""SM-clock / throttle-reason sampler for the timed region of a benchmark (B200_PROFILING.md).

Samples NVML from a background thread (default every 20 ms, so even a 60 ms timed region gets a few
samples); falls back to an ``nvidia-smi -lms`` child process when the NVML binding is missing.  The
summary carries the median SM clock, the maximum SM clock, peak power and which of the four
throttle reasons of the recipe were seen (``hw_slowdown``, ``hw_thermal_slowdown``,
``sw_thermal_slowdown``, ``sw_power_cap``)."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
import time
from typing import Dict, List, Optional, Tuple

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

# nvmlClocksEventReasons bits (nvml.h)
_REASON_BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}


def _nvml_handle(gpu_index: int):
    import pynvml
    import torch
    pynvml.nvmlInit()
    try:        # CUDA_VISIBLE_DEVICES may renumber: resolve through the UUID of the torch device
        uuid = torch.cuda.get_device_properties(gpu_index).uuid
        return pynvml, pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode())
    except Exception:
        return pynvml, pynvml.nvmlDeviceGetHandleByIndex(gpu_index)


class ClockSampler:
    def __init__(self, gpu_index: int = 0, period_ms: int = 20):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self._samples: List[Tuple[float, float, float, int]] = []     # (sm MHz, max MHz, W, reason bits)
        self._skip = 0
        self._stop = threading.Event()
        self._thr: Optional[threading.Thread] = None
        self._proc: Optional[subprocess.Popen] = None
        self.source = "none"

    # ---- NVML thread ----
    def _run_nvml(self, nv, h) -> None:
        try:
            mx = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
        except Exception:
            mx = 0.0
        reasons_fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons", None)
        period = self.period_ms / 1e3
        while not self._stop.is_set():
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    w = nv.nvmlDeviceGetPowerUsage(h) / 1e3
                except Exception:
                    w = 0.0
                bits = int(reasons_fn(h)) if reasons_fn else 0
                self._samples.append((sm, mx, w, bits))
            except Exception:
                pass
            self._stop.wait(period)

    # ---- nvidia-smi fallback ----
    def _run_smi(self) -> None:
        for line in self._proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm, mx, w = float(parts[1]), float(parts[2]), float(parts[3])
            except ValueError:
                continue
            bits = 0
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                                 parts[5:9]):
                if val.lower().startswith("active"):
                    bits |= _REASON_BITS[name]
            self._samples.append((sm, mx, w, bits))

    def start(self):
        try:
            nv, h = _nvml_handle(self.gpu_index)
            nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            self.source = "nvml"
            self._thr = threading.Thread(target=self._run_nvml, args=(nv, h), daemon=True)
            self._thr.start()
            return self
        except Exception:
            pass
        exe = shutil.which("nvidia-smi")
        if exe:
            self.source = "nvidia-smi"
            self._proc = subprocess.Popen(
                [exe, f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i", str(self.gpu_index),
                 "-lms", str(max(self.period_ms, 20))], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self._thr = threading.Thread(target=self._run_smi, daemon=True)
            self._thr.start()
        return self

    def wait_first_sample(self, timeout_s: float = 2.0) -> bool:
        t0 = time.time()
        while not self._samples and time.time() - t0 < timeout_s:
            time.sleep(0.005)
        return bool(self._samples)

    def mark(self) -> None:
        """Samples before this point (warm-up) are dropped from the summary."""
        self._skip = len(self._samples)

    def stop(self) -> Dict:
        self._stop.set()
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self._proc.kill()
        if self._thr:
            self._thr.join(timeout=2)
        return self.summary()

    def summary(self, first: Optional[int] = None, last: Optional[int] = None) -> Dict:
        lo = self._skip if first is None else first
        rows = self._samples[lo:last] or self._samples[-1:]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": self.source}
        bits = 0
        for r in rows:
            bits |= r[3]
        return {"sm_mhz": statistics.median(r[0] for r in rows), "sm_mhz_min": min(r[0] for r in rows),
                "sm_max_mhz": max(r[1] for r in rows), "power_w_max": max(r[2] for r in rows),
                "reasons": sorted(n for n, b in _REASON_BITS.items() if bits & b), "samples": len(rows),
                "period_ms": self.period_ms, "source": self.source}

    def position(self) -> int:
        return len(self._samples)
