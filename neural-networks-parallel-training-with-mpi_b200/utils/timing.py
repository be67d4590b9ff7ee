"""Device-side timing helpers: CUDA events on the launching stream, max over ranks."""
from __future__ import annotations

import time
from contextlib import contextmanager
from typing import Dict, List

import torch


class PhaseTimer:
    """Accumulates per-phase durations.  On CUDA it records event pairs and resolves them lazily
    (no host sync inside the step); on CPU it uses perf_counter."""

    def __init__(self, device, enabled: bool = True):
        self.cuda = torch.device(device).type == "cuda"
        self.enabled = enabled
        self._pending: List = []
        self.totals_ms: Dict[str, float] = {}
        self.counts: Dict[str, int] = {}

    @contextmanager
    def phase(self, name: str):
        if not self.enabled:
            yield
            return
        if self.cuda:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            yield
            b.record()
            self._pending.append((name, a, b))
        else:
            t0 = time.perf_counter()
            yield
            self._add(name, (time.perf_counter() - t0) * 1e3)

    def _add(self, name, ms):
        self.totals_ms[name] = self.totals_ms.get(name, 0.0) + ms
        self.counts[name] = self.counts.get(name, 0) + 1

    def resolve(self) -> Dict[str, float]:
        if self.cuda and self._pending:
            torch.cuda.synchronize()
            for name, a, b in self._pending:
                self._add(name, a.elapsed_time(b))
            self._pending.clear()
        return dict(self.totals_ms)

    def mean_ms(self) -> Dict[str, float]:
        self.resolve()
        return {k: v / max(self.counts[k], 1) for k, v in self.totals_ms.items()}


def max_over_ranks(value: float, comm) -> float:
    vals = comm.allgather(float(value))
    return max(vals)
