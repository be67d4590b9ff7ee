"""Gradient synchronisation + optimizer step strategies.

Every strategy honours the same contract:

    g_avg = (sum_r w_r * g_r) / D        uniform: w_r = 1, D = N (ref.py:193-197)
                                         samples: w_r = n_r, D = sum n_r
    buf   = g_avg            (first step)      torch SGD semantics, ref.py:91,211
    buf   = mu*buf + g_avg   (later)
    p    -= lr*buf

and leaves all replicas bitwise identical.

* ``GatherRootSync``  - the reference's own protocol: pickle-gather to rank 0, Python-loop average
                        in rank order, N-1 sequential sends (ref.py:185-203).  Kept as the
                        measured baseline and as the CPU parity path.
* ``AllReduceSync``   - ``torch.distributed.all_reduce`` (NCCL on GPUs: the "library" baseline).
* ``FusedSync``       - the product: one sm_100a kernel per bucket over NVLink peer memory
                        (``ops/allreduce_sgd.py``): one-shot, two-shot or NVLS multimem.
"""
from __future__ import annotations

from typing import Sequence

import torch

from ..ops import reference as ref_ops
from .comm import Comm


class GradSync:
    name = "base"

    def __init__(self, comm: Comm, weights: Sequence[float], divisor: float):
        self.comm, self.weights, self.divisor = comm, list(weights), float(divisor)

    def sync_and_step(self, flat, lr: float, mu: float) -> None:
        raise NotImplementedError

    def close(self) -> None:
        pass

    # shared tail: replicated SGD on the averaged gradient (ref.py:206-211)
    def _apply(self, flat, g_avg: torch.Tensor, lr: float, mu: float) -> None:
        ref_ops.sgd_momentum_(flat.params, g_avg, flat.momentum, lr, mu, flat.step_count == 0)
        flat.sync_shadow()
        flat.step_count += 1


class LocalSync(GradSync):
    """World of one: the average of one gradient is that gradient (``x/1`` is exact)."""
    name = "local"

    def sync_and_step(self, flat, lr, mu):
        self._apply(flat, flat.grads, lr, mu)


class GatherRootSync(GradSync):
    name = "gather_root"

    def sync_and_step(self, flat, lr, mu):
        comm = self.comm
        mine = flat.grads.detach().to("cpu")
        grad_list = comm.gather(mine, root=0)                                  # ref.py:185
        if comm.rank == 0:
            avg = ref_ops.reduce_grads(grad_list, self.weights, self.divisor)  # ref.py:188-197
            for k in range(1, comm.size):                                      # ref.py:199
                comm.send(avg, k)
        else:
            avg = comm.recv(0)                                                 # ref.py:203
        self._apply(flat, avg.to(flat.params.device), lr, mu)                  # ref.py:206-211


class AllReduceSync(GradSync):
    name = "allreduce"

    def sync_and_step(self, flat, lr, mu):
        w = self.weights[self.comm.rank]
        g = flat.grads.to(torch.float32)
        if w != 1.0:
            g = g * w
        elif g.data_ptr() == flat.grads.data_ptr():
            g = g.clone()
        self.comm.allreduce_sum_(g)
        g /= self.divisor
        self._apply(flat, g, lr, mu)


def weights_and_divisor(counts: Sequence[int], mode: str):
    """Integer-valued per-rank weights and the divisor, see module docstring."""
    n = len(counts)
    if mode == "uniform":
        return [1.0] * n, float(n)
    if mode == "samples":
        return [float(c) for c in counts], float(sum(counts))
    raise ValueError(f"unknown grad weighting {mode!r}")


def make_host_sync(name: str, comm: Comm, counts: Sequence[int], weighting: str) -> GradSync:
    w, d = weights_and_divisor(counts, weighting)
    if comm.size == 1:
        return LocalSync(comm, w, d)
    if name in ("gather_root", "auto"):
        return GatherRootSync(comm, w, d)
    if name == "allreduce":
        return AllReduceSync(comm, w, d)
    raise ValueError(f"sync mode {name!r} needs the CUDA backend")
