"""Row sharding of the dataset across ranks.

Intended semantics of ref.py:100-143: contiguous row blocks in rank order; when the row count is
not a multiple of the world size the first ``residue`` ranks own one extra row (ref.py:117).  The
reference keeps its counts in an int8 array broadcast as MPI.INT (ref.py:118,128,133),
which breaks for most world sizes; counts here are plain Python ints / int64 and every rank
computes them locally (they are a pure function of ``(rows, world)``), so no broadcast is needed.
"""
from __future__ import annotations

from typing import List, Tuple


def shard_counts(rows: int, world: int) -> List[int]:
    if world < 1:
        raise ValueError("world must be >= 1")
    if rows < 0:
        raise ValueError("rows must be >= 0")
    base, residue = divmod(rows, world)
    return [base + 1 if p < residue else base for p in range(world)]


def shard_displs(counts: List[int]) -> List[int]:
    out, acc = [], 0
    for c in counts:
        out.append(acc)
        acc += c
    return out


def shard_range(rows: int, world: int, rank: int) -> Tuple[int, int]:
    """[start, stop) of the rows owned by ``rank``."""
    counts = shard_counts(rows, world)
    start = shard_displs(counts)[rank]
    return start, start + counts[rank]


def element_counts(rows: int, cols: int, world: int) -> Tuple[List[int], List[int]]:
    """Scatterv-style element counts and displacements of the flattened ``[rows, cols]`` matrix
    (what ref.py:117-122 builds on root)."""
    counts = [c * cols for c in shard_counts(rows, world)]
    return counts, shard_displs(counts)


def grad_weights(rows: int, world: int, mode: str = "uniform") -> List[float]:
    """Per-rank factor applied to each rank's mean-gradient before the sum.

    ``uniform`` is the reference (ref.py:197 divides the plain sum by ``nprocs``, also on uneven
    shards); ``samples`` weights by shard size, which yields the gradient of the
    global-batch mean loss.
    """
    counts = shard_counts(rows, world)
    if mode == "uniform":
        return [1.0 / world] * world
    if mode == "samples":
        total = float(sum(counts))
        return [c / total for c in counts]
    raise ValueError(f"unknown grad weighting {mode!r}")
