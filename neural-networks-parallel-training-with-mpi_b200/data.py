"""Data layer: synthetic regression set, row shards, feature scaling, batches.

Reference behaviour reproduced (with its constants as defaults):

* ``make_regression(n_samples=16, n_features=2, noise=1, random_state=42)`` on root, glued to a
  ``[rows, features+1]`` float64 matrix (ref.py:72-74).
* Contiguous row shards, first ``residue`` ranks get one more row (ref.py:100-143) -
  ``parallel/sharding.py``.
* ``RegressionDataset(X, y, scale_data=True)``: standard-scales X only, with statistics of *the rows
  it is given* (ref.py:17-24).  Because the reference builds it after scattering (ref.py:145) the
  scaling is per shard; ``scale='global'`` / ``'none'`` are the alternatives.
* One batch per epoch = the whole shard (ref.py:146), float32 inputs, targets ``[B,1]``
  (ref.py:158-160).  The f64->f32 cast happens once here, not per step.

Large synthetic shapes (the wide / deep MLP configs of BASELINE.json) use ``data_source='randn'``:
every rank generates only its own rows from a counter-based seed, so nothing is scattered.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Iterator, Optional, Tuple

import numpy as np
import torch

from .parallel.comm import Comm
from .parallel.sharding import shard_counts, shard_displs


def make_regression_xy(n_samples=16, n_features=2, noise=1.0, seed=42) -> np.ndarray:
    """``[n_samples, n_features+1]`` float64, features then target (ref.py:72-74)."""
    from sklearn.datasets import make_regression

    X, y = make_regression(n_samples=n_samples, n_features=n_features, noise=noise, random_state=seed)
    return np.concatenate((X, y.reshape(-1, 1)), axis=1)


def make_randn_rows(start: int, stop: int, n_features: int, noise: float, seed: int,
                    n_outputs: int = 1, dtype=np.float32) -> np.ndarray:
    """Rows ``[start, stop)`` of a linear-teacher regression set.  Row ``i`` depends only on
    ``(seed, i)`` so any rank can build any row range and shards agree for every world size."""
    rows = stop - start
    w_rng = np.random.Generator(np.random.Philox(key=seed))
    coef = (w_rng.standard_normal((n_features, n_outputs)) / np.sqrt(n_features)).astype(np.float32)
    out = np.empty((rows, n_features + n_outputs), dtype=dtype)
    chunk = 4096
    for lo in range(0, rows, chunk):
        hi = min(rows, lo + chunk)
        # one Philox stream per chunk-aligned block of global rows
        g0 = start + lo
        blk, off = divmod(g0, chunk)
        parts, need, b = [], hi - lo, blk
        while need > 0:
            rng = np.random.Generator(np.random.Philox(key=seed + 1, counter=[0, 0, 0, b]))
            full = rng.standard_normal((chunk, n_features + n_outputs), dtype=np.float32)
            take = full[off:off + need]
            parts.append(take)
            need -= take.shape[0]
            off, b = 0, b + 1
        z = np.concatenate(parts, axis=0) if len(parts) > 1 else parts[0]
        x = z[:, :n_features]
        y = x @ coef + noise * z[:, n_features:]
        out[lo:hi, :n_features] = x
        out[lo:hi, n_features:] = y
    return out


def standard_scale(X: np.ndarray) -> np.ndarray:
    """What ``StandardScaler().fit_transform`` computes (population std; zero-variance columns are
    left unscaled), without the sklearn dependency on the hot path."""
    X = np.asarray(X, dtype=np.float64)
    mean = X.mean(axis=0)
    std = X.std(axis=0)
    std = np.where(std == 0.0, 1.0, std)
    return (X - mean) / std


class RegressionDataset(torch.utils.data.Dataset):
    """Map-style dataset with the reference's constructor (ref.py:12-30).  Tensors are accepted as
    well (the reference silently sets nothing in that case)."""

    def __init__(self, X, y, scale_data=True):
        if torch.is_tensor(X):
            X = X.detach().cpu().numpy()
        if torch.is_tensor(y):
            y = y.detach().cpu().numpy()
        X = np.asarray(X)
        y = np.asarray(y)
        if scale_data and len(X):
            X = standard_scale(X)
        self.X = torch.from_numpy(np.ascontiguousarray(X))
        self.y = torch.from_numpy(np.ascontiguousarray(y))

    def __len__(self):
        return len(self.X)

    def __getitem__(self, i):
        return self.X[i], self.y[i]


@dataclass
class Shard:
    """This rank's rows, ready for training: float32, targets ``[rows, n_outputs]``."""
    x: torch.Tensor
    y: torch.Tensor
    start: int
    rows_global: int
    counts: Tuple[int, ...]

    @property
    def rows(self) -> int:
        return int(self.x.shape[0])


def load_shard(cfg, comm: Comm) -> Shard:
    """Builds the dataset and returns this rank's shard (ref.py:66-146)."""
    n_out = cfg.n_outputs
    rows = cfg.n_samples
    val_rows = int(round(rows * cfg.val_frac)) if cfg.val_frac > 0 else 0
    train_rows = rows - val_rows
    counts = shard_counts(train_rows, comm.size)
    displs = shard_displs(counts)
    lo, hi = displs[comm.rank], displs[comm.rank] + counts[comm.rank]
    w = cfg.n_features + n_out

    if cfg.data_source == "sklearn":
        if n_out != 1:
            raise ValueError("data_source=sklearn produces a single target")
        if cfg.data_dist == "scatter":
            # reference protocol: root generates, shape is broadcast, rows are scattered
            full = make_regression_xy(rows, cfg.n_features, cfg.noise, cfg.data_seed) \
                if comm.rank == 0 else None
            h, w_b = comm.bcast(full.shape if comm.rank == 0 else None, root=0)   # ref.py:97
            assert (h, w_b) == (rows, w)
            if comm.size == 1:
                mine = full[:train_rows]
            else:
                mine = comm.scatterv_rows(full[:train_rows] if comm.rank == 0 else None,
                                          counts, w, root=0)                   # ref.py:108,138
            stats = None
            if cfg.scale == "global":
                if comm.rank == 0:
                    gx = full[:train_rows, :cfg.n_features]
                    stats = (gx.mean(axis=0), gx.std(axis=0))
                stats = comm.bcast(stats, root=0)
        else:
            full = make_regression_xy(rows, cfg.n_features, cfg.noise, cfg.data_seed)
            mine = full[lo:hi]
            gx = full[:train_rows, :cfg.n_features]
            stats = (gx.mean(axis=0), gx.std(axis=0))
        X, y = mine[:, :cfg.n_features], mine[:, cfg.n_features:]
        if cfg.scale == "per_shard" and len(X):
            X = standard_scale(X)                                               # ref.py:21-22,145
        elif cfg.scale == "global":
            mean, std = stats
            X = (X - mean) / np.where(std == 0.0, 1.0, std)
    elif cfg.data_source == "randn":
        mine = make_randn_rows(lo, hi, cfg.n_features, cfg.noise, cfg.data_seed, n_out)
        X, y = mine[:, :cfg.n_features], mine[:, cfg.n_features:]
        if cfg.scale == "per_shard" and len(X):
            X = standard_scale(X)
        elif cfg.scale == "global":
            pass  # randn features are already zero-mean / unit-variance by construction
    else:
        raise ValueError(f"unknown data source {cfg.data_source!r}")

    x_t = torch.from_numpy(np.ascontiguousarray(X)).float()                    # ref.py:159
    y_t = torch.from_numpy(np.ascontiguousarray(y)).float().reshape(len(X), n_out)   # ref.py:160
    return Shard(x_t, y_t, lo, train_rows, tuple(counts))


def load_validation(cfg) -> Optional[Tuple[torch.Tensor, torch.Tensor]]:
    """Held-out tail rows for the rank-0 evaluation the reference only sketches in dead code
    (ref.py:213-220, 227-236)."""
    if cfg.val_frac <= 0:
        return None
    rows = cfg.n_samples
    val_rows = int(round(rows * cfg.val_frac))
    if val_rows == 0:
        return None
    if cfg.data_source == "sklearn":
        full = make_regression_xy(rows, cfg.n_features, cfg.noise, cfg.data_seed)
        tail = full[rows - val_rows:]
        train_x = full[:rows - val_rows, :cfg.n_features]
        X = tail[:, :cfg.n_features]
        if cfg.scale != "none":
            std = train_x.std(axis=0)
            X = (X - train_x.mean(axis=0)) / np.where(std == 0.0, 1.0, std)
        y = tail[:, cfg.n_features:]
    else:
        tail = make_randn_rows(rows - val_rows, rows, cfg.n_features, cfg.noise, cfg.data_seed,
                               cfg.n_outputs)
        X, y = tail[:, :cfg.n_features], tail[:, cfg.n_features:]
    return (torch.from_numpy(np.ascontiguousarray(X)).float(),
            torch.from_numpy(np.ascontiguousarray(y)).float().reshape(len(X), cfg.n_outputs))


class PinnedBuffer:
    """Page-locked host memory for input batches, from the native allocator (``csrc/host_mem.cu``:
    huge-page advised + ``cudaHostRegister``; consistently full PCIe rate, unlike ``pin_memory()``
    whose buffers vary by ~20 % from one allocation to the next on this pool's boxes).  ``tensor`` is a
    view of it with the given shape and dtype.  Without a GPU it is plain host memory."""

    def __init__(self, shape, dtype=torch.float32):
        import ctypes as C
        self.shape, self.dtype = tuple(int(s) for s in shape), dtype
        n = 1
        for s in self.shape:
            n *= s
        nbytes = max(n * torch.empty((), dtype=dtype).element_size(), 1)
        self._ptr, self._mapped = None, 0
        if torch.cuda.is_available():
            from .ops import _lib
            lib = _lib.lib()
            lib.b200dp_host_alloc.argtypes = [C.c_longlong, C.POINTER(C.c_void_p), C.POINTER(C.c_longlong)]
            lib.b200dp_host_free.argtypes = [C.c_void_p, C.c_longlong]
            p, m = C.c_void_p(), C.c_longlong()
            _lib.check(lib.b200dp_host_alloc(nbytes, C.byref(p), C.byref(m)))
            self._ptr, self._mapped, self._lib = p.value, m.value, lib
            raw = (C.c_uint8 * nbytes).from_address(self._ptr)
            self.tensor = torch.frombuffer(raw, dtype=torch.uint8)[:nbytes].view(dtype)[:n].view(self.shape)
        else:
            self.tensor = torch.empty(self.shape, dtype=dtype)

    def close(self) -> None:
        if self._ptr is not None:
            ptr, self._ptr = self._ptr, None
            self.tensor = None
            self._lib.b200dp_host_free(ptr, self._mapped)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def pinned_like(t: torch.Tensor) -> PinnedBuffer:
    """A :class:`PinnedBuffer` holding a copy of ``t``."""
    buf = PinnedBuffer(t.shape, t.dtype)
    buf.tensor.copy_(t)
    return buf


class BatchPlan:
    """Yields one row index tensor (or ``None`` = the whole shard) per step of an epoch.

    Batch == shard (the reference, ref.py:146) gives one step with no permutation (SURVEY Q9:
    shuffling a full batch is a mathematical no-op).  Smaller batches draw a seeded permutation that
    is a pure function of ``(seed, epoch)`` - a resumed run continues the sequence exactly - and cut
    it into ``steps`` nearly equal batches.  Every gradient sync is a collective, so **every rank must
    run the same number of steps**: ``steps`` is derived from the LARGEST shard
    (``ceil(max_rows / batch)``), not from this rank's own row count (ref.py:146,155 iterate a
    per-rank DataLoader; with uneven shards that would give ranks different step counts and hang the
    collective).  A rank with fewer rows than steps yields empty batches for the remainder: it
    contributes a zero gradient and keeps the collective in step."""

    def __init__(self, rows: int, batch: Optional[int], shuffle: bool, seed: int,
                 max_rows: Optional[int] = None):
        self.rows, self.shuffle, self.seed = rows, shuffle, seed
        max_rows = rows if max_rows is None else max(int(max_rows), rows)
        self.full = not batch or batch >= max_rows
        self.batch = rows if self.full else batch
        self.steps = (1 if max_rows > 0 else 0) if self.full else -(-max_rows // batch)
        self.epoch_index = 0

    def steps_per_epoch(self) -> int:
        return self.steps

    def set_epoch(self, epoch: int) -> None:
        self.epoch_index = int(epoch)

    def epoch(self) -> Iterator[Optional[torch.Tensor]]:
        e = self.epoch_index
        self.epoch_index += 1
        if self.rows == 0:                 # the trainer runs this rank's `steps` empty steps itself
            return
        if self.full:
            if self.rows > 0:
                yield None
            return
        if self.shuffle:
            gen = torch.Generator().manual_seed((self.seed * 1000003 + e) & 0x7FFFFFFFFFFFFFFF)
            order = torch.randperm(self.rows, generator=gen)
        else:
            order = torch.arange(self.rows)
        lo = 0
        for n in shard_counts(self.rows, self.steps):        # first `rows % steps` batches get one more row
            yield order[lo:lo + n]
            lo += n
