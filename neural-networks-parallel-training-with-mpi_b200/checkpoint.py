"""Checkpoint / resume.

The reference never writes a checkpoint; it only uses ``state_dict()`` for the initial broadcast
(ref.py:87-88).  BASELINE.json asks for "a torch state_dict checkpoint of the
replicated MLP", so: rank 0 writes (replicas are identical, one copy suffices), every rank loads.

File layout (``torch.save`` of a dict):
    model      OrderedDict with the reference's keys ``layers.{0,2,...}.{weight,bias}``, fp32 -
               loads straight into the reference's ``MLP`` (ref.py:41-45) / our ``models.MLP``.
    momentum   same keys, the SGD momentum buffers (needed for an exact resume).
    epoch, step_count, sizes, config, format version.

No RNG state is stored: nothing in training draws from a stateful generator.  The initial weights
come from ``manual_seed(seed)`` once, and the mini-batch permutation of epoch ``e`` on rank ``r`` is a
pure function of ``(seed, r, e)`` (``data.BatchPlan``), so a run resumed at epoch ``e`` continues the
exact sequence of the uninterrupted run on every rank.
"""
from __future__ import annotations

import os
import tempfile
from typing import Any, Dict

import torch

FORMAT_VERSION = 1


def save_checkpoint(path: str, flat, epoch: int, config: Dict[str, Any], extra: Dict[str, Any] = None):
    payload = {
        "format": FORMAT_VERSION,
        "model": flat.state_dict(),
        "momentum": flat.momentum_state(),
        "epoch": int(epoch),
        "step_count": int(flat.step_count),
        "sizes": list(flat.spec.logical_sizes),
        "config": dict(config),
    }
    if extra:
        payload["extra"] = extra
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)
    fd, tmp = tempfile.mkstemp(dir=d, prefix="ckpt-tmp-", suffix=".pt")
    os.close(fd)
    try:
        torch.save(payload, tmp)
        os.replace(tmp, path)          # atomic: a crash never leaves a torn file at `path`
    finally:
        if os.path.exists(tmp):
            os.remove(tmp)


def load_checkpoint(path: str, flat, strict_sizes: bool = True) -> Dict[str, Any]:
    payload = torch.load(path, map_location="cpu", weights_only=True)   # tensors and plain containers only
    if isinstance(payload, dict) and "model" not in payload:
        payload = {"model": payload}                     # a bare state_dict (e.g. from the reference's MLP)
    if strict_sizes and "sizes" in payload and tuple(payload["sizes"]) != tuple(flat.spec.logical_sizes):
        raise ValueError(f"checkpoint is for sizes {payload['sizes']}, model has {list(flat.spec.logical_sizes)}")
    flat.load_state_dict(payload["model"])
    if "momentum" in payload:
        flat.load_momentum_state(payload["momentum"])
    flat.step_count = int(payload.get("step_count", 0))
    return payload
