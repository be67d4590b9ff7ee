"""Training driver: the reference's ``dist_train`` (ref.py:56-236) as a reusable ``Trainer``.

Flow: discover rank -> build/shard the data -> replicate the model from rank 0
-> per epoch: banner, one step per batch (local forward/MSE/backward, gradient sync fused with
the SGD-momentum update), print the local loss of the last batch.

Two compute backends:

* ``cuda``  - the product.  Hand-written sm_100a kernels from ``libb200dp.so`` (``ops/``): fused
              SIMT kernel for tiny nets, tcgen05/TMA GEMMs + fused MSE for real widths, fused
              peer-memory all-reduce + SGD.  Selected automatically when a GPU is present; it
              raises if the native library is missing (no silent fallback on a GPU box).
* ``torch`` - plain PyTorch ops (``ops/reference.py``) for boxes without a GPU (CPU plumbing
              config, parity tests, the gloo multi-process tests).
"""
from __future__ import annotations

import math
import time
from typing import Any, Dict, List, Optional

import torch

from . import checkpoint as ckpt
from .config import TrainConfig
from .data import BatchPlan, Shard, load_shard, load_validation
from .models.mlp import FlatMLP, MLPSpec
from .ops import reference as ref_ops
from .parallel.comm import Comm, SingleComm
from .parallel.grad_sync import make_host_sync
from .utils.logging import JsonlLogger, emit, epoch_banner, loss_line
from .utils.timing import PhaseTimer


class TorchCompute:
    """Local forward + loss + backward with PyTorch ops (CPU path / oracle)."""
    name = "torch"

    def __init__(self, flat: FlatMLP, cfg: TrainConfig):
        self.flat = flat

    def prepare_batch(self, x: torch.Tensor, y: torch.Tensor):
        return x, y

    def loss_and_grads(self, x, y) -> torch.Tensor:
        return ref_ops.mlp_loss_and_grads(self.flat, x, y)

    def predict(self, x) -> torch.Tensor:
        return ref_ops.mlp_forward(self.flat, x).to(torch.float32)


def resolve_backend(cfg: TrainConfig, device: torch.device) -> str:
    if cfg.backend != "auto":
        if cfg.backend == "cuda" and device.type != "cuda":
            raise RuntimeError("backend=cuda needs a GPU")
        return cfg.backend
    return "cuda" if device.type == "cuda" else "torch"


class Trainer:
    def __init__(self, cfg: TrainConfig, comm: Optional[Comm] = None,
                 device: Optional[torch.device] = None, shard: Optional[Shard] = None):
        self.cfg = cfg.validate()
        self.comm = comm if comm is not None else SingleComm()
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.backend = resolve_backend(cfg, self.device)
        self.rank, self.world = self.comm.rank, self.comm.size
        self.log = JsonlLogger(cfg.log_jsonl if self.rank == 0 else None)
        self.timer = PhaseTimer(self.device, enabled=bool(cfg.log_jsonl))
        self.history: List[Dict[str, Any]] = []
        self.start_epoch = 0

        # ---- data (ref.py:66-146) ----
        self.shard = shard if shard is not None else load_shard(cfg, self.comm)
        self.counts = list(self.shard.counts)
        self.val = load_validation(cfg) if (cfg.val_frac > 0 and self.rank == 0) else None

        # ---- model replicas (ref.py:83-88) + sync strategy ----
        self.spec = MLPSpec(cfg.layer_sizes())
        if self.backend == "cuda":
            from .ops.cuda_backend import build_cuda_runtime   # raises if libb200dp.so is missing
            rt = build_cuda_runtime(self)
            self.flat, self.compute, self.sync = rt.flat, rt.compute, rt.sync
            self._runtime = rt
        else:
            self.flat = FlatMLP(self.spec, self.device, cfg.dtype)
            self.compute = TorchCompute(self.flat, cfg)
            self.sync = make_host_sync(cfg.sync, self.comm, self.counts, cfg.grad_weighting)
            self._runtime = None
        self._replicate_initial_state()
        if cfg.resume:
            self.load_checkpoint(cfg.resume)

        batch = cfg.batch_size if cfg.use_batch_size else None
        self.plan = BatchPlan(self.shard.rows, batch, cfg.shuffle, seed=cfg.seed + 1 + self.rank,
                              max_rows=max(self.counts))
        self.plan.set_epoch(self.start_epoch)
        self.x, self.y = self.compute.prepare_batch(self.shard.x.to(self.device),
                                                    self.shard.y.to(self.device))
        if self.world > 1 and self.device.type == "cuda":
            # the first fused sync kernel spins on its peers with a bounded budget (watchdog): start the
            # ranks together so set-up skew (data generation, allocation) cannot eat into that budget
            torch.cuda.synchronize()
            self.comm.barrier()

    # ------------------------------------------------------------------ init
    def _replicate_initial_state(self):
        """Root initialises with ``manual_seed(seed)`` and everybody adopts root's parameters
        (ref.py:66-69, 84-88)."""
        cfg = self.cfg
        if cfg.init == "device" and self.device.type == "cuda":
            # same seed on every rank -> identical replicas without moving any bytes
            self.flat.init_on_device(cfg.seed)
            return
        if self.rank == 0:
            self.flat.init_like_reference(cfg.seed)
        if self.world > 1:
            if self._runtime is not None and self._runtime.device_broadcast is not None:
                self._runtime.device_broadcast(self.flat.params)
                self.flat.sync_shadow()
            else:
                sd = self.comm.bcast(self.flat.state_dict() if self.rank == 0 else None, root=0)
                self.flat.load_state_dict(sd)

    # ------------------------------------------------------------------ step
    def step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """One synchronous data-parallel step (ref.py:164-211).  Returns the local pre-update loss
        as a 0-dim device tensor (no host sync)."""
        begin = getattr(self.sync, "begin_step", None)
        if begin is not None:                      # fused sync: buckets are queued from inside backward
            begin(self.cfg.lr, self.cfg.momentum)
        nvtx = self.device.type == "cuda" and self.cfg.log_jsonl is not None
        if nvtx:
            torch.cuda.nvtx.range_push("fwd_bwd")
        with self.timer.phase("fwd_bwd"):
            loss = self.compute.loss_and_grads(x, y)
        if nvtx:
            torch.cuda.nvtx.range_pop()
            torch.cuda.nvtx.range_push("sync_update")
        with self.timer.phase("sync_update"):
            self.sync.sync_and_step(self.flat, self.cfg.lr, self.cfg.momentum)
        if nvtx:
            torch.cuda.nvtx.range_pop()
        return loss

    def _empty_step(self) -> None:
        """A step of a rank that has no rows for it: zero gradients into the same sync."""
        begin = getattr(self.sync, "begin_step", None)
        if begin is not None:
            begin(self.cfg.lr, self.cfg.momentum)
        self.flat.grads.zero_()
        self.sync.sync_and_step(self.flat, self.cfg.lr, self.cfg.momentum)

    def _minibatch(self, idx: torch.Tensor):
        """Rows ``idx`` of the resident shard (ref.py:146,155: DataLoader shuffle + collate)."""
        idx = idx.to(self.x.device)
        if self.backend == "cuda":
            from .ops import kernels as K
            return K.gather_rows(self.x, idx), self.y.index_select(0, idx)
        return self.x[idx], self.y[idx]

    # ------------------------------------------------------------------ CUDA graph
    def capture_full_shard_step(self) -> bool:
        """Captures one full-shard step (all of this step's kernels, both streams) into a CUDA graph
        so that launch-bound models - the reference's 13-parameter net is two ~3 us kernels - replay
        with a single driver call.  Only for full-shard batches (static input buffers)."""
        if self.device.type != "cuda" or self.shard.rows == 0 or getattr(self, "_graph", None) is not None:
            return getattr(self, "_graph", None) is not None
        if self.plan.batch < self.shard.rows:
            return False
        flat = self.flat
        keep = (flat.params.clone(), flat.momentum.clone(), flat.step_count,
                flat.shadow.clone() if flat.shadow is not None else None)
        torch.cuda.synchronize()
        self.comm.barrier()
        self.step(self.x, self.y)                       # warm-up: lazy one-time set-up happens here
        torch.cuda.synchronize()
        self.comm.barrier()
        flat.params.copy_(keep[0]); flat.momentum.copy_(keep[1]); flat.step_count = keep[2]
        if keep[3] is not None:
            flat.shadow.copy_(keep[3])
        torch.cuda.synchronize()
        self.comm.barrier()
        from .ops import _lib as _native
        g = torch.cuda.CUDAGraph()
        n0 = _native.launches
        with torch.cuda.graph(g):
            self._graph_loss = self.step(self.x, self.y)
        self._graph_launches = _native.launches - n0     # kernels of ours inside one replay
        flat.step_count = keep[2]
        self._graph = g
        return True

    def graph_step(self) -> torch.Tensor:
        self._graph.replay()
        self.flat.step_count += 1
        return self._graph_loss

    def train(self) -> List[Dict[str, Any]]:
        cfg = self.cfg
        use_graph = cfg.cuda_graph and self.capture_full_shard_step()
        for epoch in range(self.start_epoch, cfg.nepochs):
            emit(epoch_banner(epoch), cfg.quiet)                                # ref.py:152
            t0 = time.perf_counter()
            loss = None
            self.plan.set_epoch(epoch)
            for idx in self.plan.epoch():                                       # ref.py:155
                if idx is not None and idx.numel() == 0:
                    # this rank has run out of rows while a larger shard still has a batch: a zero
                    # gradient keeps the collective in step (every rank syncs once per step)
                    self._empty_step()
                    continue
                xb, yb = (self.x, self.y) if idx is None else self._minibatch(idx)
                loss = self.graph_step() if (use_graph and idx is None) else self.step(xb, yb)
            if self.shard.rows == 0:                                            # rank owns no rows at all
                for _ in range(self.plan.steps_per_epoch()):
                    self._empty_step()
            loss_val = float("nan") if loss is None else float(loss.item())     # ref.py:224
            check = getattr(self.sync, "check_errors", None)
            if check is not None:
                check()                                                         # watchdog
            emit(loss_line(self.rank, loss_val), cfg.quiet)
            rec = {"epoch": epoch, "rank": self.rank, "loss": loss_val,
                   "epoch_s": time.perf_counter() - t0}
            if cfg.global_loss or cfg.log_jsonl:
                rec["global_loss"] = self._global_loss(loss_val)
                if cfg.global_loss and self.rank == 0:
                    emit(f"global loss (sample-weighted): {rec['global_loss']}", cfg.quiet)
            if self.val is not None:
                rec["val_loss"] = self.evaluate(*self.val)
                emit(f"validation loss (rank 0): {rec['val_loss']}", cfg.quiet)
            if cfg.check_replicas_every and (epoch + 1) % cfg.check_replicas_every == 0:
                self.assert_replicas_identical()
            if cfg.checkpoint and cfg.checkpoint_every and (epoch + 1) % cfg.checkpoint_every == 0:
                self.save_checkpoint(cfg.checkpoint, epoch + 1)
            self.history.append(rec)
            if self.rank == 0 and cfg.log_jsonl:
                self.log.log(**rec, phases_ms=self.timer.mean_ms(),
                             samples_per_s=sum(self.counts) / max(rec["epoch_s"], 1e-12))
        if cfg.checkpoint:
            self.save_checkpoint(cfg.checkpoint, cfg.nepochs)
        return self.history

    # ------------------------------------------------------------------ host-fed training
    def fit_host_batches(self, batches, steps: int, on_loss=None) -> List[float]:
        """Trains ``steps`` steps on batches that live in (pinned) HOST memory - the end-to-end path
        a user with a real dataset takes.  ``batches`` yields ``(x_host, y_host)``; every step's
        inputs cross PCIe in that step (prefetched one step ahead on a copy stream into a double
        buffer) and every step's loss is copied back to pinned host memory and read by the host
        (one step behind the GPU, so the queue never drains).  Returns the per-step losses."""
        if self.device.type != "cuda":
            out = []
            for _, (xh, yh) in zip(range(steps), batches):
                xb, yb = self.compute.prepare_batch(xh.to(self.device), yh.to(self.device))
                out.append(float(self.step(xb, yb).item()))
            return out
        dev = self.device
        main = torch.cuda.current_stream()
        copy_stream = torch.cuda.Stream(device=dev)
        it = iter(batches)
        slots = [None, None]
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        consumed = [torch.cuda.Event(), torch.cuda.Event()]
        loss_host = torch.zeros(max(steps, 1), dtype=torch.float32).pin_memory()
        loss_evt = [torch.cuda.Event() for _ in range(max(steps, 1))]
        self.h2d_bytes_per_step = 0
        self.d2h_bytes_per_step = 4

        def stage(i):
            xh, yh = next(it)
            k = i & 1
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[k])               # the step that used this slot is done
                if slots[k] is None or slots[k][0].shape != xh.shape or slots[k][0].dtype != xh.dtype:
                    slots[k] = (torch.empty(xh.shape, dtype=xh.dtype, device=dev),
                                torch.empty(yh.shape, dtype=yh.dtype, device=dev))
                slots[k][0].copy_(xh, non_blocking=True)
                slots[k][1].copy_(yh, non_blocking=True)
                ready[k].record(copy_stream)
            self.h2d_bytes_per_step = xh.numel() * xh.element_size() + yh.numel() * yh.element_size()

        losses: List[float] = []
        if steps > 0:
            stage(0)
        for i in range(steps):
            k = i & 1
            if i + 1 < steps:
                stage(i + 1)
            main.wait_event(ready[k])
            xd, yd = slots[k]
            xb, yb = (xd, yd) if (xd.dtype == self.x.dtype and yd.dtype == self.y.dtype
                                  and xd.shape[1] == self.x.shape[1]) else self.compute.prepare_batch(xd, yd)
            loss = self.step(xb, yb)
            consumed[k].record(main)
            loss_host[i:i + 1].copy_(loss.reshape(1), non_blocking=True)
            loss_evt[i].record(main)
            if i >= 1:                                            # read step i-1's loss on the host
                loss_evt[i - 1].synchronize()
                losses.append(float(loss_host[i - 1]))
                if on_loss is not None:
                    on_loss(i - 1, losses[-1])
        if steps > 0:
            loss_evt[steps - 1].synchronize()
            losses.append(float(loss_host[steps - 1]))
            if on_loss is not None:
                on_loss(steps - 1, losses[-1])
        return losses

    # ------------------------------------------------------------------ extras
    def _global_loss(self, local: float) -> float:
        pairs = self.comm.allgather((local, self.shard.rows))
        tot = sum(n for _, n in pairs)
        return sum(l * n for l, n in pairs if n > 0 and not math.isnan(l)) / max(tot, 1)

    @torch.no_grad()
    def evaluate(self, x: torch.Tensor, y: torch.Tensor) -> float:
        """MSE on held-out rows (the reference's dead validation/test stubs, ref.py:213-236)."""
        pred = self.compute.predict(x.to(self.device))
        return float(((pred - y.to(self.device)) ** 2).mean().item())

    @torch.no_grad()
    def predict(self, x: torch.Tensor) -> torch.Tensor:
        return self.compute.predict(x.to(self.device))

    def assert_replicas_identical(self) -> None:
        """Debug-mode race detector: every rank must hold the same parameter bytes."""
        prints = self.comm.allgather(self.flat.fingerprint())
        if len(set(prints)) != 1:
            raise RuntimeError(f"replicas diverged: parameter fingerprints {prints}")

    def state_dict(self):
        """Full fp32 ``state_dict`` (reference key names) valid on every rank: collects the master
        slices first when the optimizer state is sharded across ranks (bf16 two-shot / NVLS)."""
        if self._runtime is not None:
            self._runtime.gather_optimizer_state()
        return self.flat.state_dict()

    def save_checkpoint(self, path: str, epoch: int) -> None:
        check = getattr(self.sync, "check_errors", None)
        if check is not None:
            check()                              # never persist parameters a timed-out sync left behind
        if self._runtime is not None:
            self._runtime.gather_optimizer_state()
        if self.rank == 0:
            ckpt.save_checkpoint(path, self.flat, epoch, self.cfg.asdict())
        self.comm.barrier()

    def load_checkpoint(self, path: str) -> None:
        payload = ckpt.load_checkpoint(path, self.flat)
        self.start_epoch = int(payload.get("epoch", 0))

    def close(self):
        self.log.close()
        self.sync.close()
        if self._runtime is not None:
            self._runtime.close()


def dist_train(args, comm: Optional[Comm] = None, device=None) -> Trainer:
    """Schedule a distributed training job - the reference's entry point (ref.py:56).  ``args`` may
    be the reference's own argparse Namespace (lr, momentum, batch_size, nepochs) or a
    ``TrainConfig``."""
    from .config import config_from_args
    from .launch import bind_device, discover, init_process_group
    from .parallel.comm import world_comm

    cfg = args if isinstance(args, TrainConfig) else config_from_args(args)
    if comm is None:
        info = discover()
        dev = bind_device(info) if device is None else torch.device(device)
        if info.world_size > 1:
            init_process_group(info)
        comm = world_comm()
    else:
        dev = torch.device(device) if device is not None else torch.device("cpu")
    trainer = Trainer(cfg, comm, dev)
    trainer.train()
    return trainer
