#!/usr/bin/env python
"""Headline benchmark (BASELINE.json): MLP training samples/sec on N B200s, device-timed, max over
ranks.  Prints ONE JSON line on rank 0.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 20 --warmup 5

--config wide (default)  BASELINE.json configs[2]: MLP 12288 -> 4x4096 -> 1 (100.7 M parameters), bf16,
                         8192 rows per GPU (weak scaling), synthetic linear-teacher data, random init.
--config deep            configs[3]: 4096 -> 16x8192 -> 1 (1.04 B parameters), bf16, 8192 rows per GPU.
--config ref             configs[0-1]: the reference's own 2-3-1 MLP on make_regression(16, 2) (strong
                         scaling: 16 rows in total, as hard-coded at ref.py:72).
--config ref_uneven      configs[1]: the same net on 17 rows (uneven shards, sample-weighted gradients).

--impl ours (default)    this framework: kernel-only `value` (inputs resident on the device) and the
                         end-to-end number `e2e` through Trainer.fit_host_batches (pinned host batches,
                         H2D every step, loss read back every step).
--impl torch             NCCL all-reduce + cuBLAS (torch.nn / autograd / torch.optim.SGD) build of the
                         same MLP: the library baseline.
--impl reference         the UNMODIFIED reference script from baseline/_ref, run through its own
                         dist_train(args) over the gloo mpi4py shim (real MPI is not on this image).
                         It hard-codes its model and data (ref.py:41-45,72), so it can only run
                         --config ref; for the other configs it reports `unavailable`.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    "wide": dict(n_features=12288, hidden=[4096] * 4, n_outputs=1, rows_per_gpu=8192, dtype="bf16",
                 scaling="weak", name="wide MLP 12288-4x4096-1 (100.7M params)"),
    "deep": dict(n_features=4096, hidden=[8192] * 16, n_outputs=1, rows_per_gpu=8192, dtype="bf16",
                 scaling="weak", name="deep MLP 4096-16x8192-1 (1.04B params)"),
    "ref": dict(n_features=2, hidden=[3], n_outputs=1, rows_total=16, dtype="fp32", scaling="strong",
                name="reference MLP 2-3-1 on make_regression(16,2)"),
    "ref_uneven": dict(n_features=2, hidden=[3], n_outputs=1, rows_total=17, dtype="fp32", scaling="strong",
                       name="reference MLP 2-3-1 on make_regression(17,2): rows not a multiple of the world size"),
}


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", choices=("ours", "torch", "reference"), default="ours")
    p.add_argument("--config", choices=tuple(CONFIGS), default="wide")
    p.add_argument("--sync", default="auto")
    p.add_argument("--rows-per-gpu", type=int, default=None)
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-overlap", action="store_true")
    p.add_argument("--bucket-mb", type=float, default=0.0)
    p.add_argument("--cuda-graph", choices=("auto", "on", "off"), default="auto",
                   help="replay the step from a CUDA graph (auto: on for the launch-bound ref config)")
    return p.parse_args()


def dist_setup(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if torch.cuda.is_available():
        torch.cuda.set_device(local % torch.cuda.device_count())
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = "cpu:gloo,cuda:nccl" if (torch.cuda.is_available() and args.impl != "reference") else "gloo"
        kw = {}
        if "nccl" in backend:
            kw["device_id"] = torch.device("cuda", torch.cuda.current_device())
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, world, local


def max_over_ranks(x: float, world: int) -> float:
    if world == 1:
        return x
    import torch.distributed as dist
    out = [None] * world
    dist.all_gather_object(out, float(x))
    return max(out)


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def emit(rank, rec):
    if rank == 0:
        print(json.dumps(rec))
        sys.stdout.flush()


# ------------------------------------------------------------------------------------------- ours
def make_trainer(args, cfgd, rank, world):
    import torch
    import b200dp  # noqa: F401
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    from b200dp.parallel.comm import world_comm
    if args.config in ("ref", "ref_uneven"):
        cfg = TrainConfig(quiet=True, nepochs=0, sync=args.sync, data_dist="local",
                          n_samples=cfgd["rows_total"],
                          grad_weighting="samples" if args.config == "ref_uneven" else "uniform")
    else:
        rows = args.rows_per_gpu or cfgd["rows_per_gpu"]
        cfg = TrainConfig(quiet=True, nepochs=0, n_samples=rows * world, n_features=cfgd["n_features"],
                          hidden=cfgd["hidden"], n_outputs=cfgd["n_outputs"], dtype="bf16", sync=args.sync,
                          data_source="randn", data_dist="local", scale="none", init="device", lr=1e-4,
                          noise=0.1, overlap=not args.no_overlap, bucket_mb=args.bucket_mb)
    return Trainer(cfg, world_comm(), torch.device("cuda", torch.cuda.current_device()))


def run_ours(args, rank, world):
    import torch
    from b200dp.ops import _lib
    from b200dp.utils.clocks import ClockSampler
    cfgd = CONFIGS[args.config]
    t = make_trainer(args, cfgd, rank, world)
    rows_local = t.shard.rows
    global_batch = sum(t.counts)
    K, W = args.steps, max(args.warmup, 3)

    # ---- kernel-only: inputs resident on the device -------------------------------------------
    graphed = (args.cuda_graph == "on" or (args.cuda_graph == "auto" and args.config.startswith("ref"))) \
        and t.capture_full_shard_step()
    step = t.graph_step if graphed else (lambda: t.step(t.x, t.y))
    # nvidia-smi needs ~100 ms to come up: start it before the warm-up and wait for its first line, so
    # that it is sampling (every 50 ms) while the timed region runs
    sampler = ClockSampler(torch.cuda.current_device(), period_ms=100).start() if rank == 0 else None
    torch.cuda.synchronize()
    barrier(world)                       # ranks enter the first cross-GPU kernel together
    for _ in range(W):
        step()
    torch.cuda.synchronize()
    if sampler is not None:
        sampler.wait_first_sample(2.0)
        sampler.mark()
    barrier(world)
    l0 = _lib.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(K):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    barrier(world)
    launches = (_lib.launches - l0) if not graphed else K * t._graph_launches
    ms = max_over_ranks(e0.elapsed_time(e1), world)
    clocks = sampler.stop() if sampler else {}
    chk = getattr(t.sync, "check_errors", None)
    if chk:
        chk()
    ms_per_step = ms / K
    value = global_batch / (ms_per_step * 1e-3)
    final_loss = float(loss.item())

    # ---- end to end: pinned host batches, H2D every step, loss read back every step -----------
    e2e = None
    if not args.no_e2e:
        xh = [t.shard.x.to(t.x.dtype).pin_memory(), t.shard.x.to(t.x.dtype).clone().pin_memory()]
        yh = [t.shard.y.pin_memory(), t.shard.y.clone().pin_memory()]

        def batches():
            i = 0
            while True:
                yield xh[i & 1], yh[i & 1]
                i += 1
        torch.cuda.synchronize()
        barrier(world)                   # pinning the host batches takes a rank-dependent time
        t.fit_host_batches(batches(), W)
        torch.cuda.synchronize()
        barrier(world)
        torch.cuda.synchronize()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        a0.record()
        t.fit_host_batches(batches(), K)
        a1.record()
        torch.cuda.synchronize()
        wall_ms = (time.perf_counter() - w0) * 1e3
        barrier(world)
        ems = max_over_ranks(max(a0.elapsed_time(a1), wall_ms), world) / K
        e2e = {"value": global_batch / (ems * 1e-3), "unit": "samples/s", "ms_per_step": ems,
               "h2d_bytes_per_step": int(t.h2d_bytes_per_step), "d2h_bytes_per_step": int(t.d2h_bytes_per_step)}

    spec = t.spec
    flops = 6.0 * rows_local * sum(spec.sizes[i] * spec.sizes[i + 1] for i in range(spec.num_layers))
    rec = {
        "metric": "MLP train samples/sec (device-timed, max over ranks)", "value": value, "unit": "samples/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": cfgd["scaling"], "vs_baseline": None, "dtype": cfgd["dtype"], "data": "synthetic",
        "impl": "ours",
        "config": {"model": cfgd["name"], "global_batch": global_batch, "seq_len": None,
                   "parallelism": f"dp{world}", "sync": getattr(t._runtime, "sync_name", t.sync.name),
                   "rows_per_gpu": rows_local, "params": spec.num_params, "cuda_graph": bool(graphed),
                   "l2": "inputs+weights larger than L2 (no flush needed)" if not args.config.startswith("ref")
                   else "working set is 13 parameters: cache-resident by nature"},
        "clocks": {"sm_mhz": clocks.get("sm_mhz"), "sm_max_mhz": clocks.get("sm_max_mhz"),
                   "reasons": clocks.get("reasons", [])},
        "e2e": e2e, "gpu_launches": launches,
        "tflops_per_gpu": flops / (ms_per_step * 1e-3) / 1e12, "final_loss": final_loss,
    }
    emit(rank, rec)
    t.close()


# ------------------------------------------------------------------------------------------ torch
def run_torch(args, rank, world):
    """NCCL all-reduce + cuBLAS + torch.optim.SGD build of the same MLP (library baseline)."""
    import torch
    import torch.distributed as dist
    from torch import nn
    cfgd = CONFIGS[args.config]
    dev = torch.device("cuda", torch.cuda.current_device())
    sizes = [cfgd["n_features"], *cfgd["hidden"], cfgd["n_outputs"]]
    if args.config.startswith("ref"):
        rows_total = cfgd["rows_total"]
        rows = rows_total // world + (1 if rank < rows_total % world else 0)
        cdt = torch.float32
    else:
        rows = args.rows_per_gpu or cfgd["rows_per_gpu"]
        rows_total = rows * world
        cdt = torch.bfloat16
    torch.manual_seed(0)
    mods = []
    for i in range(len(sizes) - 1):
        mods.append(nn.Linear(sizes[i], sizes[i + 1]))
        if i < len(sizes) - 2:
            mods.append(nn.ReLU())
    model = nn.Sequential(*mods).to(dev)
    params = list(model.parameters())
    opt = torch.optim.SGD(params, lr=1e-4, momentum=0.9)
    x = torch.randn(rows, sizes[0], device=dev).to(cdt)
    y = torch.randn(rows, sizes[-1], device=dev)
    loss_fn = nn.MSELoss()

    def step():
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=(cdt == torch.bfloat16)):
            out = model(x)
        loss = loss_fn(out.float(), y)
        loss.backward()
        if world > 1:
            flat = torch._utils._flatten_dense_tensors([p.grad for p in params])
            dist.all_reduce(flat, op=dist.ReduceOp.AVG)
            for p, g in zip(params, torch._utils._unflatten_dense_tensors(flat, [p.grad for p in params])):
                p.grad.copy_(g)
        opt.step()
        return loss

    K, W = args.steps, max(args.warmup, 3)
    for _ in range(W):
        step()
    torch.cuda.synchronize(); barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        loss = step()
    e1.record()
    torch.cuda.synchronize(); barrier(world)
    ms = max_over_ranks(e0.elapsed_time(e1), world) / K
    emit(rank, {"metric": "MLP train samples/sec (device-timed, max over ranks)",
                "value": rows_total / (ms * 1e-3), "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms, "higher_is_better": True, "scaling": cfgd["scaling"], "vs_baseline": None,
                "dtype": cfgd["dtype"], "data": "synthetic", "impl": "torch (NCCL all-reduce + cuBLAS + torch SGD)",
                "config": {"model": cfgd["name"], "global_batch": rows_total, "parallelism": f"dp{world}"},
                "final_loss": float(loss.item())})


# -------------------------------------------------------------------------------------- reference
def run_reference(args, rank, world):
    cfgd = CONFIGS[args.config]
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    script = os.path.join(ref_dir, "dataParallelTraining_NN_MPI.py")
    if args.config != "ref":
        emit(rank, {"impl": "reference", "unavailable":
                    "the reference hard-codes a 2-3-1 MLP on 16 make_regression rows (ref.py:41-45,72) and "
                    f"cannot run config '{args.config}'; compare with: bench.py --config ref [--impl reference]"})
        return
    if not os.path.exists(script):
        emit(rank, {"impl": "reference", "unavailable":
                    "baseline/_ref is empty: run baseline/install_reference.sh (the reference has no "
                    "setup.py/pyproject.toml, so the plain pip command cannot install it)"})
        return
    if world > 1 and 16 % world != 0:
        emit(rank, {"impl": "reference", "unavailable":
                    f"the reference's uneven-shard path fails for {world} ranks: its int8 count buffer is "
                    "broadcast as MPI.INT (ref.py:118,128,133)"})
        return
    sys.path.insert(0, os.path.join(ROOT, "baseline", "mpi4py_shim"))
    sys.path.insert(0, ref_dir)
    import contextlib
    import io
    from mpi4py import MPI                     # the gloo shim (real mpi4py is not installed on this image)
    import dataParallelTraining_NN_MPI as ref  # the unmodified reference module

    K, W = args.steps, max(args.warmup, 3)
    ns = argparse.Namespace(lr=0.001, momentum=0.9, batch_size=4, nepochs=W + K)   # its own defaults
    with contextlib.redirect_stdout(io.StringIO()):
        ref.dist_train(ns)                     # the reference's public entry point (ref.py:56)
    stamps = MPI.COMM_WORLD.gather_times       # one comm.gather per training step (ref.py:185)
    assert len(stamps) == W + K, (len(stamps), W + K)
    ms = (stamps[-1] - stamps[W - 1]) * 1e3 / K if W >= 1 else (stamps[-1] - stamps[0]) * 1e3 / (K - 1)
    ms = max_over_ranks(ms, world)
    emit(rank, {"metric": "MLP train samples/sec (host-timed: the reference is CPU-only)",
                "value": 16 / (ms * 1e-3), "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "fp32", "data": "synthetic", "impl": "reference",
                "transport": "gloo shim of the mpi4py API (no MPI runtime on this image)",
                "config": {"model": cfgd["name"], "global_batch": 16, "parallelism": f"dp{world}"}})


def main():
    args = parse()
    if args.impl == "reference" and args.config != "ref":
        run_reference(args, int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))
        return
    rank, world, _ = dist_setup(args)
    try:
        if args.impl == "ours":
            run_ours(args, rank, world)
        elif args.impl == "torch":
            run_torch(args, rank, world)
        else:
            run_reference(args, rank, world)
    finally:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            try:
                dist.barrier()
                dist.destroy_process_group()
            except Exception:
                pass


if __name__ == "__main__":
    main()
