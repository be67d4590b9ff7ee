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

--impl ours (default)    this framework.  `value` = kernel-only (inputs resident on the device); `e2e` =
                         the same metric through Trainer.fit_host_batches (pinned host batches, H2D every
                         step, loss read back every step).  Unless switched off the same process also
                         measures, on the same GPUs right after its own arm:
                           torch_baseline    the library bar BASELINE.json names (see --impl torch), with
                                             vs_torch / e2e_vs_torch;
                           reference_config  the UNMODIFIED reference from baseline/_ref at the only problem
                                             it can run (2-3-1 MLP, 16 rows) against this framework on that
                                             same problem, same N;
                           sustained         the step time over a >= 3 s loop under the power cap (the
                                             K-step timed region of a 3 ms step is a burst measurement);
                           grad_sync_us      busy and exposed (not hidden by backward) time of the fused
                                             gradient sync + update kernels per step.
--impl torch             DDP-grade library build of the same MLP: cuBLAS GEMMs, bf16 weights with fp32
                         master copy (no per-step autocast), NCCL all-reduce bucketed and overlapped with
                         backward (DistributedDataParallel), fused SGD (torch.optim.SGD(fused=True)), with
                         the same clock record and the same end-to-end (host-fed) measurement.
--impl reference         the UNMODIFIED reference script from baseline/_ref, run through its own
                         dist_train(args) over the gloo mpi4py shim (real MPI is not on this image).
                         It hard-codes its model and data (ref.py:41-45,72), so it can only run
                         --config ref; for the other configs it reports `unavailable`.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
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
METRIC = "MLP train samples/sec (device-timed, max over ranks)"


def parse(argv=None):
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
    p.add_argument("--sustained-s", type=float, default=3.0,
                   help="also time a loop of at least this many seconds (0 = skip)")
    p.add_argument("--no-torch-arm", action="store_true", help="skip the in-process library baseline")
    p.add_argument("--no-ref-config", action="store_true", help="skip the in-process reference comparison")
    return p.parse_args(argv)


def dist_setup(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if torch.cuda.is_available() and args.impl != "reference":      # the reference arm is CPU-only: no GPU, none of our code
        import b200dp  # noqa: F401
        from b200dp.launch import bind_cpu_near_gpu, device_index_for
        dev = device_index_for(local, int(os.environ.get("LOCAL_WORLD_SIZE", str(world))))
        torch.cuda.set_device(dev)
        bind_cpu_near_gpu(dev)          # before any pinned allocation: host batches land on the GPU's NUMA node
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = "cpu:gloo,cuda:nccl" if (torch.cuda.is_available() and args.impl != "reference") else "gloo"
        kw = {}
        if "nccl" in backend:
            kw["device_id"] = torch.device("cuda", torch.cuda.current_device())
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, world, local


def gather_ranks(x, world):
    if world == 1:
        return [x]
    import torch.distributed as dist
    out = [None] * world
    dist.all_gather_object(out, x)
    return out


def max_over_ranks(x: float, world: int) -> float:
    return max(gather_ranks(float(x), world))


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def emit(rank, rec):
    if rank == 0:
        print(json.dumps(rec))
        sys.stdout.flush()


def executed_flops(sizes, rows):
    """FLOPs one step really executes: forward and wgrad for every layer, dgrad for every layer but
    the first (the input needs no gradient) - not the 6*N*D convention, which counts that dgrad too."""
    full = sum(sizes[i] * sizes[i + 1] for i in range(len(sizes) - 1))
    return 2.0 * rows * (3 * full - sizes[0] * sizes[1])


class Timed:
    """K steps bracketed by barrier + synchronize, CUDA events on the launching stream, max over ranks,
    NVML clock samples taken inside the bracket."""

    def __init__(self, world, sampler):
        self.world, self.sampler = world, sampler

    def run(self, step, K):
        import torch
        torch.cuda.synchronize()
        barrier(self.world)
        lo = self.sampler.position() if self.sampler else 0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = None
        for _ in range(K):
            out = step()
        e1.record()
        torch.cuda.synchronize()
        hi = self.sampler.position() if self.sampler else 0
        barrier(self.world)
        ms = max_over_ranks(e0.elapsed_time(e1), self.world)
        clocks = self.sampler.summary(lo, max(hi, lo + 1)) if self.sampler else {}
        return ms, out, clocks


def clocks_field(c):
    return {"sm_mhz": c.get("sm_mhz"), "sm_max_mhz": c.get("sm_max_mhz"), "reasons": c.get("reasons", []),
            "samples": c.get("samples", 0), "period_ms": c.get("period_ms"), "power_w_max": c.get("power_w_max")}


def bare_h2d_ms(nbytes_x, nbytes_y, world, iters=6):
    """All ranks copy one step's input bytes from pinned host memory at the same time, nothing else
    running: the PCIe floor of the end-to-end step on this box at this N (max over ranks)."""
    import torch
    from b200dp.data import PinnedBuffer
    bx, by = PinnedBuffer((nbytes_x,), torch.uint8), PinnedBuffer((max(nbytes_y, 1),), torch.uint8)
    hx, hy = bx.tensor, by.tensor
    dx = torch.empty(nbytes_x, dtype=torch.uint8, device="cuda")
    dy = torch.empty(max(nbytes_y, 1), dtype=torch.uint8, device="cuda")
    for _ in range(2):
        dx.copy_(hx, non_blocking=True); dy.copy_(hy, non_blocking=True)
    torch.cuda.synchronize()
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        dx.copy_(hx, non_blocking=True); dy.copy_(hy, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    barrier(world)
    ms = e0.elapsed_time(e1) / iters
    del hx, hy
    bx.close(); by.close()
    return max_over_ranks(ms, world)


# ------------------------------------------------------------------------------------------- ours
def make_trainer(args, cfgd, config_name, world):
    import torch
    import b200dp  # noqa: F401
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    from b200dp.parallel.comm import world_comm
    if config_name in ("ref", "ref_uneven"):
        cfg = TrainConfig(quiet=True, nepochs=0, sync=args.sync, data_dist="local",
                          n_samples=cfgd["rows_total"],
                          grad_weighting="samples" if config_name == "ref_uneven" else "uniform")
    else:
        rows = args.rows_per_gpu or cfgd["rows_per_gpu"]
        cfg = TrainConfig(quiet=True, nepochs=0, n_samples=rows * world, n_features=cfgd["n_features"],
                          hidden=cfgd["hidden"], n_outputs=cfgd["n_outputs"], dtype="bf16", sync=args.sync,
                          data_source="randn", data_dist="local", scale="none", init="device", lr=1e-4,
                          noise=0.1, overlap=not args.no_overlap, bucket_mb=args.bucket_mb)
    return Trainer(cfg, world_comm(), torch.device("cuda", torch.cuda.current_device()))


def measure_ours(args, config_name, rank, world, sampler, extras=True):
    """Kernel-only and end-to-end numbers of this framework for one config.  Returns the record."""
    import torch
    from b200dp.ops import _lib
    cfgd = CONFIGS[config_name]
    t = make_trainer(args, cfgd, config_name, world)
    rows_local = t.shard.rows
    global_batch = sum(t.counts)
    K, W = args.steps, max(args.warmup, 3)
    timed = Timed(world, sampler)

    # ---- kernel-only: inputs resident on the device -------------------------------------------
    graphed = (args.cuda_graph == "on" or (args.cuda_graph == "auto" and config_name.startswith("ref"))) \
        and t.capture_full_shard_step()
    step = t.graph_step if graphed else (lambda: t.step(t.x, t.y))
    torch.cuda.synchronize()
    barrier(world)                       # ranks enter the first cross-GPU kernel together
    for _ in range(W):
        step()
    l0 = _lib.launches
    ms, loss, clocks = timed.run(step, K)
    launches = (_lib.launches - l0) if not graphed else K * t._graph_launches
    chk = getattr(t.sync, "check_errors", None)
    if chk:
        chk()
    ms_per_step = ms / K
    value = global_batch / (ms_per_step * 1e-3)
    final_loss = float(loss.item())

    # ---- end to end: pinned host batches, H2D every step, loss read back every step -----------
    e2e = None
    if not args.no_e2e:
        # host batches in the framework's pinned staging buffers (data.PinnedBuffer: huge-page advised,
        # cudaHostRegister'ed) - what a user of the host-fed API is told to allocate
        from b200dp.data import pinned_like
        hb = [pinned_like(t.shard.x.to(t.x.dtype)), pinned_like(t.shard.x.to(t.x.dtype)),
              pinned_like(t.shard.y), pinned_like(t.shard.y)]
        xh, yh = [hb[0].tensor, hb[1].tensor], [hb[2].tensor, hb[3].tensor]

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
        h2d = int(t.h2d_bytes_per_step)
        e2e = {"value": global_batch / (ems * 1e-3), "unit": "samples/s", "ms_per_step": ems,
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(t.d2h_bytes_per_step),
               "h2d_gbs_per_rank": h2d / (ems * 1e-3) / 1e9}
        if extras and h2d > (1 << 20):
            floor = bare_h2d_ms(xh[0].numel() * xh[0].element_size(), yh[0].numel() * yh[0].element_size(), world)
            e2e["bare_h2d_ms"] = floor                       # all ranks copying at once, nothing else running
            e2e["bound"] = "pcie" if floor > 0.9 * ms_per_step else "compute"
            e2e["fraction_of_bound"] = max(floor, ms_per_step) / ems
        del xh, yh
        for b in hb:
            b.close()

    # ---- grad-sync us/step: busy and exposed time of the fused sync + update kernels ------------
    grad_sync = None
    if extras and hasattr(t.sync, "start_profile") and not graphed:
        t.sync.start_profile()
        for _ in range(10):
            step()
        prof = t.sync.stop_profile()
        if prof:
            mine = {"busy_us": statistics.median(p["busy_us"] for p in prof),
                    "exposed_us": statistics.median(p["exposed_us"] for p in prof),
                    "kernels_per_step": prof[0]["kernels"]}
            allr = gather_ranks(mine, world)
            grad_sync = {"busy_us": max(a["busy_us"] for a in allr), "exposed_us": max(a["exposed_us"] for a in allr),
                         "kernels_per_step": mine["kernels_per_step"],
                         "how": "CUDA events around every sync(+update) kernel on the side stream (busy) and "
                                "around the main stream's final join (exposed); median of 10 steps, max over ranks"}

    # ---- sustained: the same step for >= sustained_s seconds (power-capped steady state); last, so
    # the end-to-end measurement above still sees the clocks the kernel-only one saw --------
    sustained = None
    if extras and args.sustained_s > 0 and not config_name.startswith("ref"):
        n = max(K, int(args.sustained_s * 1e3 / ms_per_step) + 1)
        sms, _, sclk = timed.run(step, n)
        sustained = {"steps": n, "seconds": sms / 1e3, "ms_per_step": sms / n,
                     "value": global_batch / (sms / n * 1e-3), "clocks": clocks_field(sclk)}

    spec = t.spec
    rec = {
        "metric": METRIC, "value": value, "unit": "samples/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": cfgd["scaling"], "vs_baseline": None, "dtype": cfgd["dtype"], "data": "synthetic",
        "impl": "ours",
        "config": {"model": cfgd["name"], "global_batch": global_batch, "seq_len": None,
                   "parallelism": f"dp{world}", "sync": getattr(t._runtime, "sync_name", t.sync.name),
                   "rows_per_gpu": rows_local, "params": spec.num_params, "cuda_graph": bool(graphed),
                   "device": torch.cuda.current_device(),
                   "l2": "inputs+weights larger than L2 (no flush needed)" if not config_name.startswith("ref")
                   else "working set is 13 parameters: cache-resident by nature"},
        "clocks": clocks_field(clocks),
        "e2e": e2e, "gpu_launches": launches,
        "tflops_per_gpu": executed_flops(spec.sizes, rows_local) / (ms_per_step * 1e-3) / 1e12,
        "tflops_note": "executed FLOPs (no dgrad for the first layer)", "final_loss": final_loss,
    }
    if sustained:
        sustained["tflops_per_gpu"] = executed_flops(spec.sizes, rows_local) / (sustained["ms_per_step"] * 1e-3) / 1e12
        rec["sustained"] = sustained
    if grad_sync:
        rec["grad_sync_us"] = grad_sync
    t.close()
    del t
    torch.cuda.empty_cache()
    return rec


def run_ours(args, rank, world):
    import torch
    from b200dp.utils.clocks import ClockSampler
    period = int(os.environ.get("B200DP_CLOCK_PERIOD_MS", "20"))
    sampler = ClockSampler(torch.cuda.current_device(), period_ms=period).start() if (rank == 0 and period > 0) else None
    if sampler is not None:
        sampler.wait_first_sample(2.0)
    rec = measure_ours(args, args.config, rank, world, sampler)
    wide = not args.config.startswith("ref")
    if wide and not args.no_torch_arm:
        try:
            tb = measure_torch(args, args.config, rank, world, sampler)
            rec["torch_baseline"] = {k: tb[k] for k in ("impl", "value", "ms_per_step", "e2e", "clocks", "final_loss")}
            rec["vs_torch"] = rec["value"] / tb["value"]
            if rec.get("e2e") and tb.get("e2e"):
                rec["e2e_vs_torch"] = rec["e2e"]["value"] / tb["e2e"]["value"]
        except Exception as ex:                                   # the product number must survive a baseline failure
            rec["torch_baseline"] = {"error": f"{type(ex).__name__}: {ex}"[:300]}
    if wide and not args.no_ref_config:
        try:
            rec["reference_config"] = reference_config_block(args, rank, world, sampler)
        except Exception as ex:
            rec["reference_config"] = {"error": f"{type(ex).__name__}: {ex}"[:300]}
    if sampler is not None:
        sampler.stop()
    emit(rank, rec)


def reference_config_block(args, rank, world, sampler):
    """Same box, same N: the unmodified reference at the only problem it can run (its hard-coded 2-3-1
    MLP on 16 rows, ref.py:41-45,72) against this framework on that same problem."""
    sub = argparse.Namespace(**vars(args))
    sub.sync, sub.no_e2e = "auto", False
    ours = measure_ours(sub, "ref", rank, world, sampler, extras=False)
    ref = measure_reference(args, rank, world)
    out = {"config": CONFIGS["ref"]["name"], "n_gpus": world,
           "ours": {"kernel_only_us_per_step": ours["ms_per_step"] * 1e3, "value": ours["value"],
                    "e2e_us_per_step": ours["e2e"]["ms_per_step"] * 1e3, "e2e_value": ours["e2e"]["value"],
                    "cuda_graph": ours["config"]["cuda_graph"], "sync": ours["config"]["sync"]},
           "reference": ref}
    if "ms_per_step" in ref:
        out["e2e_ratio"] = ours["e2e"]["value"] / ref["value"]
        out["kernel_only_ratio"] = ours["value"] / ref["value"]
    return out


# ------------------------------------------------------------------------------------------ torch
def measure_torch(args, config_name, rank, world, sampler):
    """The library bar: cuBLAS + NCCL (DDP, buckets overlapped with backward) + fused SGD, bf16
    weights with an fp32 master copy.  Same timing bracket, clock record and end-to-end path as ours."""
    import torch
    import torch.distributed as dist
    from torch import nn
    cfgd = CONFIGS[config_name]
    dev = torch.device("cuda", torch.cuda.current_device())
    sizes = [cfgd["n_features"], *cfgd["hidden"], cfgd["n_outputs"]]
    ref_like = config_name.startswith("ref")
    if ref_like:
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
    lr, mu = 1e-4, 0.9
    mixed = cdt == torch.bfloat16
    if mixed:
        master = [p.detach().clone().float() for p in model.parameters()]
        model = model.to(torch.bfloat16)
        master_grads = [torch.zeros_like(m) for m in master]
        for m, g in zip(master, master_grads):
            m.grad = g
        opt = torch.optim.SGD(master, lr=lr, momentum=mu, fused=True)
    else:
        opt = torch.optim.SGD(model.parameters(), lr=lr, momentum=mu, fused=True)
    net = model
    if world > 1:
        net = nn.parallel.DistributedDataParallel(model, device_ids=[dev.index], gradient_as_bucket_view=True,
                                                  bucket_cap_mb=32)
    params = list(model.parameters())
    loss_fn = nn.MSELoss()
    x = torch.randn(rows, sizes[0], device=dev).to(cdt)
    y = torch.randn(rows, sizes[-1], device=dev)

    def step_on(xb, yb):
        for p in params:
            p.grad = None
        loss = loss_fn(net(xb).float(), yb)
        loss.backward()                                        # DDP: bucketed all-reduce (mean) under backward
        if mixed:
            with torch.no_grad():
                torch._foreach_copy_(master_grads, [p.grad for p in params])
                opt.step()
                torch._foreach_copy_(params, master)
        else:
            opt.step()
        return loss

    K, W = args.steps, max(args.warmup, 3)
    timed = Timed(world, sampler)
    for _ in range(W):
        step_on(x, y)
    ms, loss, clocks = timed.run(lambda: step_on(x, y), K)
    ms_per_step = ms / K
    rec = {"metric": METRIC, "value": rows_total / (ms_per_step * 1e-3), "unit": "samples/s", "n_gpus": world,
           "steps": K, "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True,
           "scaling": cfgd["scaling"], "vs_baseline": None, "dtype": cfgd["dtype"], "data": "synthetic",
           "impl": "torch (cuBLAS + DDP/NCCL overlapped buckets + fused SGD, bf16 weights + fp32 master)",
           "config": {"model": cfgd["name"], "global_batch": rows_total, "parallelism": f"dp{world}"},
           "clocks": clocks_field(clocks), "final_loss": float(loss.item()), "e2e": None}

    if not args.no_e2e:
        # host-fed: pinned batches, H2D of step i+1 under step i on a copy stream, loss read back per step
        from b200dp.data import pinned_like          # the same staging allocator as the other arm
        hb = [pinned_like(x.cpu()), pinned_like(x.cpu()), pinned_like(y.cpu()), pinned_like(y.cpu())]
        xh, yh = [hb[0].tensor, hb[1].tensor], [hb[2].tensor, hb[3].tensor]
        slots = [(torch.empty_like(x), torch.empty_like(y)) for _ in range(2)]
        copy_stream = torch.cuda.Stream(device=dev)
        main = torch.cuda.current_stream()

        def fit(n):
            ready = [torch.cuda.Event(), torch.cuda.Event()]
            used = [torch.cuda.Event(), torch.cuda.Event()]
            lh = torch.zeros(n, dtype=torch.float32).pin_memory()
            levt = [torch.cuda.Event() for _ in range(n)]

            def stage(i):
                k = i & 1
                with torch.cuda.stream(copy_stream):
                    copy_stream.wait_event(used[k])
                    slots[k][0].copy_(xh[k], non_blocking=True)
                    slots[k][1].copy_(yh[k], non_blocking=True)
                    ready[k].record(copy_stream)
            stage(0)
            out = []
            for i in range(n):
                k = i & 1
                if i + 1 < n:
                    stage(i + 1)
                main.wait_event(ready[k])
                l = step_on(*slots[k])
                used[k].record(main)
                lh[i:i + 1].copy_(l.detach().reshape(1), non_blocking=True)
                levt[i].record(main)
                if i >= 1:
                    levt[i - 1].synchronize()
                    out.append(float(lh[i - 1]))
            levt[n - 1].synchronize()
            out.append(float(lh[n - 1]))
            return out

        torch.cuda.synchronize(); barrier(world)
        fit(W)
        torch.cuda.synchronize(); barrier(world)
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        a0.record()
        fit(K)
        a1.record()
        torch.cuda.synchronize()
        wall_ms = (time.perf_counter() - w0) * 1e3
        barrier(world)
        ems = max_over_ranks(max(a0.elapsed_time(a1), wall_ms), world) / K
        h2d = x.numel() * x.element_size() + y.numel() * y.element_size()
        rec["e2e"] = {"value": rows_total / (ems * 1e-3), "unit": "samples/s", "ms_per_step": ems,
                      "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4}
        del xh, yh
        for b in hb:
            b.close()
    del net, model, opt, params
    torch.cuda.empty_cache()
    return rec


def run_torch(args, rank, world):
    import torch
    from b200dp.utils.clocks import ClockSampler
    sampler = ClockSampler(torch.cuda.current_device(), period_ms=20).start() if rank == 0 else None
    if sampler is not None:
        sampler.wait_first_sample(2.0)
    rec = measure_torch(args, args.config, rank, world, sampler)
    if sampler is not None:
        sampler.stop()
    emit(rank, rec)


# -------------------------------------------------------------------------------------- reference
def reference_unavailable(args, world):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    script = os.path.join(ref_dir, "dataParallelTraining_NN_MPI.py")
    if not os.path.exists(script):
        return ("baseline/_ref is empty: run baseline/install_reference.sh (the reference has no "
                "setup.py/pyproject.toml, so the plain pip command cannot install it)")
    if world > 1 and 16 % world != 0:
        return (f"the reference's uneven-shard path fails for {world} ranks: its int8 count buffer is "
                "broadcast as MPI.INT (ref.py:118,128,133)")
    return None


def measure_reference(args, rank, world):
    """Runs the unmodified reference through its own dist_train(args) (ref.py:56) over the gloo shim;
    returns its record, or {"unavailable": why}."""
    why = reference_unavailable(args, world)
    if why:
        return {"impl": "reference", "unavailable": why}
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    for p in (os.path.join(ROOT, "baseline", "mpi4py_shim"), ref_dir):
        if p not in sys.path:
            sys.path.insert(0, p)
    import contextlib
    import io
    from mpi4py import MPI                     # the gloo shim (real mpi4py is not installed on this image)
    import dataParallelTraining_NN_MPI as ref  # the unmodified reference module

    K, W = args.steps, max(args.warmup, 3)
    ns = argparse.Namespace(lr=0.001, momentum=0.9, batch_size=4, nepochs=W + K)   # its own defaults
    n0 = len(MPI.COMM_WORLD.gather_times)
    with contextlib.redirect_stdout(io.StringIO()):
        ref.dist_train(ns)                     # the reference's public entry point (ref.py:56)
    stamps = MPI.COMM_WORLD.gather_times[n0:]  # one comm.gather per training step (ref.py:185)
    assert len(stamps) == W + K, (len(stamps), W + K)
    ms = (stamps[-1] - stamps[W - 1]) * 1e3 / K
    ms = max_over_ranks(ms, world)
    cfgd = CONFIGS["ref"]
    return {"metric": "MLP train samples/sec (host-timed: the reference is CPU-only)",
            "value": 16 / (ms * 1e-3), "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "fp32", "data": "synthetic", "impl": "reference",
            "transport": "gloo shim of the mpi4py API (no MPI runtime on this image)",
            "config": {"model": cfgd["name"], "global_batch": 16, "parallelism": f"dp{world}"}}


def run_reference(args, rank, world):
    if args.config != "ref":
        emit(rank, {"impl": "reference", "unavailable":
                    "the reference hard-codes a 2-3-1 MLP on 16 make_regression rows (ref.py:41-45,72) and "
                    f"cannot run config '{args.config}'; the same-box comparison at its own problem is the "
                    "`reference_config` block of `bench.py --impl ours` (or: bench.py --config ref --impl reference)"})
        return
    emit(rank, measure_reference(args, rank, world))


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
