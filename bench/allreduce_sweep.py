"""BASELINE.json configs[4]: gradient all-reduce + SGD sweep, 1 KB - 1 GB, ours (one-shot / two-shot /
NVLS) vs NCCL all-reduce + SGD kernel vs the reference's gather/root-average/send pattern.

    torchrun --nproc-per-node N --master-addr 127.0.0.1 bench/allreduce_sweep.py [--max-mb 1024] [--dtype bf16]

Device-timed (CUDA events), max over ranks.  Reports us/step, algorithmic bus bandwidth
(2(N-1)/N * S / t) and the fraction of the NVLink roofline (measured 770 GB/s per direction)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import b200dp  # noqa: E402,F401
from b200dp.launch import bind_device, discover, init_process_group  # noqa: E402
from b200dp.ops import allreduce_sgd as ar, kernels as K  # noqa: E402
from b200dp.ops.cuda_backend import SymmLayout  # noqa: E402
from b200dp.parallel.comm import world_comm  # noqa: E402

LINK_GBS = 770.0          # B200_PROFILING.md: measured peer copy per direction; re-measured here (peer_copy rows)


class _Raw:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def nvlink_bytes(dev_index):
    """(tx, rx) payload bytes this GPU has moved over all its NVLinks so far (NVML throughput counters,
    KiB granularity), or None when the counters are not exposed."""
    try:
        import pynvml
        pynvml.nvmlInit()
        uuid = torch.cuda.get_device_properties(dev_index).uuid
        h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode())
        tx_id = getattr(pynvml, "NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX", 138)
        rx_id = getattr(pynvml, "NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX", 139)
        try:
            vals = pynvml.nvmlDeviceGetFieldValues(h, [(tx_id, 0xFFFFFFFF), (rx_id, 0xFFFFFFFF)])   # all links
            out = [int(v.value.ullVal) for v in vals]
            if all(v.nvmlReturn == 0 for v in vals):
                return out[0] * 1024, out[1] * 1024
        except Exception:
            pass
        tx = rx = 0
        for link in range(18):
            vals = pynvml.nvmlDeviceGetFieldValues(h, [(tx_id, link), (rx_id, link)])
            if any(v.nvmlReturn != 0 for v in vals):
                continue
            tx += int(vals[0].value.ullVal); rx += int(vals[1].value.ullVal)
        return tx * 1024, rx * 1024
    except Exception:
        return None


def timed(fn, iters, world):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    out = [None] * world
    dist.all_gather_object(out, a.elapsed_time(b) / iters)
    return max(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=float, default=1024)
    ap.add_argument("--dtype", choices=("fp32", "bf16"), default="bf16")
    ap.add_argument("--blocks", type=int, default=0)
    ap.add_argument("--out", default=None)
    ap.add_argument("--with-gather-root", action="store_true")
    args = ap.parse_args()
    info = discover(); dev = bind_device(info); init_process_group(info)
    comm = world_comm(); rank, world = comm.rank, comm.size
    bf16 = args.dtype == "bf16"
    esz = 2 if bf16 else 4
    max_elems = int(args.max_mb * (1 << 20)) // esz
    max_elems = max_elems // 1024 * 1024
    symm = SymmLayout(comm, dev, max_elems, bf16, bf16, multicast=True)
    have_mc = symm.buf.mc_ptr is not None
    gdt = torch.bfloat16 if bf16 else torch.float32
    grads = symm.tensor("grads", gdt); params = symm.tensor("params", torch.float32)
    shadow = symm.tensor("shadow", torch.bfloat16) if bf16 else None
    mom = torch.zeros(max_elems, device=dev)
    epochs = torch.zeros(ar.SYNC_MAX_BLOCKS, dtype=torch.int32, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    grads.copy_(torch.randn(max_elems, device=dev).to(gdt))
    w = [1.0] * world
    rows = []
    sizes = []
    nbytes = 1024
    while nbytes <= args.max_mb * (1 << 20):
        sizes.append(nbytes); nbytes *= 4
    for nbytes in sizes:
        n = max(nbytes // esz // 8 * 8, 8)
        iters = 50 if nbytes <= (16 << 20) else 10
        rec = {"bytes": n * esz, "n_gpus": world, "dtype": args.dtype}
        for name, mode in (("oneshot", 0), ("twoshot", 1), ("nvls", 2)):
            if mode == 2 and not have_mc:
                continue
            if mode == 0 and nbytes > (64 << 20):
                continue
            a = ar.make_args(rank, world, symm.ptrs("grads"), symm.ptrs("params"),
                             symm.ptrs("shadow") if bf16 else None, symm.ptrs("flags"), mom.data_ptr(),
                             epochs.data_ptr(), err.data_ptr(), w, float(world), bf16, mode, not bf16,
                             args.blocks, 20000, symm.mc("grads"), symm.mc("shadow") if bf16 else None,
                             symm.mc("params"))
            c0 = nvlink_bytes(dev.index) if nbytes >= (16 << 20) else None
            ms = timed(lambda: ar.launch(a, 0, n, 1e-6, 0.9, False), iters, world)
            rec[name + "_us"] = ms * 1e3
            rec[name + "_busbw_gbs"] = 2 * (world - 1) / world * n * esz / (ms * 1e-3) / 1e9
            if c0 is not None:
                torch.cuda.synchronize()
                c1 = nvlink_bytes(dev.index)
                if c1 is not None:                      # timed() runs 3 warm-up + iters launches
                    calls = iters + 3
                    rec[name + "_nvlink_tx_bytes_per_call"] = (c1[0] - c0[0]) / calls
                    rec[name + "_nvlink_rx_bytes_per_call"] = (c1[1] - c0[1]) / calls
        # NCCL all-reduce + our local SGD kernel
        g = grads[:n]
        def nccl():
            dist.all_reduce(g, op=dist.ReduceOp.SUM)
            K.sgd_momentum(params[:n], g, mom[:n], shadow[:n] if bf16 else None, 1e-6, 0.9, False, 1.0 / world)
        ms = timed(nccl, iters, world)
        rec["nccl_sgd_us"] = ms * 1e3
        rec["nccl_busbw_gbs"] = 2 * (world - 1) / world * n * esz / (ms * 1e-3) / 1e9
        if args.with_gather_root and nbytes <= (64 << 20):
            def gather_root():
                lst = comm.gather(g.detach().to("cpu"), root=0)
                if rank == 0:
                    acc = torch.zeros(n)
                    for t in lst:
                        acc += t.float()
                    acc /= world
                    for k in range(1, world):
                        comm.send(acc, k)
                else:
                    acc = comm.recv(0)
                K.sgd_momentum(params[:n], acc.to(dev), mom[:n], shadow[:n] if bf16 else None, 1e-6, 0.9, False)
            ms = timed(gather_root, 3, world)
            rec["gather_root_us"] = ms * 1e3
        # plain peer copy over NVLink (every rank pulls from its right neighbour at the same time): the
        # link-rate denominator SURVEY 6.2 asks to be measured on this box
        if nbytes >= (1 << 20):
            peer = (rank + 1) % world
            src = torch.as_tensor(_Raw(symm.ptrs("grads")[peer], n * esz), device=dev)
            dst = torch.empty(n * esz, dtype=torch.uint8, device=dev)
            ms = timed(lambda: dst.copy_(src, non_blocking=True), iters, world)
            rec["peer_copy_us"] = ms * 1e3
            rec["peer_copy_gbs"] = n * esz / (ms * 1e-3) / 1e9
            del dst
        best = min(v for k, v in rec.items() if k.endswith("_us") and not k.startswith(("nccl", "gather", "peer")))
        link_bytes = 2 * (world - 1) / world * n * esz
        rec["algorithmic_link_bytes_twoshot"] = link_bytes
        rec["roofline_us"] = link_bytes / (LINK_GBS * 1e9) * 1e6
        rec["best_frac_of_link_roofline"] = rec["roofline_us"] / best
        rows.append(rec)
        if rank == 0:
            print(json.dumps(rec)); sys.stdout.flush()
    if int(err.item()):
        raise RuntimeError("sync watchdog fired")
    if rank == 0 and args.out:
        with open(args.out, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")
    torch.cuda.synchronize(); dist.barrier()
    symm.buf.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
