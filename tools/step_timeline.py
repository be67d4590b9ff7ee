"""Kernel timeline of one training step (torch.profiler / CUPTI): per-kernel start, duration, stream,
and the idle gaps on the main stream.  python tools/step_timeline.py [--config wide] > profiles/timeline.md"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="wide")
    ap.add_argument("--gpus", type=int, default=1, help="under torchrun: world size (rank 0 prints its own timeline)")
    ap.add_argument("--sync", default="auto")
    ap.add_argument("--no-overlap", action="store_true")
    a = ap.parse_args()
    args = argparse.Namespace(config=a.config, sync=a.sync, rows_per_gpu=None, no_overlap=a.no_overlap, bucket_mb=0.0,
                              gpus=a.gpus, impl="ours")
    rank, world, _ = bench.dist_setup(args)
    if rank != 0:
        sys.stdout = open(os.devnull, "w")
    t = bench.make_trainer(args, bench.CONFIGS[a.config], a.config, world)
    for _ in range(5):
        t.step(t.x, t.y)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            t.step(t.x, t.y)
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    # keep the middle step: kernels between the 2nd and 3rd head_fwd
    heads = [i for i, e in enumerate(evs) if "head_fwd" in e.name]
    fwd0 = [i for i, e in enumerate(evs) if "gemm" in e.name]
    t0 = evs[0].time_range.start
    print("| start us | dur us | stream | kernel |\n|---|---|---|---|")
    main_stream = None
    rows = []
    for e in evs:
        import re
        name = re.sub(r"^(void )?b200dp::\(anonymous namespace\)::", "", e.name)
        name = re.sub(r"\(.*$", "", name)[:60]
        rows.append((e.time_range.start - t0, e.time_range.end - e.time_range.start, getattr(e, "stream", -1), name))
    for r in rows:
        print(f"| {r[0]:.1f} | {r[1]:.1f} | {r[2]} | {r[3]} |")
    total = rows[-1][0] + rows[-1][1] - rows[0][0]
    busy = {}
    for r in rows:
        busy[r[3]] = busy.get(r[3], 0.0) + r[1]
    print(f"\nspan of 3 steps: {total:.1f} us ({total/3:.1f} us/step)\n")
    print("| kernel | total us over 3 steps |\n|---|---|")
    for k, v in sorted(busy.items(), key=lambda kv: -kv[1]):
        print(f"| {k} | {v:.1f} |")
    sys.stdout.flush()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        t.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
