"""One of every kernel a training step launches, at the benchmark's own shapes, between
cudaProfilerStart/Stop - the workload behind profiles/ncu_*.md:

    python tools/ncu_workload.py && ncu --set full --clock-control none --import-source on \
        --profile-from-start off -o gpurun_out/prof_step python tools/ncu_workload.py

Profiled region: one step of the wide MLP (12288 -> 4 x 4096 -> 1, 8192 rows: pair GEMMs fwd / dgrad with
fused column sums / wgrad with half-tile tail, 1-CTA wgrad chunk, GEMV head forward+MSE, head backward,
head reduce, SGD-momentum updates), one fused all-reduce+SGD of a 64 MB bf16 bucket for a world of 4
emulated on this GPU (two-shot), the reference's own 2-3-1 step (tiny_mlp kernel + update) and one
TF32 GEMM with its transpose kernel."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import b200dp  # noqa: E402,F401
from b200dp.config import TrainConfig  # noqa: E402
from b200dp.engine import Trainer  # noqa: E402
from b200dp.ops import allreduce_sgd as ar, gemm, kernels as K  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="all", help="all | step | sync | small")
    a = ap.parse_args()
    dev = torch.device("cuda")
    args = argparse.Namespace(config="wide", sync="auto", rows_per_gpu=None, no_overlap=True, bucket_mb=0.0)
    t = bench.make_trainer(args, bench.CONFIGS["wide"], "wide", 1)      # serial schedule: one kernel at a time
    for _ in range(3):
        t.step(t.x, t.y)
    # fused sync: world of 4 on one GPU, 64 MB of bf16 gradients
    world, n = 4, 32 * 1024 * 1024
    p0 = torch.randn(n, device=dev)
    params = [p0.clone() for _ in range(world)]
    shadow = [p0.bfloat16() for _ in range(world)]
    grads = [torch.randn(n, device=dev).bfloat16() for _ in range(world)]
    flags = [torch.zeros(ar.SYNC_MAX_BLOCKS * 8, device=dev, dtype=torch.int32) for _ in range(world)]
    mom = [torch.zeros(n, device=dev) for _ in range(world)]
    ep = [torch.zeros(ar.SYNC_MAX_BLOCKS, device=dev, dtype=torch.int32) for _ in range(world)]
    err = [torch.zeros(1, device=dev, dtype=torch.int32) for _ in range(world)]
    sargs = [ar.make_args(r, world, [g.data_ptr() for g in grads], [p.data_ptr() for p in params],
                          [s.data_ptr() for s in shadow], [f.data_ptr() for f in flags], mom[r].data_ptr(),
                          ep[r].data_ptr(), err[r].data_ptr(), [1.0] * world, float(world), True, 1, False, 37, 5000)
             for r in range(world)]
    dargs = torch.zeros(ar.device_args_bytes() * world, device=dev, dtype=torch.uint8)
    ar.launch_emulated(sargs, 0, n, 0.1, 0.9, False, dargs)
    tiny = Trainer(TrainConfig(quiet=True, nepochs=0), device=dev)
    tiny.step(tiny.x, tiny.y)
    xf = torch.randn(4096, 2048, device=dev)
    wf = torch.randn(2048, 2048, device=dev) / 45
    wT = torch.empty(2048, 2048, device=dev)
    of = gemm.linear_fwd(xf, wf, None, relu=True)
    torch.cuda.synchronize()

    torch.cuda.profiler.start()
    if a.only in ("all", "step"):
        t.step(t.x, t.y)
    if a.only in ("all", "sync"):
        ar.launch_emulated(sargs, 0, n, 0.1, 0.9, False, dargs)
    if a.only in ("all", "small"):
        tiny.step(tiny.x, tiny.y)
        K.transpose_f32(wf, wT)
        gemm.linear_fwd(xf, wf, None, relu=True, out=of)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    assert all(int(e.item()) == 0 for e in err)
    print("ncu workload ok")


if __name__ == "__main__":
    main()
