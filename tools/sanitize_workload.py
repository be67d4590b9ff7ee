"""Small single-GPU workload that touches every kernel once; run under compute-sanitizer:
    compute-sanitizer --tool memcheck python tools/sanitize_workload.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200dp  # noqa: E402,F401
from b200dp.config import TrainConfig  # noqa: E402
from b200dp.engine import Trainer  # noqa: E402
from b200dp.ops import allreduce_sgd as ar, gemm  # noqa: E402

dev = torch.device("cuda")
# tiny path (SIMT whole-net kernel + local SGD), eager and multi-block
Trainer(TrainConfig(quiet=True, nepochs=2), device=dev).train()
Trainer(TrainConfig(quiet=True, nepochs=1, n_samples=4000, data_dist="local"), device=dev).train()
# tensor-core path: both GEMM kernels (CTA pair needs >= 55 tiles of 256x256), GEMV head, column sums, SGD
t = Trainer(TrainConfig(quiet=True, nepochs=2, n_samples=512, n_features=64, hidden=[256, 128], dtype="bf16",
                        data_source="randn", data_dist="local", scale="none"), device=dev)
t.train()
t8 = Trainer(TrainConfig(quiet=True, nepochs=1, n_samples=256, n_features=64, hidden=[128], n_outputs=8, dtype="bf16",
                         data_source="randn", data_dist="local", scale="none", use_batch_size=True, batch_size=64),
             device=dev)
t8.train()
x = torch.randn(2048, 256, device=dev).bfloat16()
w = torch.randn(2048, 256, device=dev).bfloat16()
for cg in (1, 2):
    y = gemm.linear_fwd(x, w, torch.zeros(2048, device=dev), True, cta_group=cg)
    gemm.linear_dgrad(y, w, x, cta_group=cg)
    gemm.linear_wgrad(y, x, cta_group=cg)
# fused sync, emulated world of 2, both modes
n = 4096
for mode in (0, 1):
    params = [torch.zeros(n, device=dev) for _ in range(2)]
    grads = [torch.randn(n, device=dev) for _ in range(2)]
    flags = [torch.zeros(ar.SYNC_MAX_BLOCKS * 8, device=dev, dtype=torch.int32) for _ in range(2)]
    mom = [torch.zeros(n, device=dev) for _ in range(2)]
    ep = [torch.zeros(ar.SYNC_MAX_BLOCKS, device=dev, dtype=torch.int32) for _ in range(2)]
    err = [torch.zeros(1, device=dev, dtype=torch.int32) for _ in range(2)]
    args = [ar.make_args(r, 2, [g.data_ptr() for g in grads], [p.data_ptr() for p in params], None,
                         [f.data_ptr() for f in flags], mom[r].data_ptr(), ep[r].data_ptr(), err[r].data_ptr(),
                         [1.0, 1.0], 2.0, False, mode, True, 0, 5000) for r in range(2)]
    dargs = torch.zeros(ar.device_args_bytes() * 2, device=dev, dtype=torch.uint8)
    ar.launch_emulated(args, 0, n, 0.1, 0.9, False, dargs)
torch.cuda.synchronize()
print("sanitize workload ok")
