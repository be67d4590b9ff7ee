"""Short GEMM run for ncu: `python bench/profile_gemm.py [fwd|dgrad|wgrad] [M N K] [iters]`."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200dp  # noqa: E402,F401
from b200dp.ops import gemm  # noqa: E402

op = sys.argv[1] if len(sys.argv) > 1 else "fwd"
M, N, K = (int(v) for v in sys.argv[2:5]) if len(sys.argv) >= 5 else (8192, 4096, 4096)
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 4
cg = int(sys.argv[6]) if len(sys.argv) > 6 else 0
x = torch.randn(M, K, device="cuda").bfloat16()
w = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
dy = torch.randn(M, N, device="cuda").bfloat16()
b = torch.randn(N, device="cuda")
y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
dx = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
dw = torch.empty(N, K, device="cuda", dtype=torch.bfloat16)
for _ in range(iters):
    if op == "fwd":
        gemm.linear_fwd(x, w, b, True, out=y, cta_group=cg)
    elif op == "dgrad":
        gemm.linear_dgrad(dy, w, x, out=dx, cta_group=cg)
    else:
        gemm.linear_wgrad(dy, x, out=dw, cta_group=cg)
torch.cuda.synchronize()
print("ok", op, M, N, K)
