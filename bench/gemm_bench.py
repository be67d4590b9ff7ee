"""Times the three Linear GEMMs (CUDA events, L2 flushed between iterations) against torch.matmul
(cuBLAS) and the measured peak in MEASURED_PEAKS.json.  python bench/gemm_bench.py [M N K]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200dp  # noqa: E402,F401
from b200dp.ops import gemm  # noqa: E402


def timeit(fn, iters=10, warm=3):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    M, N, K = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (8192, 4096, 4096)
    peaks = {}
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peaks = json.load(open(p))
    peak = peaks.get("bf16_tflops", 1590.0)
    x = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    dy = torch.randn(M, N, device="cuda").bfloat16()
    b = torch.randn(N, device="cuda")
    y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    dx = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
    dw = torch.empty(N, K, device="cuda", dtype=torch.float32)
    fl = 2.0 * M * N * K
    rows = []
    variants = [("1cta_n256", dict(cta_group=1, tile_n=256)), ("1cta_n128", dict(cta_group=1, tile_n=128)),
                ("2cta", dict(cta_group=2))]
    dwb = torch.empty(N, K, device="cuda", dtype=torch.bfloat16)
    for name, ours, lib in [
        ("fwd  ", lambda **kw: gemm.linear_fwd(x, w, b, True, out=y, **kw), lambda: torch.relu(torch.addmm(b.bfloat16(), x, w.t()))),
        ("dgrad", lambda **kw: gemm.linear_dgrad(dy, w, x, out=dx, **kw), lambda: (dy @ w) * (x > 0)),
        ("wgrad", lambda **kw: gemm.linear_wgrad(dy, x, out=dwb, **kw), lambda: dy.t() @ x),
    ]:
        lmed, lbest = timeit(lib)
        rec = {"op": name.strip(), "M": M, "N": N, "K": K, "torch_ms": lmed, "torch_tflops": fl / lmed / 1e9}
        for vname, kw in variants:
            try:
                med, best = timeit(lambda: ours(**kw))
                rec[vname + "_tflops"] = fl / med / 1e9
                rec[vname + "_frac_of_measured_peak"] = fl / med / 1e9 / peak
            except Exception as e:  # noqa: BLE001
                rec[vname + "_error"] = str(e)[:200]
        rows.append(rec)
        print(json.dumps(rec))
    return rows


if __name__ == "__main__":
    main()
