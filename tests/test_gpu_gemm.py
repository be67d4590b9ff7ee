"""tcgen05 GEMM numerics vs a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rand(shape, scale=1.0, dtype=torch.bfloat16):
    return (torch.randn(*shape, device="cuda") * scale).to(dtype)


def _close(got, want):
    """Outputs are O(1).  bf16 operands are exact in the fp32 reference, products are exact in fp32,
    so an fp32 result differs only by accumulation order (held to 1e-3); a bf16 result adds one
    rounding (half an ulp = 2^-9 relative)."""
    if got.dtype == torch.float32:
        torch.testing.assert_close(got, want, rtol=1e-3, atol=1e-3)
    else:
        torch.testing.assert_close(got.float(), want, rtol=8e-3, atol=1e-3)


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 256), (1024, 768, 1024), (200, 136, 72),
                                   (8192, 4096, 4096)])
@pytest.mark.parametrize("tile_n", [256, 128])
def test_forward_bias_relu(M, N, K, tile_n):
    from b200dp.ops import gemm
    torch.manual_seed(0)
    x, w = _rand((M, K)), _rand((N, K), K ** -0.5)
    b = torch.randn(N, device="cuda")
    got = gemm.linear_fwd(x, w, b, relu=True, tile_n=tile_n)
    want = torch.relu(x.float() @ w.float().t() + b)
    _close(got, want)
    got2 = gemm.linear_fwd(x, w, None, relu=False, tile_n=tile_n)
    _close(got2, x.float() @ w.float().t())


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (512, 256, 384), (200, 136, 72), (4096, 4096, 4096)])
def test_dgrad_mask(M, N, K):
    """dX[M,K] = (dY[M,N] @ W[N,K]) * [X > 0]"""
    from b200dp.ops import gemm
    torch.manual_seed(1)
    dy, w = _rand((M, N)), _rand((N, K), N ** -0.5)
    x = torch.relu(_rand((M, K)))
    got = gemm.linear_dgrad(dy, w, x, cta_group=1)
    want = (dy.float() @ w.float()) * (x.float() > 0)
    _close(got, want)
    got2 = gemm.linear_dgrad(dy, w, None, cta_group=1)
    _close(got2, dy.float() @ w.float())


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (512, 256, 384), (328, 136, 72), (8192, 4096, 4096)])
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_wgrad(M, N, K, out_dtype):
    """dW[N,K] = dY[M,N]^T @ X[M,K]  (reduction over the batch rows M)"""
    from b200dp.ops import gemm
    torch.manual_seed(2)
    dy, x = _rand((M, N), M ** -0.5), _rand((M, K))
    out = torch.empty(N, K, dtype=out_dtype, device="cuda")
    got = gemm.linear_wgrad(dy, x, out=out, cta_group=1)
    want = dy.float().t() @ x.float()
    _close(got, want)


def test_strided_views_and_errors():
    from b200dp.ops import gemm
    from b200dp.ops._lib import B200dpError
    big = _rand((256, 512))
    x = big[:, :256]                       # lda = 512
    w = _rand((128, 256), 1 / 16)
    got = gemm.linear_fwd(x, w, None, relu=False)
    _close(got, x.float() @ w.float().t())
    with pytest.raises(B200dpError):
        gemm.linear_fwd(_rand((64, 64)), _rand((12, 64)), None, relu=False)   # N % 8 != 0


# ---- CTA-pair (cta_group::2) variant: same contract, 256x256 tiles over two SMs ----
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (512, 512, 256), (200, 136, 72), (1000, 776, 1032),
                                   (8192, 4096, 4096)])
def test_pair_forward(M, N, K):
    from b200dp.ops import gemm
    torch.manual_seed(0)
    x, w = _rand((M, K)), _rand((N, K), K ** -0.5)
    b = torch.randn(N, device="cuda")
    got = gemm.linear_fwd(x, w, b, relu=True, cta_group=2)
    _close(got, torch.relu(x.float() @ w.float().t() + b))


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (512, 256, 384), (200, 136, 72), (4096, 4096, 4096)])
def test_pair_dgrad(M, N, K):
    from b200dp.ops import gemm
    torch.manual_seed(1)
    dy, w = _rand((M, N)), _rand((N, K), N ** -0.5)
    x = torch.relu(_rand((M, K)))
    got = gemm.linear_dgrad(dy, w, x, cta_group=2)
    _close(got, (dy.float() @ w.float()) * (x.float() > 0))


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (512, 256, 384), (328, 136, 72), (8192, 4096, 4096)])
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_pair_wgrad(M, N, K, out_dtype):
    from b200dp.ops import gemm
    torch.manual_seed(2)
    dy, x = _rand((M, N), M ** -0.5), _rand((M, K))
    out = torch.empty(N, K, dtype=out_dtype, device="cuda")
    got = gemm.linear_wgrad(dy, x, out=out, cta_group=2)
    _close(got, dy.float().t() @ x.float())


# ---- tail wave as half tiles: same bits, fewer idle SM pairs ----
@pytest.mark.parametrize("kind", ["fwd", "dgrad", "wgrad"])
def test_pair_split_tail_is_bit_identical(kind):
    """4096 x 4096 outputs = 256 tiles over 74 SM pairs: the last 34 run as 68 half tiles."""
    from b200dp.ops import gemm
    torch.manual_seed(5)
    if kind == "fwd":
        a, b = _rand((4096, 512)), _rand((4096, 512), 512 ** -0.5)
        run = lambda st: gemm.linear_fwd(a, b, None, relu=False, cta_group=2, split_tail=st)
        want = a.float() @ b.float().t()
    elif kind == "dgrad":
        a, b = _rand((4096, 512)), _rand((512, 4096), 512 ** -0.5)
        run = lambda st: gemm.linear_dgrad(a, b, None, cta_group=2, split_tail=st)
        want = a.float() @ b.float()
    else:
        a, b = _rand((1024, 4096), 1024 ** -0.5), _rand((1024, 4096))
        run = lambda st: gemm.linear_wgrad(a, b, out=torch.empty(4096, 4096, dtype=torch.bfloat16, device="cuda"),
                                           cta_group=2, split_tail=st)
        want = a.float().t() @ b.float()
    split, whole = run(0), run(-1)
    assert torch.equal(split, whole)
    _close(split, want)


# ---- fused bias gradient: column sums of the stored result, from the same epilogue ----
@pytest.mark.parametrize("M,N,K,cta_group", [(256, 256, 64, 1), (200, 136, 72, 1), (1000, 776, 264, 1),
                                             (1000, 776, 264, 2), (8192, 4096, 512, 2), (8192, 4096, 512, 1),
                                             (300, 264, 128, 2)])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
def test_dgrad_fused_colsum(M, N, K, cta_group, out_dtype):
    """dX = (dY W) * [X > 0] and, from the same launch, db_below = dX.sum(0) (the `sum` of
    AddmmBackward0, ref.py:176)."""
    from b200dp.ops import gemm
    torch.manual_seed(6)
    dy, w = _rand((M, K)), _rand((K, N), K ** -0.5)
    x = torch.relu(_rand((M, N)))
    scratch = torch.zeros(gemm.colsum_scratch_bytes(M, N), dtype=torch.uint8, device="cuda")
    cs = torch.full((N,), float("nan"), dtype=out_dtype, device="cuda")
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    gemm.linear_dgrad(dy, w, x, out=out, cta_group=cta_group, colsum_out=cs, colsum_scratch=scratch)
    want = (dy.float() @ w.float()) * (x.float() > 0)
    _close(out, want)
    want_cs = want.sum(0)
    tol = dict(rtol=8e-3, atol=2e-2) if out_dtype == torch.bfloat16 else dict(rtol=1e-3, atol=2e-3 * M ** 0.5 / 16)
    torch.testing.assert_close(cs.float(), want_cs, **tol)
    # the scratch is reusable as is (tickets are back at zero) and the result is bit-reproducible
    first = cs.clone()
    for _ in range(3):
        cs.fill_(float("nan"))
        gemm.linear_dgrad(dy, w, x, out=out, cta_group=cta_group, colsum_out=cs, colsum_scratch=scratch)
        assert torch.equal(cs, first)
    assert int(scratch[:4096].view(torch.int32).abs().sum()) == 0


# ---- fp32 operands on kind::tf32 (the reference's precision, ref.py:159) ----
def _close_tf32(got, want):
    # operands are read with 10 mantissa bits: ~1e-3 relative per product
    scale = float(want.abs().max())
    torch.testing.assert_close(got, want, rtol=5e-3, atol=4e-3 * scale)


@pytest.mark.parametrize("M,N,K", [(128, 256, 32), (256, 512, 256), (200, 136, 72), (1000, 776, 1032),
                                   (4096, 2048, 2048)])
@pytest.mark.parametrize("tile_n", [0, 128])
def test_tf32_forward(M, N, K, tile_n):
    from b200dp.ops import gemm
    torch.manual_seed(7)
    x, w = _rand((M, K), dtype=torch.float32), _rand((N, K), K ** -0.5, dtype=torch.float32)
    b = torch.randn(N, device="cuda")
    got = gemm.linear_fwd(x, w, b, relu=True, tile_n=tile_n)
    assert got.dtype == torch.float32
    _close_tf32(got, torch.relu(x @ w.t() + b))


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (512, 256, 384), (200, 136, 72), (2048, 2048, 1024)])
def test_tf32_dgrad_with_colsum(M, N, K):
    """kind::tf32 reads K-major operands only: dgrad runs on a transposed copy of W."""
    from b200dp.ops import gemm, kernels as Kk
    torch.manual_seed(8)
    dy, w = _rand((M, N), dtype=torch.float32), _rand((N, K), N ** -0.5, dtype=torch.float32)
    x = torch.relu(_rand((M, K), dtype=torch.float32))
    wT = torch.empty(K, N, device="cuda")
    Kk.transpose_f32(w, wT)
    assert torch.equal(wT, w.t())
    scratch = torch.zeros(gemm.colsum_scratch_bytes(M, K), dtype=torch.uint8, device="cuda")
    cs = torch.empty(K, device="cuda")
    got = torch.empty(M, K, device="cuda")
    gemm.gemm(dy, wT, got, mask_src=x, colsum_out=cs, colsum_scratch=scratch)
    want = (dy @ w) * (x > 0)
    _close_tf32(got, want)
    torch.testing.assert_close(cs, want.sum(0), rtol=5e-3, atol=4e-3 * float(want.sum(0).abs().max()))


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (512, 256, 384), (328, 136, 72), (4096, 1024, 2048)])
def test_tf32_wgrad(M, N, K):
    from b200dp.ops import gemm, kernels as Kk
    torch.manual_seed(9)
    dy, x = _rand((M, N), M ** -0.5, dtype=torch.float32), _rand((M, K), dtype=torch.float32)
    dyT, xT = torch.empty(N, M, device="cuda"), torch.empty(K, M, device="cuda")
    Kk.transpose_f32(dy, dyT)
    Kk.transpose_f32(x, xT)
    out = torch.empty(N, K, device="cuda")
    gemm.gemm(dyT, xT, out)
    _close_tf32(out, dy.t() @ x)


def test_tf32_refuses_transposed_operands():
    from b200dp.ops import gemm
    from b200dp.ops._lib import B200dpError
    a, b = _rand((64, 128), dtype=torch.float32), _rand((64, 256), dtype=torch.float32)
    with pytest.raises(B200dpError, match="K-major"):
        gemm.gemm(a, b, torch.empty(128, 256, device="cuda"), a_mn_major=True, b_mn_major=True)


@pytest.mark.parametrize("rows,cols", [(32, 32), (100, 36), (1000, 776), (4096, 2048)])
def test_transpose_f32(rows, cols):
    from b200dp.ops import kernels as Kk
    src = torch.randn(rows, cols + 4, device="cuda")[:, :cols]           # strided source
    dst = torch.empty(cols, rows + 4, device="cuda")[:, :rows]           # strided destination
    Kk.transpose_f32(src, dst)
    assert torch.equal(dst, src.t())
