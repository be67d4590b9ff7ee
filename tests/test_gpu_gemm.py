"""tcgen05 GEMM numerics vs a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rand(shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).to(torch.bfloat16)


def _close(got, want, K):
    # bf16 inputs, fp32 accumulate: error ~ sqrt(K) * eps_fp32 * |a||b|, plus bf16 output rounding
    torch.testing.assert_close(got.float(), want, rtol=1.6e-2, atol=2e-2 * (K ** 0.5) * 0.05 + 1e-3)


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
    _close(got, want, K)
    got2 = gemm.linear_fwd(x, w, None, relu=False, tile_n=tile_n)
    _close(got2, x.float() @ w.float().t(), K)


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (512, 256, 384), (200, 136, 72), (4096, 4096, 4096)])
def test_dgrad_mask(M, N, K):
    """dX[M,K] = (dY[M,N] @ W[N,K]) * [X > 0]"""
    from b200dp.ops import gemm
    torch.manual_seed(1)
    dy, w = _rand((M, N)), _rand((N, K), N ** -0.5)
    x = torch.relu(_rand((M, K)))
    got = gemm.linear_dgrad(dy, w, x)
    want = (dy.float() @ w.float()) * (x.float() > 0)
    _close(got, want, N)
    got2 = gemm.linear_dgrad(dy, w, None)
    _close(got2, dy.float() @ w.float(), N)


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (512, 256, 384), (328, 136, 72), (8192, 4096, 4096)])
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_wgrad(M, N, K, out_dtype):
    """dW[N,K] = dY[M,N]^T @ X[M,K]  (reduction over the batch rows M)"""
    from b200dp.ops import gemm
    torch.manual_seed(2)
    dy, x = _rand((M, N), M ** -0.5), _rand((M, K))
    out = torch.empty(N, K, dtype=out_dtype, device="cuda")
    got = gemm.linear_wgrad(dy, x, out=out)
    want = dy.float().t() @ x.float()
    _close(got, want, M)


def test_strided_views_and_errors():
    from b200dp.ops import gemm
    from b200dp.ops._lib import B200dpError
    big = _rand((256, 512))
    x = big[:, :256]                       # lda = 512
    w = _rand((128, 256), 1 / 16)
    got = gemm.linear_fwd(x, w, None, relu=False)
    _close(got, x.float() @ w.float().t(), 256)
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
    _close(got, torch.relu(x.float() @ w.float().t() + b), K)


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (512, 256, 384), (200, 136, 72), (4096, 4096, 4096)])
def test_pair_dgrad(M, N, K):
    from b200dp.ops import gemm
    torch.manual_seed(1)
    dy, w = _rand((M, N)), _rand((N, K), N ** -0.5)
    x = torch.relu(_rand((M, K)))
    got = gemm.linear_dgrad(dy, w, x, cta_group=2)
    _close(got, (dy.float() @ w.float()) * (x.float() > 0), N)


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (512, 256, 384), (328, 136, 72), (8192, 4096, 4096)])
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_pair_wgrad(M, N, K, out_dtype):
    from b200dp.ops import gemm
    torch.manual_seed(2)
    dy, x = _rand((M, N), M ** -0.5), _rand((M, K))
    out = torch.empty(N, K, dtype=out_dtype, device="cuda")
    got = gemm.linear_wgrad(dy, x, out=out, cta_group=2)
    _close(got, dy.float().t() @ x.float(), M)
