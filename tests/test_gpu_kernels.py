"""SIMT kernels vs plain PyTorch fp32 references of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _flat(sizes, dtype="fp32", seed=0):
    from b200dp.models.mlp import FlatMLP, MLPSpec
    flat = FlatMLP(MLPSpec(sizes), "cuda", dtype)
    flat.init_like_reference(seed)
    return flat


@pytest.mark.parametrize("sizes,rows", [([2, 3, 1], 16), ([2, 3, 1], 5), ([5, 16, 8, 2], 300),
                                        ([7, 64, 64, 3], 5000), ([2, 3, 1], 100000)])
def test_tiny_mlp_matches_autograd(sizes, rows):
    from b200dp.ops import kernels as K
    torch.manual_seed(1)
    flat = _flat(sizes)
    x = torch.randn(rows, sizes[0], device="cuda")
    y = torch.randn(rows, sizes[-1], device="cuda")
    scratch = K.scratch(K.tiny_mlp_scratch_bytes(flat.spec, rows), "cuda")
    loss = torch.zeros((), device="cuda")
    K.tiny_mlp_fwd_bwd(K.mlp_desc(flat.spec), flat.params, x, y, flat.grads, loss, scratch)
    first = flat.grads.clone()
    K.tiny_mlp_fwd_bwd(K.mlp_desc(flat.spec), flat.params, x, y, flat.grads, loss, scratch)
    assert torch.equal(first, flat.grads)                     # deterministic, scratch reusable
    m = flat.as_module().cuda().double()
    ref_loss = torch.nn.functional.mse_loss(m(x.double()), y.double())
    ref_loss.backward()
    torch.testing.assert_close(loss.double(), ref_loss.detach(), rtol=1e-5, atol=1e-6)
    for seg, p in zip(flat.spec.segments, m.parameters()):
        got = flat.grads[seg.offset:seg.offset + seg.numel].view(seg.shape)
        torch.testing.assert_close(got.double(), p.grad, rtol=2e-4, atol=1e-6)


@pytest.mark.parametrize("n", [1, 7, 4096, 1 << 20])
@pytest.mark.parametrize("pred_dtype,dy_dtype", [(torch.float32, torch.float32), (torch.bfloat16, torch.bfloat16)])
def test_mse_loss_grad(n, pred_dtype, dy_dtype):
    from b200dp.ops import kernels as K
    torch.manual_seed(2)
    pred = torch.randn(n, device="cuda").to(pred_dtype)
    tgt = torch.randn(n, device="cuda")
    dy = torch.empty(n, device="cuda", dtype=dy_dtype)
    loss = torch.zeros((), device="cuda")
    scratch = K.scratch(K.mse_scratch_bytes(n), "cuda")
    for _ in range(2):                                         # second call reuses the scratch
        K.mse_loss_grad(pred, tgt, dy, loss, scratch)
    d = pred.float() - tgt
    torch.testing.assert_close(loss, (d * d).mean(), rtol=1e-5, atol=1e-7)
    tol = dict(rtol=1e-2, atol=1e-8) if dy_dtype == torch.bfloat16 else dict(rtol=1e-6, atol=1e-9)
    torch.testing.assert_close(dy.float(), 2 * d / n, **tol)


@pytest.mark.parametrize("rows,K_,nout", [(64, 256, 1), (1000, 4096, 1), (333, 512, 3), (8192, 4096, 1)])
@pytest.mark.parametrize("edt", [torch.bfloat16, torch.float32])
def test_head_fwd_bwd(rows, K_, nout, edt):
    """GEMV head + MSE + dY, and its backward with the fused column sums of dH (= the bias gradient
    of the layer below); bf16 elements (bf16 training) and fp32 elements (fp32 training)."""
    from b200dp.ops import kernels as K
    torch.manual_seed(3)
    H = torch.relu(torch.randn(rows, K_, device="cuda")).to(edt)
    W = (torch.randn(nout, K_, device="cuda") / K_ ** 0.5).to(edt)
    b = torch.randn(nout, device="cuda")
    y = torch.randn(rows, nout, device="cuda")
    yhat = torch.empty(rows, nout, device="cuda")
    dy = torch.empty(rows, nout, device="cuda")
    loss = torch.zeros((), device="cuda")
    scratch = K.scratch(K.head_scratch_bytes(rows, K_, nout), "cuda")
    K.head_fwd_mse(H, W, b, y, yhat, dy, loss, scratch)
    ref = H.float() @ W.float().t() + b
    torch.testing.assert_close(yhat, ref, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(loss, ((ref - y) ** 2).mean(), rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(dy, 2 * (ref - y) / (rows * nout), rtol=1e-3, atol=1e-7)
    # a rank's all-reduce weight folded into the loss gradient (the loss itself is not scaled)
    dy3, loss3 = torch.empty_like(dy), torch.zeros((), device="cuda")
    K.head_fwd_mse(H, W, b, y, yhat, dy3, loss3, scratch, grad_scale=3.0)
    torch.testing.assert_close(dy3, 3.0 * dy, rtol=1e-6, atol=0)
    assert torch.equal(loss3, loss)
    grad_dtypes = (torch.float32, torch.bfloat16) if edt == torch.bfloat16 else (torch.float32,)
    for gdt in grad_dtypes:
        dH = torch.empty(rows, K_, device="cuda", dtype=edt)
        dW = torch.empty(nout, K_, device="cuda", dtype=gdt)
        db = torch.empty(nout, device="cuda", dtype=gdt)
        dbp = torch.full((K_,), float("nan"), device="cuda", dtype=gdt)
        K.head_bwd(dy, W, H, dH, dW, db, scratch, db_prev=dbp)
        tol = dict(rtol=2e-2, atol=1e-5) if gdt == torch.bfloat16 else dict(rtol=1e-3, atol=1e-6)
        torch.testing.assert_close(dW.float(), dy.t() @ H.float(), **tol)
        torch.testing.assert_close(db.float(), dy.sum(0), **tol)
        want_dh = (dy @ W.float()) * (H.float() > 0)
        torch.testing.assert_close(dH.float(), want_dh, rtol=1e-2 if edt == torch.bfloat16 else 1e-4, atol=1e-7)
        torch.testing.assert_close(dbp.float(), want_dh.sum(0), **tol)
        # without db_prev nothing else changes
        dW2, db2 = torch.empty_like(dW), torch.empty_like(db)
        K.head_bwd(dy, W, H, dH, dW2, db2, scratch)
        assert torch.equal(dW2, dW) and torch.equal(db2, db)


@pytest.mark.parametrize("rows,n", [(64, 8), (1000, 4096), (8192, 1024)])
@pytest.mark.parametrize("edt", [torch.bfloat16, torch.float32])
def test_colsum(rows, n, edt):
    from b200dp.ops import kernels as K
    A = torch.randn(rows, n, device="cuda").to(edt)
    out = torch.empty(n, device="cuda")
    K.colsum(A, out, K.scratch(K.colsum_scratch_bytes(rows, n), "cuda"))
    torch.testing.assert_close(out, A.float().sum(0), rtol=1e-4, atol=1e-3)


def test_mse_grad_scale_and_denominator():
    from b200dp.ops import kernels as K
    torch.manual_seed(5)
    n = 4096
    pred, tgt = torch.randn(n, device="cuda"), torch.randn(n, device="cuda")
    dy, loss = torch.empty(n, device="cuda"), torch.zeros((), device="cuda")
    scratch = K.scratch(K.mse_scratch_bytes(n), "cuda")
    K.mse_loss_grad(pred, tgt, dy, loss, scratch, grad_scale=5.0, denom=n // 2)
    d = pred - tgt
    torch.testing.assert_close(loss, (d * d).sum() / (n // 2), rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(dy, 5.0 * 2 * d / (n // 2), rtol=1e-6, atol=1e-9)


@pytest.mark.parametrize("gdt", [torch.float32, torch.bfloat16])
def test_sgd_momentum_matches_torch_optim(gdt):
    from b200dp.ops import kernels as K
    torch.manual_seed(4)
    n = 4096 + 128
    p = torch.randn(n, device="cuda")
    ref_p = torch.nn.Parameter(p.clone())
    opt = torch.optim.SGD([ref_p], lr=0.01, momentum=0.9)
    m = torch.zeros(n, device="cuda")
    shadow = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    for step in range(4):
        g = torch.randn(n, device="cuda").to(gdt)
        ref_p.grad = g.float().clone()
        opt.step()
        K.sgd_momentum(p, g, m, shadow, 0.01, 0.9, first_step=(step == 0))
        torch.testing.assert_close(p, ref_p.detach(), rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(shadow.float(), p.bfloat16().float())


def test_cast():
    from b200dp.ops import kernels as K
    src = torch.randn(1000003, device="cuda")[:1000001]
    src = src[: src.numel() // 4 * 4 + 1].clone()
    dst = torch.empty(src.numel(), device="cuda", dtype=torch.bfloat16)
    K.cast_f32_bf16(src, dst)
    assert torch.equal(dst, src.bfloat16())


@pytest.mark.parametrize("dtype,cols", [(torch.bfloat16, 256), (torch.float32, 12), (torch.float32, 2)])
def test_gather_rows(dtype, cols):
    from b200dp.ops import kernels as K
    src = torch.randn(1000, cols, device="cuda").to(dtype)
    idx = torch.randperm(1000, device="cuda")[:333]
    assert torch.equal(K.gather_rows(src, idx), src[idx])


def test_minibatch_training_on_gpu_matches_cpu():
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    kw = dict(quiet=True, nepochs=3, batch_size=4, use_batch_size=True, hidden=[8], n_features=4, n_samples=32,
              lr=0.0003, scale="global", data_dist="local")
    g = Trainer(TrainConfig(**kw), device=torch.device("cuda"))
    c = Trainer(TrainConfig(**kw, backend="torch"), device=torch.device("cpu"))
    assert [h["loss"] for h in g.train()] == pytest.approx([h["loss"] for h in c.train()], rel=1e-4)
