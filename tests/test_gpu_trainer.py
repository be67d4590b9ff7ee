"""End-to-end on one GPU: the CUDA backend reproduces the reference's loss trajectory (tiny path) and
tracks a PyTorch autograd model (tensor-core path)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_native_library_is_loaded():
    from b200dp.ops import _lib
    assert _lib.available(), "libb200dp.so missing: run python build.py"
    assert _lib.lib().b200dp_version() >= 100


def test_reference_config_golden_losses_on_gpu():
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    t = Trainer(TrainConfig(quiet=True), device=torch.device("cuda"))
    assert t.backend == "cuda" and t.compute.name == "cuda-tiny"
    losses = [h["loss"] for h in t.train()]
    assert losses == pytest.approx([2530.8320, 2526.6858, 2518.3696], rel=5e-6)
    t.close()


def test_gpu_matches_cpu_engine_over_many_epochs():
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    kw = dict(quiet=True, nepochs=50, hidden=[16, 8], n_features=5, n_samples=200, lr=0.0003,
              scale="global", data_dist="local")
    g = Trainer(TrainConfig(**kw), device=torch.device("cuda"))
    c = Trainer(TrainConfig(**kw, backend="torch"), device=torch.device("cpu"))
    lg = [h["loss"] for h in g.train()]
    lc = [h["loss"] for h in c.train()]
    assert lg == pytest.approx(lc, rel=2e-4)
    for k, v in c.flat.state_dict().items():
        torch.testing.assert_close(g.flat.state_dict()[k], v, rtol=2e-3, atol=2e-5)


@pytest.mark.parametrize("nout", [1, 8])
def test_tensor_core_path_tracks_autograd(nout):
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    cfg = TrainConfig(quiet=True, nepochs=1, hidden=[256, 128], n_features=64, n_samples=1024, n_outputs=nout,
                      dtype="bf16", data_source="randn", data_dist="local", scale="none", lr=0.01, noise=0.1)
    t = Trainer(cfg, device=torch.device("cuda"))
    assert t.compute.name == "cuda-tcgen05"
    # one step by hand so gradients can be inspected
    m = t.flat.as_module().cuda()
    for p in m.parameters():
        p.data = p.data.bfloat16().float()
    x32, y32 = t.shard.x.cuda(), t.shard.y.cuda()
    ref_loss = torch.nn.functional.mse_loss(m(x32.bfloat16().float()), y32)
    ref_loss.backward()
    loss = t.compute.loss_and_grads(t.x, t.y)
    torch.cuda.synchronize()
    torch.testing.assert_close(loss, ref_loss.detach(), rtol=2e-2, atol=1e-3)
    for seg, p in zip(t.spec.segments, m.parameters()):
        got = t.flat.grads[seg.offset:seg.offset + seg.numel].view(seg.shape).float()
        scale = p.grad.abs().max().item() + 1e-12
        assert (got - p.grad).abs().max().item() <= 5e-2 * scale, seg.key
    # and training reduces the loss
    cfg2 = TrainConfig(**{**cfg.asdict(), "nepochs": 30})
    t2 = Trainer(cfg2, device=torch.device("cuda"))
    hist = t2.train()
    assert hist[-1]["loss"] < 0.95 * hist[0]["loss"]


def test_checkpoint_from_gpu_loads_into_plain_module(tmp_path):
    from b200dp import MLP
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    path = str(tmp_path / "ck.pt")
    t = Trainer(TrainConfig(quiet=True, nepochs=2, checkpoint=path), device=torch.device("cuda"))
    t.train()
    m = MLP()
    m.load_state_dict(torch.load(path, weights_only=False)["model"])
    x = torch.randn(4, 2)
    torch.testing.assert_close(m(x), t.predict(x).cpu(), rtol=1e-5, atol=1e-6)


def test_cuda_graph_replay_matches_eager():
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    t = Trainer(TrainConfig(quiet=True, cuda_graph=True, nepochs=5), device=torch.device("cuda"))
    lg = [h["loss"] for h in t.train()]
    assert t._graph is not None and t._graph_launches == 2
    e = Trainer(TrainConfig(quiet=True, nepochs=5), device=torch.device("cuda"))
    le = [h["loss"] for h in e.train()]
    assert lg == le and lg[:3] == pytest.approx([2530.8320, 2526.6858, 2518.3696], rel=5e-6)
    assert t.flat.fingerprint() == e.flat.fingerprint()


def test_overlapped_and_serial_schedules_agree_bitwise():
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    kw = dict(quiet=True, nepochs=4, hidden=[256, 128, 64], n_features=64, n_samples=2048, dtype="bf16",
              data_source="randn", data_dist="local", scale="none", lr=0.01, noise=0.1)
    a = Trainer(TrainConfig(**kw, overlap=True), device=torch.device("cuda"))
    b = Trainer(TrainConfig(**kw, overlap=False), device=torch.device("cuda"))
    la = [h["loss"] for h in a.train()]
    lb = [h["loss"] for h in b.train()]
    assert la == lb
    assert a.flat.fingerprint() == b.flat.fingerprint()


def test_host_fed_training_equals_resident_training():
    """fit_host_batches (pinned host batches, H2D per step, lagged loss read-back) must produce exactly
    the losses and parameters of stepping on the resident shard."""
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    kw = dict(quiet=True, nepochs=0, hidden=[256, 128], n_features=64, n_samples=1024, dtype="bf16",
              data_source="randn", data_dist="local", scale="none", lr=0.01, noise=0.1)
    a = Trainer(TrainConfig(**kw), device=torch.device("cuda"))
    b = Trainer(TrainConfig(**kw), device=torch.device("cuda"))
    xh = a.shard.x.to(torch.bfloat16).pin_memory()
    yh = a.shard.y.pin_memory()
    la = a.fit_host_batches(iter([(xh, yh)] * 6), 6)
    lb = []
    for _ in range(6):
        lb.append(float(b.step(b.x, b.y).item()))
    assert la == lb and len(la) == 6
    assert a.flat.fingerprint() == b.flat.fingerprint()
    assert a.h2d_bytes_per_step == xh.numel() * 2 + yh.numel() * 4 and a.d2h_bytes_per_step == 4
    # fp32 host batches are accepted too (cast on the device)
    c = Trainer(TrainConfig(**kw), device=torch.device("cuda"))
    lc = c.fit_host_batches(iter([(a.shard.x.pin_memory(), yh)] * 2), 2)
    assert lc == la[:2]


def test_tensor_core_path_with_unaligned_widths():
    """hidden=100, 52 and 30 input features are padded to multiples of 8 inside the flat buffers."""
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    cfg = TrainConfig(quiet=True, nepochs=0, hidden=[100, 52], n_features=30, n_samples=512, dtype="bf16",
                      data_source="randn", data_dist="local", scale="none", lr=0.01, noise=0.1)
    t = Trainer(cfg, device=torch.device("cuda"))
    assert t.compute.name == "cuda-tcgen05" and t.spec.sizes == (32, 104, 56, 1)
    m = t.flat.as_module().cuda()
    for p in m.parameters():
        p.data = p.data.bfloat16().float()
    x32, y32 = t.shard.x.cuda(), t.shard.y.cuda()
    ref_loss = torch.nn.functional.mse_loss(m(x32.bfloat16().float()), y32)
    ref_loss.backward()
    loss = t.compute.loss_and_grads(t.x, t.y)
    torch.testing.assert_close(loss, ref_loss.detach(), rtol=2e-2, atol=1e-3)
    sd_keys = list(t.state_dict().items())
    assert [tuple(v.shape) for _, v in sd_keys] == [(100, 30), (100,), (52, 100), (52,), (1, 52), (1,)]
    for seg, p in zip(t.spec.segments, m.parameters()):
        got = t.flat._logical(t.flat.grads, seg).float()
        scale = p.grad.abs().max().item() + 1e-12
        assert (got - p.grad).abs().max().item() <= 5e-2 * scale, seg.key
        full = t.flat._view(t.flat.grads, seg).float()
        assert abs(float(full.abs().sum()) - float(got.abs().sum())) <= 1e-3 * float(got.abs().sum()) + 1e-6
    first = float(loss.item())            # the loss tensor is a persistent device scalar: read it now
    for _ in range(20):
        last = t.step(t.x, t.y)
    assert float(last.item()) < first
    pred = t.predict(t.shard.x[:8])
    assert pred.shape == (8, 1)


def test_cuda_graph_on_the_tensor_core_path_with_two_streams():
    """Graph capture of a step that forks to the side stream (bucketed update, column sums) and joins."""
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    kw = dict(quiet=True, nepochs=4, hidden=[256, 128, 64], n_features=64, n_samples=2048, dtype="bf16",
              data_source="randn", data_dist="local", scale="none", lr=0.01, noise=0.1)
    g = Trainer(TrainConfig(**kw, cuda_graph=True), device=torch.device("cuda"))
    e = Trainer(TrainConfig(**kw), device=torch.device("cuda"))
    lg = [h["loss"] for h in g.train()]
    le = [h["loss"] for h in e.train()]
    assert g._graph is not None and g._graph_launches > 5
    assert lg == le and g.flat.fingerprint() == e.flat.fingerprint()


def test_row_chunked_wgrad_and_buckets_are_bitwise_neutral():
    """A small bucket limit forces the first layer's weight into two row chunks (two wgrad GEMMs, two
    update launches); the result must be bit-identical to the unchunked schedule."""
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    kw = dict(quiet=True, nepochs=3, hidden=[2048, 64], n_features=4096, n_samples=1024, dtype="bf16",
              data_source="randn", data_dist="local", scale="none", lr=0.01, noise=0.1)
    a = Trainer(TrainConfig(**kw, bucket_mb=4.0), device=torch.device("cuda"))
    b = Trainer(TrainConfig(**kw, bucket_mb=1000.0), device=torch.device("cuda"))
    assert len(a._runtime.sync.chunks_of[0]) == 2 and len(b._runtime.sync.chunks_of[0]) == 1
    la = [h["loss"] for h in a.train()]
    lb = [h["loss"] for h in b.train()]
    assert la == lb and a.flat.fingerprint() == b.flat.fingerprint()


def _autograd_reference(t, round_bf16):
    m = t.flat.as_module().cuda()
    if round_bf16:
        for p in m.parameters():
            p.data = p.data.bfloat16().float()
    x32, y32 = t.shard.x.cuda(), t.shard.y.cuda()
    if round_bf16:
        x32 = x32.bfloat16().float()
    loss = torch.nn.functional.mse_loss(m(x32), y32)
    loss.backward()
    return m, loss.detach()


def test_fp32_model_beyond_the_simt_kernel_runs_on_tf32_tensor_cores():
    """`--dtype fp32` (the reference's only precision, ref.py:159) for a model far beyond the SIMT
    kernel's 64-wide limit: kind::tf32 GEMMs on the fp32 parameters, fp32 gradients and update."""
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    cfg = TrainConfig(quiet=True, nepochs=0, hidden=[512, 200], n_features=96, n_samples=1024, dtype="fp32",
                      data_source="randn", data_dist="local", scale="none", lr=0.01, noise=0.1)
    t = Trainer(cfg, device=torch.device("cuda"))
    assert t.compute.name == "cuda-tcgen05" and t.flat.shadow is None and t.flat.grads.dtype == torch.float32
    m, ref_loss = _autograd_reference(t, round_bf16=False)
    loss = t.compute.loss_and_grads(t.x, t.y)
    torch.testing.assert_close(loss, ref_loss, rtol=5e-3, atol=1e-4)
    for seg, p in zip(t.spec.segments, m.parameters()):
        got = t.flat._logical(t.flat.grads, seg)
        scale = p.grad.abs().max().item() + 1e-12
        # TF32 operands carry 10 mantissa bits (~1e-3 per product); three layers forward and back
        assert (got - p.grad).abs().max().item() <= 3e-2 * scale, seg.key
    first = float(loss.item())
    for _ in range(30):
        last = t.step(t.x, t.y)
    assert float(last.item()) < 0.9 * first
    # evaluation runs on the same kernels
    pred = t.predict(t.shard.x[:64])
    want = t.flat.as_module().cuda()(t.shard.x[:64].cuda())
    torch.testing.assert_close(pred, want, rtol=1e-2, atol=1e-2)
    t.close()


@pytest.mark.parametrize("nout", [5, 12])
def test_heads_with_5_to_15_outputs_are_padded_to_the_gemm_head(nout):
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    cfg = TrainConfig(quiet=True, nepochs=0, hidden=[256, 128], n_features=64, n_samples=1024, n_outputs=nout,
                      dtype="bf16", data_source="randn", data_dist="local", scale="none", lr=0.01, noise=0.1)
    t = Trainer(cfg, device=torch.device("cuda"))
    assert t.spec.sizes[-1] == (nout + 7) // 8 * 8 and t.spec.logical_sizes[-1] == nout
    m, ref_loss = _autograd_reference(t, round_bf16=True)
    loss = t.compute.loss_and_grads(t.x, t.y)
    torch.testing.assert_close(loss, ref_loss, rtol=2e-2, atol=1e-3)
    for seg, p in zip(t.spec.segments, m.parameters()):
        got = t.flat._logical(t.flat.grads, seg).float()
        scale = p.grad.abs().max().item() + 1e-12
        assert (got - p.grad).abs().max().item() <= 5e-2 * scale, seg.key
        full = t.flat._view(t.flat.grads, seg).float()                   # padding rows receive exactly zero
        assert float(full.abs().sum()) == pytest.approx(float(got.abs().sum()), rel=1e-6)
    assert tuple(t.state_dict()[f"layers.{2 * (t.spec.num_layers - 1)}.weight"].shape) == (nout, 128)
    assert t.predict(t.shard.x[:8]).shape == (8, nout)
    t.close()


def test_headline_shape_three_steps_against_bf16_autograd():
    """The benchmark's own configuration (12288 -> 4 x 4096 -> 1, 8192 rows, wgrad of layer 0 in row
    chunks, CTA-pair GEMMs with fused column sums, half-tile tail waves) for three steps: before every
    step the loss and all gradients are compared with an autograd model built from the trainer's
    current parameters (weights rounded to bf16, fp32 biases)."""
    from b200dp.config import TrainConfig
    from b200dp.engine import Trainer
    cfg = TrainConfig(quiet=True, nepochs=0, n_samples=8192, n_features=12288, hidden=[4096] * 4, dtype="bf16",
                      data_source="randn", data_dist="local", scale="none", init="device", lr=1e-4, noise=0.1)
    t = Trainer(cfg, device=torch.device("cuda"))
    assert len(t._runtime.sync.chunks_of[0]) > 1                          # chunked wgrad of the 12288-wide layer
    x32, y32 = t.x.float(), t.y
    losses = []
    for step in range(3):
        m = t.flat.as_module().cuda()
        with torch.no_grad():
            for p in m.parameters():
                if p.dim() == 2:
                    p.copy_(p.bfloat16().float())
        ref_loss = torch.nn.functional.mse_loss(m(x32), y32)
        ref_loss.backward()
        loss = t.compute.loss_and_grads(t.x, t.y)                          # gradients only (no update queued)
        torch.cuda.synchronize()
        assert float(loss.item()) == pytest.approx(float(ref_loss.item()), rel=2e-2), step
        for seg, p in zip(t.spec.segments, m.parameters()):
            got = t.flat._logical(t.flat.grads, seg).float()
            scale = p.grad.abs().max().item() + 1e-12
            # the autograd model keeps fp32 activations; ours rounds them to bf16 at every layer, and a
            # bias gradient is a sum of 8192 signed terms that mostly cancel: allow it 10 % of its peak
            # (five layers deep, every weight gradient passes through up to eight bf16 roundings)
            assert (got - p.grad).abs().max().item() <= 1e-1 * scale, (step, seg.key)
        del m, ref_loss
        losses.append(float(t.step(t.x, t.y).item()))                      # the real step: same gradients + update
    assert losses[2] < losses[0]
    t.close()


def test_pinned_staging_buffers_feed_training():
    """data.PinnedBuffer (native huge-page + cudaHostRegister allocator): torch sees it as pinned, so
    the per-step H2D copies of fit_host_batches are truly asynchronous; training from it matches
    training from pin_memory() buffers bit for bit."""
    from b200dp.config import TrainConfig
    from b200dp.data import PinnedBuffer, pinned_like
    from b200dp.engine import Trainer
    kw = dict(quiet=True, nepochs=0, hidden=[256, 128], n_features=64, n_samples=1024, dtype="bf16",
              data_source="randn", data_dist="local", scale="none", lr=0.01, noise=0.1)
    a = Trainer(TrainConfig(**kw), device=torch.device("cuda"))
    b = Trainer(TrainConfig(**kw), device=torch.device("cuda"))
    px, py = pinned_like(a.shard.x.to(torch.bfloat16)), pinned_like(a.shard.y)
    assert px.tensor.is_pinned() and py.tensor.is_pinned() and px.tensor.data_ptr() % (1 << 21) == 0
    la = a.fit_host_batches(iter([(px.tensor, py.tensor)] * 4), 4)
    lb = b.fit_host_batches(iter([(a.shard.x.to(torch.bfloat16).pin_memory(), a.shard.y.pin_memory())] * 4), 4)
    assert la == lb and a.flat.fingerprint() == b.flat.fingerprint()
    big = PinnedBuffer((3, 1000, 7), torch.float32)
    assert big.tensor.shape == (3, 1000, 7) and float(big.tensor.abs().sum()) == 0.0
    for buf in (px, py, big):
        buf.close()
