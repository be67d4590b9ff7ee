"""Fused all-reduce + SGD: a whole world emulated on one GPU in a single cooperative launch
(the blocks really wait on each other through the signal pads), checked against
(reduce in rank order -> torch.optim.SGD) for one-shot and two-shot, fp32 and bf16 gradients,
uniform and sample weighting, several consecutive steps with no host sync in between."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(world, n, mode, gdt, weights, div, steps=3, lr=0.05, mu=0.9, max_blocks=0):
    from b200dp.ops import allreduce_sgd as ar
    dev = "cuda"
    torch.manual_seed(world * 1000 + n % 997)
    bf16 = gdt == torch.bfloat16
    p0 = torch.randn(n, device=dev)
    params = [p0.clone() for _ in range(world)]
    shadow = [p0.bfloat16() for _ in range(world)] if bf16 else None
    grads = [torch.zeros(n, device=dev, dtype=gdt) for _ in range(world)]
    flags = [torch.zeros(ar.SYNC_MAX_BLOCKS * 8, device=dev, dtype=torch.int32) for _ in range(world)]
    mom = [torch.zeros(n, device=dev) for _ in range(world)]
    epochs = [torch.zeros(ar.SYNC_MAX_BLOCKS, device=dev, dtype=torch.int32) for _ in range(world)]
    err = [torch.zeros(1, device=dev, dtype=torch.int32) for _ in range(world)]
    push_master = not bf16
    args = [ar.make_args(r, world, [g.data_ptr() for g in grads], [p.data_ptr() for p in params],
                         [s.data_ptr() for s in shadow] if bf16 else None, [f.data_ptr() for f in flags],
                         mom[r].data_ptr(), epochs[r].data_ptr(), err[r].data_ptr(), weights, div, bf16,
                         mode, push_master, max_blocks, 5000) for r in range(world)]
    dargs = torch.zeros(ar.device_args_bytes() * world, device=dev, dtype=torch.uint8)

    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SGD([ref_p], lr=lr, momentum=mu)
    all_g = [[torch.randn(n, device=dev).to(gdt) for _ in range(world)] for _ in range(steps)]
    for s in range(steps):
        for r in range(world):
            grads[r].copy_(all_g[s][r])
        ar.launch_emulated(args, 0, n, lr, mu, False, dargs)      # momentum starts at zero: first_step not needed
    torch.cuda.synchronize()
    for s in range(steps):
        acc = torch.zeros(n, device=dev)
        for r in range(world):
            g = all_g[s][r].float()
            acc = acc + g if weights[r] == 1.0 else torch.addcmul(acc, g, torch.tensor(weights[r], device=dev))
        ref_p.grad = acc / div
        opt.step()
    assert all(int(e.item()) == 0 for e in err)
    vn = 8 if bf16 else 4
    units = n // vn
    for r in range(world):
        if mode == 0 or push_master:
            torch.testing.assert_close(params[r], ref_p.detach(), rtol=2e-5, atol=2e-6)
        else:   # bf16 two-shot: the fp32 master of a slice lives on its owner
            lo, hi = units * r // world * vn, units * (r + 1) // world * vn
            torch.testing.assert_close(params[r][lo:hi], ref_p.detach()[lo:hi], rtol=2e-5, atol=2e-6)
        if bf16:
            torch.testing.assert_close(shadow[r].float(), ref_p.detach().bfloat16().float(), rtol=1e-2, atol=1e-6)
    # replicas bitwise identical (what they compute with)
    ref_rep = shadow if bf16 else params
    for r in range(1, world):
        assert torch.equal(ref_rep[0], ref_rep[r])
    assert int(epochs[0][0].item()) == 2 * steps


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("gdt", [torch.float32, torch.bfloat16])
def test_fused_sync_uniform(world, mode, gdt):
    _run(world, 128 * 40, mode, gdt, [1.0] * world, float(world))


def test_fused_sync_reference_sized_bucket():
    # the reference's 13 parameters live in 4 padded segments of 128 floats
    _run(2, 512, 0, torch.float32, [1.0, 1.0], 2.0, steps=5)


def test_fused_sync_sample_weights():
    _run(3, 128 * 9, 1, torch.float32, [6.0, 5.0, 5.0], 16.0)
    _run(3, 128 * 9, 0, torch.bfloat16, [6.0, 5.0, 5.0], 16.0)


@pytest.mark.parametrize("mode", [0, 1])
def test_fused_sync_large_multiblock(mode):
    _run(4, 1 << 20, mode, torch.bfloat16, [1.0] * 4, 4.0, steps=2, max_blocks=16)


def test_watchdog_flags_a_peer_that_never_arrives():
    """Failure detection: a dead rank hangs the reference forever at comm.gather /
    comm.recv (ref.py:185,203).  Here the spin is bounded: rank 0 of a 2-rank world is launched
    alone (its peer never shows up), the kernel gives up after timeout_ms and raises the flag."""
    from b200dp.ops import allreduce_sgd as ar
    dev, n = "cuda", 1024
    bufs = [torch.zeros(n, device=dev) for _ in range(2)]
    flags = [torch.zeros(ar.SYNC_MAX_BLOCKS * 8, device=dev, dtype=torch.int32) for _ in range(2)]
    mom = torch.zeros(n, device=dev)
    epochs = torch.zeros(ar.SYNC_MAX_BLOCKS, device=dev, dtype=torch.int32)
    err = torch.zeros(1, device=dev, dtype=torch.int32)
    a = ar.make_args(0, 2, [b.data_ptr() for b in bufs], [b.data_ptr() for b in bufs], None,
                     [f.data_ptr() for f in flags], mom.data_ptr(), epochs.data_ptr(), err.data_ptr(),
                     [1.0, 1.0], 2.0, False, 0, True, 1, timeout_ms=50)
    ar.launch(a, 0, n, 0.1, 0.9, False)
    torch.cuda.synchronize()
    assert int(err.item()) == 2          # 1 + index of the missing peer
