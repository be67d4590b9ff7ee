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
    bufs[0].fill_(1.0)                   # "gradients" that must NOT be applied
    ar.launch(a, 0, n, 0.1, 0.9, False)
    torch.cuda.synchronize()
    assert int(err.item()) == 2          # 1 + index of the missing peer
    # a timed-out launch changes nothing: no update from stale peer memory, no momentum, no epoch bump
    assert float(bufs[0].min()) == 1.0 and float(bufs[0].max()) == 1.0
    assert float(mom.abs().max()) == 0.0 and int(epochs.abs().max()) == 0


@pytest.mark.parametrize("world", [2, 3, 8])
def test_replicated_tail_keeps_fp32_biases_identical_on_every_rank(world):
    """bf16 two-shot keeps the fp32 master of a slice on its owner only - except the tail of the
    bucket (a layer's biases), which kernels read in fp32 on every rank: with replicate_tail the
    owner writes those fp32 values into all peers (round-1 bug: non-owners kept their initial bias)."""
    from b200dp.ops import allreduce_sgd as ar
    dev, n, tail, steps, lr, mu = "cuda", 128 * 24, 128 * 3, 4, 0.05, 0.9
    torch.manual_seed(11)
    p0 = torch.randn(n, device=dev)
    params = [p0.clone() for _ in range(world)]
    shadow = [p0.bfloat16() for _ in range(world)]
    grads = [torch.zeros(n, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    flags = [torch.zeros(ar.SYNC_MAX_BLOCKS * 8, device=dev, dtype=torch.int32) for _ in range(world)]
    mom = [torch.zeros(n, device=dev) for _ in range(world)]
    epochs = [torch.zeros(ar.SYNC_MAX_BLOCKS, device=dev, dtype=torch.int32) for _ in range(world)]
    err = [torch.zeros(1, device=dev, dtype=torch.int32) for _ in range(world)]
    args = [ar.make_args(r, world, [g.data_ptr() for g in grads], [p.data_ptr() for p in params],
                         [s_.data_ptr() for s_ in shadow], [f.data_ptr() for f in flags], mom[r].data_ptr(),
                         epochs[r].data_ptr(), err[r].data_ptr(), [1.0] * world, float(world), True, 1, False, 0, 5000)
            for r in range(world)]
    dargs = torch.zeros(ar.device_args_bytes() * world, device=dev, dtype=torch.uint8)
    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SGD([ref_p], lr=lr, momentum=mu)
    for s_ in range(steps):
        gs = [torch.randn(n, device=dev).bfloat16() for _ in range(world)]
        for r in range(world):
            grads[r].copy_(gs[r])
        ar.launch_emulated(args, 0, n, lr, mu, False, dargs, replicate_tail=tail)
        acc = torch.zeros(n, device=dev)
        for r in range(world):
            acc = acc + gs[r].float()
        ref_p.grad = acc / world
        opt.step()
    torch.cuda.synchronize()
    assert all(int(e.item()) == 0 for e in err)
    for r in range(world):
        torch.testing.assert_close(params[r][n - tail:], ref_p.detach()[n - tail:], rtol=2e-5, atol=2e-6)
        assert torch.equal(params[r][n - tail:], params[0][n - tail:])       # bit-equal fp32 tail everywhere
        assert torch.equal(shadow[r], shadow[0])
    # outside the tail a non-owner's master is untouched (still the initial value): sharded on purpose
    vn, units = 8, n // 8
    lo0, hi0 = 0, units // world * vn
    assert torch.equal(params[world - 1][lo0:hi0], p0[lo0:hi0])


def test_nvls_mode_refuses_unequal_weights():
    """The NVSwitch sums its inputs as they are; per-rank weights must be folded into the gradients
    by the caller (round-1 bug: they were silently ignored)."""
    from b200dp.ops import allreduce_sgd as ar
    from b200dp.ops._lib import B200dpError
    dev, n = "cuda", 1024
    bufs = [torch.zeros(n, device=dev) for _ in range(2)]
    flags = [torch.zeros(ar.SYNC_MAX_BLOCKS * 8, device=dev, dtype=torch.int32) for _ in range(2)]
    mom = torch.zeros(n, device=dev)
    epochs = torch.zeros(ar.SYNC_MAX_BLOCKS, device=dev, dtype=torch.int32)
    err = torch.zeros(1, device=dev, dtype=torch.int32)
    a = ar.make_args(0, 2, [b.data_ptr() for b in bufs], [b.data_ptr() for b in bufs], None,
                     [f.data_ptr() for f in flags], mom.data_ptr(), epochs.data_ptr(), err.data_ptr(),
                     [9.0, 8.0], 17.0, False, 2, True, 1, timeout_ms=50,
                     mc_grads=bufs[0].data_ptr(), mc_params=bufs[0].data_ptr())
    with pytest.raises(B200dpError, match="equal per-rank weights"):
        ar.launch(a, 0, n, 0.1, 0.9, False)


def test_sync_kernel_fits_beside_a_persistent_gemm_cta():
    """The fused sync runs on a second stream while backward's GEMMs run, and its cross-GPU barrier
    needs every one of its CTAs resident: one GEMM CTA + one sync CTA must fit an SM's 64 K registers.
    (A GEMM epilogue that grew to 180 registers/thread once pushed the sync kernels out: they ran only
    in the gaps between GEMMs and the 2-GPU step lost 0.2 ms.)"""
    import ctypes as C
    from b200dp.ops import _lib
    lib = _lib.lib()
    g, s = C.c_int(), C.c_int()
    lib.b200dp_overlap_register_budget.argtypes = [C.POINTER(C.c_int), C.POINTER(C.c_int)]
    _lib.check(lib.b200dp_overlap_register_budget(C.byref(g), C.byref(s)))
    assert 0 < g.value and 0 < s.value and g.value + s.value <= 65536, (g.value, s.value)
