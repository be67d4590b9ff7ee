"""The allocation-free entry point ``b200dp_tiny_mlp_fwd_bwd_noalloc``.

Some users of libb200dp cannot allocate device memory at the point where they train: a host loop
with a hard latency budget (cudaMalloc synchronises the device), code hosted inside another
framework whose allocator owns the GPU, a job that received a fixed memory budget at start-up.
For them the library carries the kernel's cross-block scratch itself, in a built-in device array.

This test drives the entry point through the C ABI exactly as such a caller would (ctypes, raw
device pointers) and checks, on two calls in a row with different data, that it produces what the
scratch-taking entry point ``b200dp_tiny_mlp_fwd_bwd`` produces - the second call proves that the
leftovers of the first one in the shared array do no harm.  It also checks that a problem whose
scratch need exceeds the built-in array is refused with a readable error instead of overrunning.

Collected last on purpose (file name): if it fails under ``pytest -x`` everything else has run.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

NOALLOC_SCRATCH_BYTES = 1 << 20


def _setup(sizes, rows, seed):
    from b200dp.models.mlp import FlatMLP, MLPSpec
    from b200dp.ops import kernels as K
    torch.manual_seed(seed)
    flat = FlatMLP(MLPSpec(sizes), "cuda")
    flat.init_like_reference(seed)
    x = torch.randn(rows, sizes[0], device="cuda")
    y = torch.randn(rows, sizes[-1], device="cuda")
    return flat, K.mlp_desc(flat.spec), x, y


def _lib():
    from b200dp.ops import _lib, kernels as K
    lib = K.L()
    fn = lib.b200dp_tiny_mlp_fwd_bwd_noalloc
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    return lib, _lib


def test_noalloc_matches_scratch_entry_point_twice_in_a_row():
    from b200dp.ops import kernels as K
    lib, _l = _lib()
    stream = torch.cuda.current_stream().cuda_stream
    sizes = [2, 3, 1]
    for call, (rows, seed) in enumerate([(100000, 11), (70001, 12)]):      # many blocks: scratch is really used
        flat, desc, x, y = _setup(sizes, rows, seed)
        need = K.tiny_mlp_scratch_bytes(flat.spec, rows)
        assert 256 < need <= NOALLOC_SCRATCH_BYTES
        # existing entry point, caller-allocated scratch
        want_g = torch.zeros_like(flat.grads)
        want_l = torch.zeros((), device="cuda")
        K.tiny_mlp_fwd_bwd(desc, flat.params, x, y, want_g, want_l, K.scratch(need, "cuda"))
        # new entry point: inputs and outputs only
        got_g = torch.full_like(flat.grads, float("nan"))
        got_l = torch.full((), float("nan"), device="cuda")
        # Free device memory as the DRIVER sees it: torch.cuda.memory_allocated() only counts torch's own
        # caching allocator and would not notice a cudaMalloc made inside the library.  The first call on a
        # device may still pay for lazy module loading (the library's kernels + its built-in scratch array
        # become resident), so the strict comparison is made on the second call of this process.
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()
        torch_before = torch.cuda.memory_allocated()
        rc = lib.b200dp_tiny_mlp_fwd_bwd_noalloc(C.byref(desc), flat.params.data_ptr(), x.data_ptr(), y.data_ptr(),
                                                 rows, got_g.data_ptr(), got_l.data_ptr(), stream)
        assert rc == 0, lib.b200dp_last_error().decode()
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert torch.cuda.memory_allocated() == torch_before
        if call > 0:
            assert free_after >= free_before, (free_before, free_after)
        torch.testing.assert_close(got_l, want_l, rtol=1e-5, atol=1e-6)
        for seg in flat.spec.segments:
            sl = slice(seg.offset, seg.offset + seg.numel)
            torch.testing.assert_close(got_g[sl], want_g[sl], rtol=2e-4, atol=1e-6)
        # and both agree with autograd
        m = flat.as_module().cuda().double()
        ref = torch.nn.functional.mse_loss(m(x.double()), y.double())
        torch.testing.assert_close(got_l.double(), ref.detach(), rtol=1e-5, atol=1e-6)


def test_noalloc_single_block_needs_no_scratch():
    lib, _l = _lib()
    flat, desc, x, y = _setup([2, 3, 1], 16, 3)
    g = torch.zeros_like(flat.grads)
    l = torch.zeros((), device="cuda")
    rc = lib.b200dp_tiny_mlp_fwd_bwd_noalloc(C.byref(desc), flat.params.data_ptr(), x.data_ptr(), y.data_ptr(), 16,
                                             g.data_ptr(), l.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.b200dp_last_error().decode()
    m = flat.as_module().cuda()
    torch.testing.assert_close(l, torch.nn.functional.mse_loss(m(x), y).detach(), rtol=1e-5, atol=1e-6)


def test_noalloc_refuses_problems_beyond_its_scratch():
    from b200dp.ops import kernels as K
    lib, _l = _lib()
    sizes, rows = [7, 64, 64, 3], 100000                 # ~5k parameters x up to 148 blocks > 1 MiB
    flat, desc, x, y = _setup(sizes, rows, 5)
    assert K.tiny_mlp_scratch_bytes(flat.spec, rows) > NOALLOC_SCRATCH_BYTES
    g = torch.zeros_like(flat.grads)
    l = torch.zeros((), device="cuda")
    rc = lib.b200dp_tiny_mlp_fwd_bwd_noalloc(C.byref(desc), flat.params.data_ptr(), x.data_ptr(), y.data_ptr(),
                                             rows, g.data_ptr(), l.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.b200dp_last_error().decode()
    assert "scratch" in msg and str(NOALLOC_SCRATCH_BYTES) in msg
    torch.cuda.synchronize()
    assert float(g.abs().sum()) == 0.0                   # nothing was launched
