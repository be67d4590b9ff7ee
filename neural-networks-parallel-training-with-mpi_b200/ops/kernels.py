"""Python faces of the SIMT kernels in libb200dp (tiny MLP, MSE, regression head, column sums,
local SGD, casts).  Thin, allocation-free on the hot path: callers own outputs and scratch."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import c_float, c_int, c_ll, c_void_p


class MlpDesc(C.Structure):
    _fields_ = [("num_layers", c_int), ("sizes", c_int * (_lib.MAX_LAYERS + 1)),
                ("w_off", c_ll * _lib.MAX_LAYERS), ("b_off", c_ll * _lib.MAX_LAYERS),
                ("flat_numel", c_ll)]


_lib.register("b200dp_tiny_mlp_fwd_bwd", [C.POINTER(MlpDesc), c_void_p, c_void_p, c_void_p, c_int, c_void_p,
                                          c_void_p, c_void_p, c_ll, c_void_p])
_lib.register("b200dp_mse_loss_grad", [c_void_p, c_int, c_void_p, c_ll, c_ll, c_void_p, c_int, c_void_p, c_float,
                                       c_void_p, c_ll, c_void_p])
_lib.register("b200dp_head_fwd_mse", [c_void_p, c_ll, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_int,
                                      c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_ll, c_void_p])
_lib.register("b200dp_head_bwd", [c_void_p, c_void_p, c_void_p, c_ll, c_int, c_int, c_int, c_int, c_void_p, c_ll,
                                  c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_ll, c_void_p])
_lib.register("b200dp_colsum", [c_void_p, c_int, c_ll, c_int, c_int, c_void_p, c_int, c_void_p, c_ll, c_void_p])
_lib.register("b200dp_sgd_momentum", [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_ll, c_float, c_float,
                                      c_float, c_int, c_void_p])
_lib.register("b200dp_sgd_momentum_ex", [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_ll, c_float, c_float,
                                         c_float, c_int, c_int, c_void_p])
_lib.register("b200dp_cast_f32_bf16", [c_void_p, c_void_p, c_ll, c_void_p])
_lib.register("b200dp_transpose_f32", [c_void_p, c_ll, c_int, c_int, c_void_p, c_ll, c_void_p])
_lib.register("b200dp_gather_rows", [c_void_p, c_ll, c_void_p, c_int, c_ll, c_void_p, c_void_p])


def _fix_restypes(lib):
    for name in ("b200dp_tiny_mlp_scratch_bytes", "b200dp_mse_scratch_bytes", "b200dp_head_scratch_bytes",
                 "b200dp_colsum_scratch_bytes", "b200dp_sync_args_device_bytes"):
        getattr(lib, name).restype = c_ll
    lib.b200dp_tiny_mlp_scratch_bytes.argtypes = [C.POINTER(MlpDesc), c_int]
    lib.b200dp_mse_scratch_bytes.argtypes = [c_ll]
    lib.b200dp_head_scratch_bytes.argtypes = [c_int, c_int, c_int]
    lib.b200dp_colsum_scratch_bytes.argtypes = [c_int, c_int]
    return lib


def L():
    lib = _lib.lib()
    if not getattr(lib, "_b200dp_kernels_ready", False):
        _fix_restypes(lib)
        lib._b200dp_kernels_ready = True
    return lib


def mlp_desc(spec) -> MlpDesc:
    d = MlpDesc()
    d.num_layers = spec.num_layers
    for i, s in enumerate(spec.sizes):
        d.sizes[i] = s
    for l in range(spec.num_layers):
        w, b = spec.layer_segments(l)
        d.w_off[l], d.b_off[l] = w.offset, b.offset
    d.flat_numel = spec.flat_numel
    return d


def scratch(nbytes: int, device) -> torch.Tensor:
    """Zero-initialised scratch (the ticket counters the kernels keep in it must start at zero)."""
    return torch.zeros(max(int(nbytes), 256), dtype=torch.uint8, device=device)


# ------------------------------------------------------------------------------------------ tiny
def tiny_mlp_scratch_bytes(spec, rows: int) -> int:
    n = L().b200dp_tiny_mlp_scratch_bytes(C.byref(mlp_desc(spec)), rows)
    if n < 0:
        raise _lib.B200dpError(_lib.lib().b200dp_last_error().decode())
    return int(n)


def tiny_mlp_fwd_bwd(desc: MlpDesc, params, x, y, grads, loss, scratch_buf, stream=None):
    rows = x.shape[0]
    _lib.check(L().b200dp_tiny_mlp_fwd_bwd(C.byref(desc), params.data_ptr(), x.data_ptr(), y.data_ptr(), rows,
                                           grads.data_ptr(), loss.data_ptr(), scratch_buf.data_ptr(),
                                           scratch_buf.numel(), _lib.stream_ptr(stream)))
    _lib.count()


# ------------------------------------------------------------------------------------------ loss
def mse_scratch_bytes(numel: int) -> int:
    return int(L().b200dp_mse_scratch_bytes(numel))


def mse_loss_grad(pred, target, dy, loss, scratch_buf, stream=None, grad_scale: float = 1.0, denom: int = 0):
    """``loss = sum((pred-target)^2) / n``, ``dy = grad_scale * 2 (pred-target) / n``; ``n = denom`` when
    given (padded outputs), else ``numel``."""
    assert pred.is_contiguous() and target.is_contiguous() and dy.is_contiguous()
    assert target.dtype == torch.float32 and pred.numel() == target.numel() == dy.numel()
    _lib.check(L().b200dp_mse_loss_grad(pred.data_ptr(), int(pred.dtype == torch.bfloat16), target.data_ptr(),
                                        pred.numel(), denom, dy.data_ptr(), int(dy.dtype == torch.bfloat16),
                                        loss.data_ptr(), grad_scale, scratch_buf.data_ptr(), scratch_buf.numel(),
                                        _lib.stream_ptr(stream)))
    _lib.count()


def head_scratch_bytes(rows: int, K: int, nout: int) -> int:
    return int(L().b200dp_head_scratch_bytes(rows, K, nout))


def head_fwd_mse(H, W, bias, target, yhat, dy, loss, scratch_buf, stream=None, grad_scale: float = 1.0):
    """H, W: bf16 (bf16 training) or fp32 (fp32 training); everything else fp32."""
    rows, K = H.shape
    nout = W.shape[0]
    assert H.dtype == W.dtype and H.dtype in (torch.bfloat16, torch.float32)
    _lib.check(L().b200dp_head_fwd_mse(H.data_ptr(), H.stride(0), W.data_ptr(), int(H.dtype == torch.float32),
                                       bias.data_ptr(), target.data_ptr(), rows, K, nout, yhat.data_ptr(),
                                       dy.data_ptr(), loss.data_ptr(), grad_scale, scratch_buf.data_ptr(),
                                       scratch_buf.numel(), _lib.stream_ptr(stream)))
    _lib.count()


def head_bwd(dy, W, H, dH: Optional[torch.Tensor], dW, db, scratch_buf, stream=None,
             db_prev: Optional[torch.Tensor] = None):
    """``db_prev`` (needs ``dH``): receives the column sums of dH = the bias gradient of the layer
    below, fused into the same pass."""
    rows, K = H.shape
    nout = W.shape[0]
    assert dW.dtype == db.dtype and H.dtype == W.dtype and (dH is None or dH.dtype == H.dtype)
    assert db_prev is None or (dH is not None and db_prev.dtype == dW.dtype and db_prev.numel() >= K)
    _lib.check(L().b200dp_head_bwd(dy.data_ptr(), W.data_ptr(), H.data_ptr(), H.stride(0),
                                   int(H.dtype == torch.float32), rows, K, nout,
                                   _lib.ptr(dH), dH.stride(0) if dH is not None else 0, dW.data_ptr(),
                                   db.data_ptr(), _lib.ptr(db_prev), int(dW.dtype == torch.bfloat16),
                                   scratch_buf.data_ptr(), scratch_buf.numel(), _lib.stream_ptr(stream)))
    _lib.count(2)


def colsum_scratch_bytes(rows: int, n: int) -> int:
    return int(L().b200dp_colsum_scratch_bytes(rows, n))


def colsum(A, out, scratch_buf, stream=None):
    rows, n = A.shape
    _lib.check(L().b200dp_colsum(A.data_ptr(), int(A.dtype == torch.float32), A.stride(0), rows, n, out.data_ptr(),
                                 int(out.dtype == torch.bfloat16), scratch_buf.data_ptr(),
                                 scratch_buf.numel(), _lib.stream_ptr(stream)))
    _lib.count(2)


# ------------------------------------------------------------------------------------------ optim
def sgd_momentum(params, grads, momentum, shadow, lr, mu, first_step, grad_scale=1.0, stream=None, max_blocks=0):
    _lib.check(L().b200dp_sgd_momentum_ex(params.data_ptr(), grads.data_ptr(), int(grads.dtype == torch.bfloat16),
                                          momentum.data_ptr(), _lib.ptr(shadow), params.numel(), lr, mu,
                                          grad_scale, int(first_step), max_blocks, _lib.stream_ptr(stream)))
    _lib.count()


def cast_f32_bf16(src, dst, stream=None):
    _lib.check(L().b200dp_cast_f32_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _lib.stream_ptr(stream)))
    _lib.count()


def transpose_f32(src, dst, stream=None):
    """``dst[c, r] = src[r, c]``; both 2-D fp32 with unit inner stride (row strides are honoured)."""
    rows, cols = src.shape
    assert src.dtype == dst.dtype == torch.float32 and tuple(dst.shape) == (cols, rows)
    assert src.stride(1) == 1 and dst.stride(1) == 1
    _lib.check(L().b200dp_transpose_f32(src.data_ptr(), src.stride(0), rows, cols, dst.data_ptr(), dst.stride(0),
                                        _lib.stream_ptr(stream)))
    _lib.count()
    return dst


def gather_rows(src, idx, dst=None, stream=None):
    """``dst[r] = src[idx[r]]`` for a 2-D row-major tensor whose rows are a multiple of 16 bytes; other
    shapes fall back to torch indexing (tiny rows such as the reference's 2 features)."""
    row_bytes = src.shape[1] * src.element_size()
    if src.dim() != 2 or row_bytes % 16 or (src.stride(0) * src.element_size()) % 16 or src.stride(1) != 1:
        out = src.index_select(0, idx)
        if dst is not None:
            dst.copy_(out)
            return dst
        return out
    if dst is None:
        dst = torch.empty(idx.numel(), src.shape[1], dtype=src.dtype, device=src.device)
    _lib.check(L().b200dp_gather_rows(src.data_ptr(), src.stride(0) * src.element_size(), idx.data_ptr(),
                                      idx.numel(), row_bytes, dst.data_ptr(), _lib.stream_ptr(stream)))
    _lib.count()
    return dst
