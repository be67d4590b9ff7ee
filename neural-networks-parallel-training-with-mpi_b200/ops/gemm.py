"""Python face of the tcgen05 Linear GEMMs (``csrc/gemm_tcgen05.cu``)."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib


def _check_operand(t: torch.Tensor, name: str, dtype):
    if t.dtype != dtype or not t.is_cuda or t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{name} must be a 2-D {dtype} CUDA tensor with unit inner stride")


def colsum_scratch_bytes(M: int, N: int) -> int:
    return int(_lib.lib().b200dp_gemm_colsum_scratch_bytes(M, N))


def gemm(A: torch.Tensor, B: torch.Tensor, out: torch.Tensor, *, a_mn_major=False, b_mn_major=False,
         bias: Optional[torch.Tensor] = None, relu=False, mask_src: Optional[torch.Tensor] = None,
         tile_n=0, max_ctas=0, cta_group=0, split_tail=0, colsum_out: Optional[torch.Tensor] = None,
         colsum_scratch: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
    """``out[M,N] = epi(A[M,K] @ B[N,K]^T)``.  ``a_mn_major``: A is given as its transpose ``[K,M]``;
    ``b_mn_major``: B is given as ``[K,N]``.  bf16 operands run ``kind::f16``, fp32 operands
    ``kind::tf32`` (fp32 output).  ``colsum_out[N]``: also receives the column sums of the result
    (fused bias gradient; needs ``colsum_scratch`` of ``colsum_scratch_bytes(M, N)`` zero-initialised
    bytes)."""
    fp32 = A.dtype == torch.float32
    _check_operand(A, "A", A.dtype if fp32 else torch.bfloat16)
    _check_operand(B, "B", A.dtype)
    if mask_src is not None and mask_src.dtype != A.dtype:
        raise ValueError("mask_src must have the operands' dtype")
    M, K = (A.shape[1], A.shape[0]) if a_mn_major else (A.shape[0], A.shape[1])
    N, Kb = (B.shape[1], B.shape[0]) if b_mn_major else (B.shape[0], B.shape[1])
    if K != Kb:
        raise ValueError(f"reduction sizes differ: {K} vs {Kb}")
    if tuple(out.shape) != (M, N) or out.stride(1) != 1 or out.dtype not in (torch.bfloat16, torch.float32):
        raise ValueError("out must be [M,N] bf16/fp32 with unit inner stride")
    a = _lib.GemmArgs(M=M, N=N, K=K, A=A.data_ptr(), lda=A.stride(0), a_mn_major=int(a_mn_major),
                      B=B.data_ptr(), ldb=B.stride(0), b_mn_major=int(b_mn_major),
                      D=out.data_ptr(), ldd=out.stride(0), d_fp32=int(out.dtype == torch.float32),
                      bias=_lib.ptr(bias), relu=int(relu), mask_src=_lib.ptr(mask_src),
                      ldmask=mask_src.stride(0) if mask_src is not None else 0,
                      tile_n=tile_n, max_ctas=max_ctas, cta_group=cta_group, split_tail=split_tail,
                      colsum_out=_lib.ptr(colsum_out),
                      colsum_out_bf16=int(colsum_out is not None and colsum_out.dtype == torch.bfloat16),
                      colsum_scratch=_lib.ptr(colsum_scratch),
                      colsum_scratch_bytes=colsum_scratch.numel() * colsum_scratch.element_size()
                      if colsum_scratch is not None else 0)
    fn = _lib.lib().b200dp_gemm_tf32 if fp32 else _lib.lib().b200dp_gemm_bf16
    _lib.check(fn(C.byref(a), _lib.stream_ptr(stream)))
    _lib.count()
    return out


def linear_fwd(x, w, bias, relu, out=None, **kw):
    """``relu?(x @ w.T + bias)``; x [M,K] bf16, w [N,K] bf16, bias fp32 [N]."""
    if out is None:
        out = torch.empty(x.shape[0], w.shape[0], dtype=x.dtype, device=x.device)
    return gemm(x, w, out, bias=bias, relu=relu, **kw)


def linear_dgrad(dy, w, relu_src, out=None, **kw):
    """``(dy @ w) * [relu_src > 0]``; dy [M,N] bf16, w [N,K] bf16 (used MN-major, no transpose copy)."""
    if out is None:
        out = torch.empty(dy.shape[0], w.shape[1], dtype=dy.dtype, device=dy.device)
    return gemm(dy, w, out, b_mn_major=True, mask_src=relu_src, **kw)


def linear_wgrad(dy, x, out=None, **kw):
    """``dy.T @ x`` -> [N,K]; both operands used MN-major straight from their row-major storage."""
    if out is None:
        out = torch.empty(dy.shape[1], x.shape[1], dtype=torch.float32, device=dy.device)
    return gemm(dy, x, out, a_mn_major=True, b_mn_major=True, **kw)
