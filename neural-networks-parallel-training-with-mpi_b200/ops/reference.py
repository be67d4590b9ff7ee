"""Plain-PyTorch statement of every op the CUDA kernels implement.

Two uses: (1) the oracle the GPU numerics tests compare against (always evaluated in fp32 or
fp64), (2) the ``torch`` backend of the trainer, which is what runs on a box without a GPU (CPU
plumbing config of BASELINE.json).  On a GPU box the trainer refuses to fall back to this file.

Math per SURVEY §2.4 (ATen ops the reference's step triggers, ref.py:170-176, 211).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch


def linear_fwd(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], relu: bool) -> torch.Tensor:
    """``relu?(x @ w.T + b)``   (addmm + clamp_min; ref.py:42-44)."""
    y = x @ w.t()
    if b is not None:
        y = y + b
    return torch.relu(y) if relu else y


def linear_dgrad(dy: torch.Tensor, w: torch.Tensor, relu_src: Optional[torch.Tensor]) -> torch.Tensor:
    """``dx = (dy @ w) * [relu_src > 0]``; ``relu_src`` is the layer's input activation (itself a
    ReLU output), so the mask is that of the ReLU feeding this layer (threshold_backward)."""
    dx = dy @ w
    if relu_src is not None:
        dx = dx * (relu_src > 0).to(dx.dtype)
    return dx


def linear_wgrad(dy: torch.Tensor, x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``dW = dy.T @ x``, ``db = sum_rows(dy)``."""
    return dy.t() @ x, dy.sum(dim=0)


def mse_loss_grad(yhat: torch.Tensor, y: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``mean((yhat-y)^2)`` over all elements and its gradient ``2 (yhat-y) / numel``
    (``nn.MSELoss()`` default reduction, ref.py:94,173)."""
    d = yhat - y
    n = d.numel()
    return (d * d).sum() / n, d * (2.0 / n)


def sgd_momentum_(p: torch.Tensor, g: torch.Tensor, buf: torch.Tensor, lr: float, mu: float,
                  first_step: bool) -> None:
    """torch.optim.SGD with momentum, dampening 0, no Nesterov, no weight decay (ref.py:91,211):
    first step ``buf = g``; later ``buf = mu*buf + g``; ``p -= lr*buf``."""
    g = g.to(torch.float32)
    if first_step:
        buf.copy_(g)
    else:
        buf.mul_(mu).add_(g)
    p.add_(buf, alpha=-lr)


def reduce_grads(per_rank, weights, divisor: float) -> torch.Tensor:
    """``(sum_r weights[r] * g_r) / divisor`` with fp32 accumulation in rank order 0..N-1 - the
    summation order of the reference's root loop (ref.py:193-197)."""
    acc = torch.zeros_like(per_rank[0], dtype=torch.float32)
    for w, g in zip(weights, per_rank):
        g32 = g.to(torch.float32)
        acc += g32 if w == 1.0 else g32 * w
    return acc / divisor


def mlp_forward(flat, x: torch.Tensor, keep: bool = False):
    """Forward through a ``FlatMLP``.  Returns the output (and the list of layer inputs)."""
    acts = [x]
    h = x
    L = flat.spec.num_layers
    for l in range(L):
        w, b = flat.compute_weight(l), flat.compute_bias(l)
        h = linear_fwd(h.to(w.dtype), w, b.to(w.dtype), relu=(l < L - 1))
        if keep:
            acts.append(h)
    return (h, acts) if keep else h


def mlp_loss_and_grads(flat, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """One local forward + MSE + backward (ref.py:170-176) with hand-scheduled ops - the same
    schedule the CUDA backend launches.  Gradients are written into ``flat.grads`` (overwrite,
    i.e. ``zero_grad`` is implicit; ref.py:167).  Returns the scalar loss (fp32)."""
    spec = flat.spec
    L = spec.num_layers
    if x.shape[0] == 0:
        flat.grads.zero_()
        return torch.full((), float("nan"), device=x.device)
    out, acts = mlp_forward(flat, x, keep=True)
    loss, dy = mse_loss_grad(out.to(torch.float32), y.to(torch.float32))
    dy = dy.to(out.dtype)
    for l in reversed(range(L)):
        w = flat.compute_weight(l)
        xin = acts[l].to(w.dtype)
        dw, db = linear_wgrad(dy.to(torch.float32), xin.to(torch.float32))
        flat.weight(l, flat.grads).copy_(dw)
        flat.bias(l, flat.grads).copy_(db)
        if l > 0:
            dy = linear_dgrad(dy, w, acts[l])
    return loss
