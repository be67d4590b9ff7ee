"""Single-process emulation of the reference's math (SURVEY §4.3 recipe), written with autograd and
``torch.optim.SGD`` so it shares no code with the framework's hand-scheduled backward.

seed-0 ``nn.Sequential(Linear(2,3), ReLU, Linear(3,1))`` -> ``make_regression(16,2,noise=1,rs=42)`` ->
contiguous shards (+1-row rule) -> per-shard StandardScaler on X -> fp32 -> per-rank full-shard MSE ->
mean of grads (uniform) or sample-weighted -> SGD(lr, momentum)."""
import numpy as np
import torch
from sklearn.datasets import make_regression
from sklearn.preprocessing import StandardScaler
from torch import nn


def emulate(nprocs, nepochs=3, lr=0.001, momentum=0.9, weighting="uniform", scale="per_shard",
            n_samples=16, n_features=2, hidden=(3,), seed=0):
    torch.manual_seed(seed)
    mods, sizes = [], [n_features, *hidden, 1]
    for i in range(len(sizes) - 1):
        mods.append(nn.Linear(sizes[i], sizes[i + 1]))
        if i < len(sizes) - 2:
            mods.append(nn.ReLU())
    model = nn.Sequential(*mods)
    X, y = make_regression(n_samples=n_samples, n_features=n_features, noise=1, random_state=42)
    base, res = divmod(n_samples, nprocs)
    counts = [base + 1 if p < res else base for p in range(nprocs)]
    displ = np.concatenate([[0], np.cumsum(counts)[:-1]])
    shards = []
    if scale == "global":
        X = StandardScaler().fit_transform(X)
    for r in range(nprocs):
        xs, ys = X[displ[r]:displ[r] + counts[r]], y[displ[r]:displ[r] + counts[r]]
        if scale == "per_shard" and len(xs):
            xs = StandardScaler().fit_transform(xs)
        shards.append((torch.from_numpy(xs).float(), torch.from_numpy(ys).float().reshape(-1, 1)))
    opt = torch.optim.SGD(model.parameters(), lr=lr, momentum=momentum)
    loss_fn = nn.MSELoss()
    params = list(model.parameters())
    losses = []
    for _ in range(nepochs):
        per_rank, ep = [], []
        for xs, ys in shards:
            opt.zero_grad()
            if len(xs) == 0:
                per_rank.append([torch.zeros_like(p) for p in params]); ep.append(float("nan")); continue
            loss = loss_fn(model(xs), ys)
            loss.backward()
            per_rank.append([p.grad.clone() for p in params])
            ep.append(loss.item())
        for k, p in enumerate(params):
            if weighting == "uniform":
                acc = 0
                for r in range(nprocs):
                    acc = acc + per_rank[r][k]
                p.grad = acc / nprocs
            else:
                acc = 0
                for r in range(nprocs):
                    acc = acc + per_rank[r][k] * float(counts[r])
                p.grad = acc / float(sum(counts))
        opt.step()
        losses.append(ep)
    return losses, model
