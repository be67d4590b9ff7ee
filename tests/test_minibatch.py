"""Mini-batch mode (`--use-batch-size`, an extension: the reference always trains on the whole local
shard, ref.py:146) on uneven shards, and exact resume of shuffled mini-batch runs."""
import math

import pytest
import torch

from b200dp.config import TrainConfig
from b200dp.data import BatchPlan
from b200dp.engine import Trainer
from b200dp.parallel.comm import run_threads


def test_every_rank_runs_the_same_number_of_steps():
    # 17 rows on 2 ranks = 9 + 8 rows, batch 4: the larger shard needs 3 steps, so both ranks run 3
    a = BatchPlan(9, 4, True, seed=1, max_rows=9)
    b = BatchPlan(8, 4, True, seed=2, max_rows=9)
    assert a.steps_per_epoch() == b.steps_per_epoch() == 3
    ea, eb = list(a.epoch()), list(b.epoch())
    assert [len(i) for i in ea] == [3, 3, 3] and [len(i) for i in eb] == [3, 3, 2]
    assert sorted(torch.cat(ea).tolist()) == list(range(9)) and sorted(torch.cat(eb).tolist()) == list(range(8))
    # a rank with fewer rows than steps yields empty batches instead of stopping early
    c = BatchPlan(2, 4, False, seed=3, max_rows=12)
    assert [len(i) for i in c.epoch()] == [1, 1, 0]
    # no rows at all: the trainer runs the (empty) steps itself
    d = BatchPlan(0, 4, True, seed=4, max_rows=12)
    assert d.steps_per_epoch() == 3 and list(d.epoch()) == []
    # full-shard mode is one step with no permutation (reference behaviour)
    e = BatchPlan(9, None, True, seed=5, max_rows=9)
    assert e.steps_per_epoch() == 1 and list(e.epoch()) == [None]


def test_permutation_is_a_pure_function_of_seed_and_epoch():
    a, b = BatchPlan(32, 8, True, seed=7), BatchPlan(32, 8, True, seed=7)
    first = [torch.cat(list(a.epoch())).tolist() for _ in range(3)]
    b.set_epoch(2)                                   # a resumed run jumps straight to epoch 2
    assert torch.cat(list(b.epoch())).tolist() == first[2]
    assert first[0] != first[1] != first[2]


@pytest.mark.parametrize("weighting", ["uniform", "samples"])
def test_minibatches_on_uneven_shards_do_not_hang_and_keep_replicas_identical(weighting):
    cfg = TrainConfig(quiet=True, nepochs=3, n_samples=17, batch_size=4, use_batch_size=True,
                      grad_weighting=weighting, data_dist="local", scale="global", lr=3e-4)

    def fn(comm):
        t = Trainer(cfg, comm, torch.device("cpu"))
        hist = t.train()
        return [h["loss"] for h in hist], t.flat.fingerprint(), t.flat.step_count, t.plan.steps_per_epoch()

    res = run_threads(2, fn)
    assert res[0][1] == res[1][1]                                 # bitwise identical replicas
    assert res[0][2] == res[1][2] == 9 and res[0][3] == res[1][3] == 3
    assert all(math.isfinite(l) for r in res for l in r[0])


def test_rank_without_rows_keeps_the_collective_in_step():
    cfg = TrainConfig(quiet=True, nepochs=2, n_samples=2, batch_size=1, use_batch_size=True, data_dist="local",
                      scale="none")

    def fn(comm):
        t = Trainer(cfg, comm, torch.device("cpu"))
        t.train()
        return t.flat.fingerprint(), t.flat.step_count, t.shard.rows

    res = run_threads(3, fn)                                     # shards of 1, 1, 0 rows
    assert [r[2] for r in res] == [1, 1, 0]
    assert len({r[0] for r in res}) == 1 and {r[1] for r in res} == {2}


def test_resume_of_a_shuffled_minibatch_run_is_exact(tmp_path):
    kw = dict(quiet=True, n_samples=32, batch_size=8, use_batch_size=True, data_dist="local", scale="global",
              lr=3e-4, hidden=[8], n_features=4)
    full = Trainer(TrainConfig(**kw, nepochs=6), device=torch.device("cpu"))
    full.train()
    path = str(tmp_path / "ck.pt")
    first = Trainer(TrainConfig(**kw, nepochs=3, checkpoint=path), device=torch.device("cpu"))
    first.train()
    second = Trainer(TrainConfig(**kw, nepochs=6, resume=path), device=torch.device("cpu"))
    assert second.start_epoch == 3
    second.train()
    assert second.flat.fingerprint() == full.flat.fingerprint()
    assert [h["loss"] for h in second.history] == [h["loss"] for h in full.history[3:]]
