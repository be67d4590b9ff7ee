import numpy as np
import pytest
import torch
from sklearn.preprocessing import StandardScaler

from b200dp import MLP, RegressionDataset
from b200dp.checkpoint import load_checkpoint, save_checkpoint
from b200dp.config import TrainConfig
from b200dp.data import load_shard, make_randn_rows, make_regression_xy, standard_scale
from b200dp.engine import Trainer
from b200dp.models.mlp import FlatMLP, MLPSpec
from b200dp.parallel.comm import SingleComm, run_threads


def test_dataset_anchor_rows():                           # SURVEY §4.3 data sanity anchor
    xy = make_regression_xy()
    assert xy.shape == (16, 3) and xy.dtype == np.float64
    np.testing.assert_allclose(xy[0], [1.5230, 0.6477, 82.1004], atol=5e-4)
    np.testing.assert_allclose(xy[1], [-0.2258, 1.4656, 5.2281], atol=5e-4)


def test_standard_scale_equals_sklearn():
    x = np.random.RandomState(0).randn(11, 4) * 3 + 1
    x[:, 2] = 7.0                                         # zero-variance column
    np.testing.assert_allclose(standard_scale(x), StandardScaler().fit_transform(x), atol=1e-12)


def test_regression_dataset_interface():                  # ref.py:12-30
    xy = make_regression_xy()
    ds = RegressionDataset(xy[:, :2], xy[:, 2])
    assert len(ds) == 16 and ds.X.dtype == torch.float64
    x0, y0 = ds[0]
    assert x0.shape == (2,) and y0.shape == ()
    np.testing.assert_allclose(ds.X.mean(0).numpy(), 0, atol=1e-12)
    raw = RegressionDataset(xy[:, :2], xy[:, 2], scale_data=False)
    np.testing.assert_array_equal(raw.X.numpy(), xy[:, :2])
    # tensors are accepted (the reference silently breaks on them)
    assert len(RegressionDataset(torch.from_numpy(xy[:, :2]), torch.from_numpy(xy[:, 2]))) == 16


@pytest.mark.parametrize("world", [2, 3, 5])
def test_scatter_and_local_shards_agree(world):
    def fn(comm):
        a = load_shard(TrainConfig(data_dist="scatter"), comm)
        b = load_shard(TrainConfig(data_dist="local"), comm)
        torch.testing.assert_close(a.x, b.x)
        torch.testing.assert_close(a.y, b.y)
        return a.rows, a.start

    res = run_threads(world, fn)
    assert sum(r[0] for r in res) == 16
    assert [r[1] for r in res] == list(np.cumsum([0] + [r[0] for r in res])[:-1])


def test_randn_rows_are_world_size_independent():
    full = make_randn_rows(0, 10000, 7, 0.1, seed=3)
    part = make_randn_rows(4090, 8200, 7, 0.1, seed=3)
    np.testing.assert_array_equal(full[4090:8200], part)


def test_flat_layout_alignment_and_keys():
    spec = MLPSpec([2, 3, 1])
    assert [s.key for s in spec.segments] == ["layers.0.weight", "layers.0.bias", "layers.2.weight",
                                              "layers.2.bias"]
    assert spec.num_params == 13
    assert all(s.offset % 128 == 0 for s in spec.segments)
    big = MLPSpec([12288, 4096, 4096, 4096, 4096, 1])
    assert abs(big.num_params - 100.7e6) < 0.1e6


def test_seed0_init_matches_reference_values():
    flat = FlatMLP(MLPSpec([2, 3, 1]), "cpu")
    flat.init_like_reference(0)
    sd = flat.state_dict()
    torch.testing.assert_close(sd["layers.0.bias"], torch.tensor([-0.0140100336, 0.5606575012, -0.0627515167]))
    torch.testing.assert_close(sd["layers.2.bias"], torch.tensor([-0.5515707135]))
    torch.manual_seed(0)
    ref = MLP()
    for k, v in ref.state_dict().items():
        assert torch.equal(sd[k], v)


def test_checkpoint_roundtrip_and_loads_into_plain_module(tmp_path):
    path = str(tmp_path / "ck.pt")
    t = Trainer(TrainConfig(quiet=True, nepochs=2, checkpoint=path))
    t.train()
    payload = torch.load(path, weights_only=False)
    assert list(payload["model"].keys()) == ["layers.0.weight", "layers.0.bias", "layers.2.weight",
                                             "layers.2.bias"]
    assert all(v.dtype == torch.float32 for v in payload["model"].values())
    assert payload["epoch"] == 2
    # loads into a stock module shaped like the reference's MLP (ref.py:41-45)
    m = MLP()
    m.load_state_dict(payload["model"])
    x = torch.randn(4, 2)
    torch.testing.assert_close(m(x), t.predict(x))


def test_resume_is_exact(tmp_path):
    path = str(tmp_path / "ck.pt")
    full = Trainer(TrainConfig(quiet=True, nepochs=6))
    full.train()
    a = Trainer(TrainConfig(quiet=True, nepochs=3, checkpoint=path))
    a.train()
    b = Trainer(TrainConfig(quiet=True, nepochs=6, resume=path))
    assert b.start_epoch == 3
    b.train()
    assert b.flat.fingerprint() == full.flat.fingerprint()
    assert [h["loss"] for h in b.history] == [h["loss"] for h in full.history[3:]]


def test_checkpoint_size_mismatch(tmp_path):
    path = str(tmp_path / "ck.pt")
    flat = FlatMLP(MLPSpec([2, 3, 1]), "cpu")
    save_checkpoint(path, flat, 0, {})
    with pytest.raises(ValueError):
        load_checkpoint(path, FlatMLP(MLPSpec([2, 4, 1]), "cpu"))


def test_validation_and_jsonl(tmp_path):
    log = str(tmp_path / "m.jsonl")
    t = Trainer(TrainConfig(quiet=True, nepochs=2, n_samples=32, val_frac=0.25, log_jsonl=log,
                            data_dist="local"), SingleComm())
    h = t.train()
    t.close()
    assert "val_loss" in h[0] and h[0]["val_loss"] > 0
    import json
    recs = [json.loads(l) for l in open(log)]
    assert len(recs) == 2 and "samples_per_s" in recs[0] and "global_loss" in recs[0]


def test_padded_spec_keeps_logical_state_dict_and_zero_padding():
    """Widths that are not multiples of 8 are padded inside the flat buffers (tensor-core path); the
    padding is invisible in state_dict and mathematically inert."""
    from b200dp.ops import reference as ref_ops
    spec = MLPSpec([5, 10, 3, 1], pad_to=8)
    assert spec.sizes == (8, 16, 8, 1) and spec.logical_sizes == (5, 10, 3, 1)
    assert spec.num_params == 5 * 10 + 10 + 10 * 3 + 3 + 3 + 1
    flat = FlatMLP(spec, "cpu")
    flat.init_like_reference(0)
    sd = flat.state_dict()
    assert [tuple(v.shape) for v in sd.values()] == [(10, 5), (10,), (3, 10), (3,), (1, 3), (1,)]
    w0 = flat.weight(0)
    assert w0.shape == (16, 8) and float(w0[10:].abs().sum()) == 0 and float(w0[:, 5:].abs().sum()) == 0
    torch.manual_seed(0)
    ref = MLP([5, 10, 3, 1])
    x = torch.randn(7, 5)
    xp = torch.zeros(7, 8)
    xp[:, :5] = x
    torch.testing.assert_close(ref_ops.mlp_forward(flat, xp), ref(x))
    # one training step leaves the padding at zero and matches the unpadded model
    y = torch.randn(7, 1)
    ref_ops.mlp_loss_and_grads(flat, xp, y)
    g0 = flat.weight(0, flat.grads)
    assert float(g0[10:].abs().sum()) == 0 and float(g0[:, 5:].abs().sum()) == 0
    torch.nn.functional.mse_loss(ref(x), y).backward()
    torch.testing.assert_close(g0[:10, :5], ref.layers[0].weight.grad)
    flat2 = FlatMLP(MLPSpec([5, 10, 3, 1], pad_to=8), "cpu")
    flat2.load_state_dict(sd)
    assert flat2.fingerprint() == flat.fingerprint()


def test_two_rank_resume_is_exact_and_only_rank0_writes(tmp_path):
    path = str(tmp_path / "ck2.pt")

    def full(comm):
        t = Trainer(TrainConfig(quiet=True, nepochs=6), comm)
        t.train()
        return t.flat.fingerprint(), [h["loss"] for h in t.history]

    def first_half(comm):
        t = Trainer(TrainConfig(quiet=True, nepochs=3, checkpoint=path), comm)
        t.train()
        return t.flat.fingerprint()

    def second_half(comm):
        t = Trainer(TrainConfig(quiet=True, nepochs=6, resume=path), comm)
        assert t.start_epoch == 3
        t.train()
        return t.flat.fingerprint(), [h["loss"] for h in t.history]

    want = run_threads(2, full)
    run_threads(2, first_half)
    got = run_threads(2, second_half)
    for r in range(2):
        assert got[r][0] == want[r][0]
        assert got[r][1] == want[r][1][3:]


def test_global_loss_is_the_sample_weighted_mean():
    def fn(comm):
        t = Trainer(TrainConfig(quiet=True, nepochs=1, global_loss=True, n_samples=17, data_dist="local"), comm)
        h = t.train()
        return h[0]["loss"], t.shard.rows, h[0]["global_loss"]

    res = run_threads(3, fn)
    want = sum(l * n for l, n, _ in res) / sum(n for _, n, _ in res)
    assert all(abs(g - want) < 1e-9 * abs(want) for _, _, g in res)


def test_pinned_buffer_without_a_gpu_is_plain_host_memory():
    import torch
    from b200dp.data import PinnedBuffer, pinned_like
    if torch.cuda.is_available():
        import pytest
        pytest.skip("CPU fallback only")
    b = PinnedBuffer((4, 3), torch.float32)
    assert b.tensor.shape == (4, 3) and b.tensor.dtype == torch.float32
    c = pinned_like(torch.arange(6.0).reshape(2, 3))
    assert c.tensor.tolist() == [[0.0, 1.0, 2.0], [3.0, 4.0, 5.0]]
    b.close(); c.close()


def test_clock_sampler_is_harmless_without_a_gpu():
    import torch
    from b200dp.utils.clocks import ClockSampler
    if torch.cuda.is_available():
        import pytest
        pytest.skip("CPU fallback only")
    s = ClockSampler(0, period_ms=5).start()
    s.wait_first_sample(0.05)
    s.mark()
    out = s.stop()
    assert out["samples"] == 0 and out["reasons"] == [] and out["sm_mhz"] is None
