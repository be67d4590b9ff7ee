import pytest

from b200dp.parallel.sharding import (element_counts, grad_weights, shard_counts, shard_displs,
                                      shard_range)


@pytest.mark.parametrize("rows,world", [(16, 1), (16, 2), (16, 3), (16, 5), (16, 7), (16, 8), (16, 16),
                                        (16, 20), (0, 4), (1, 3), (1000003, 8)])
def test_counts_cover_all_rows_contiguously(rows, world):
    c = shard_counts(rows, world)
    d = shard_displs(c)
    assert sum(c) == rows and len(c) == world
    assert max(c) - min(c) <= 1
    # first `residue` ranks own the extra row (ref.py:117)
    res = rows % world
    assert all(c[p] == rows // world + 1 for p in range(res))
    assert all(c[p] == rows // world for p in range(res, world))
    for r in range(world):
        assert shard_range(rows, world, r) == (d[r], d[r] + c[r])
    assert d == [sum(c[:p]) for p in range(world)]         # ref.py:121


def test_reference_example_16_rows_3_ranks():
    assert shard_counts(16, 3) == [6, 5, 5]
    assert shard_counts(16, 5) == [4, 3, 3, 3, 3]
    cnt, dsp = element_counts(16, 3, 3)
    assert cnt == [18, 15, 15] and dsp == [0, 18, 33]


def test_counts_exceed_int8_range():
    # the reference stores w*count in int8 (ref.py:118); ours must not overflow
    cnt, _ = element_counts(100000, 3, 2)
    assert cnt == [150000, 150000]


def test_grad_weights():
    assert grad_weights(16, 4) == [0.25] * 4
    w = grad_weights(16, 3, "samples")
    assert w == [6 / 16, 5 / 16, 5 / 16] and abs(sum(w) - 1) < 1e-12
    with pytest.raises(ValueError):
        grad_weights(16, 3, "bogus")


def test_bad_world():
    with pytest.raises(ValueError):
        shard_counts(4, 0)


from hypothesis import given, settings, strategies as st


@settings(max_examples=200, deadline=None)
@given(rows=st.integers(0, 10**7), world=st.integers(1, 512))
def test_shard_properties(rows, world):
    c = shard_counts(rows, world)
    d = shard_displs(c)
    assert sum(c) == rows and len(c) == world and max(c) - min(c) <= 1
    assert c == sorted(c, reverse=True)                      # the extra rows go to the first ranks
    assert all(d[i + 1] == d[i] + c[i] for i in range(world - 1))
    ws = grad_weights(rows, world, "samples") if rows else None
    if ws:
        assert abs(sum(ws) - 1.0) < 1e-9


def test_spread_placement_takes_one_gpu_per_pcie_switch_first():
    """launch.spread_devices: a 2- or 4-rank run on an 8-GPU HGX board (GPU pairs share a PCIe switch and
    its x16 uplink) must not put two ranks on one switch while another switch is idle."""
    from b200dp.launch import spread_devices
    c = [[0] * 8 for _ in range(8)]
    for i in range(8):
        for j in range(8):
            if i != j:
                c[i][j] = 40 if i // 2 == j // 2 else (20 if i // 4 == j // 4 else 10)
    assert spread_devices(2, c, 8) == [0, 4]
    assert sorted(d // 2 for d in spread_devices(4, c, 8)) == [0, 1, 2, 3]
    assert spread_devices(8, c, 8) == list(range(8))
    assert spread_devices(3, None, 8) == [0, 1, 2]                   # no topology data: identity


def test_bench_counts_executed_flops():
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("bench", os.path.join(os.path.dirname(__file__), "..", "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    sizes, rows = [12288, 4096, 4096, 4096, 4096, 1], 8192
    full = sum(sizes[i] * sizes[i + 1] for i in range(5))
    # forward + wgrad for every layer, dgrad for all but the first: 4.12 TFLOP, not the 4.95 of 6*N*D
    assert bench.executed_flops(sizes, rows) == 2.0 * rows * (3 * full - sizes[0] * sizes[1])
    assert abs(bench.executed_flops(sizes, rows) / 1e12 - 4.124) < 0.01


def test_measured_placement_falls_back_without_gpus(monkeypatch, tmp_path):
    """launch.measured_spread on a box where the probe cannot run (no GPU here): the child fails, the
    answer file says so, and placement falls back to GPUs 0..N-1 instead of raising or hanging."""
    import tempfile
    from b200dp import launch
    monkeypatch.setattr(tempfile, "tempdir", str(tmp_path))
    monkeypatch.setenv("LOCAL_RANK", "0")
    monkeypatch.setenv("MASTER_PORT", "45678")
    import torch
    if torch.cuda.is_available() and torch.cuda.device_count() > 2:
        import pytest
        pytest.skip("this test is about boxes where the probe fails")
    assert launch.measured_spread(2, timeout_s=60.0) is None
    assert launch.spread_devices(2, None, 8) == [0, 1]
