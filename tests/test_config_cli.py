import argparse

import pytest

from b200dp.config import TrainConfig, build_parser, config_from_args, parse_args
from b200dp.launch import discover


def test_defaults_match_reference():
    cfg = parse_args([])
    # ref.py:245-252
    assert (cfg.lr, cfg.momentum, cfg.batch_size, cfg.nepochs) == (0.001, 0.9, 4, 3)
    # ref.py:72 and ref.py:41-45
    assert (cfg.n_samples, cfg.n_features, cfg.noise, cfg.data_seed) == (16, 2, 1.0, 42)
    assert cfg.layer_sizes() == [2, 3, 1]
    assert cfg.scale == "per_shard" and cfg.grad_weighting == "uniform" and not cfg.use_batch_size


def test_lr_and_momentum_are_floats_when_given():
    cfg = parse_args(["--lr", "0.05", "--momentum", "0.5", "--batch_size", "8", "--nepochs", "7"])
    assert cfg.lr == 0.05 and cfg.momentum == 0.5 and cfg.batch_size == 8 and cfg.nepochs == 7
    assert isinstance(cfg.lr, float)


def test_reference_namespace_is_accepted():
    ns = argparse.Namespace(lr="0.01", momentum=0.9, batch_size="4", nepochs=2)   # strings, as Q1
    cfg = config_from_args(ns)
    assert cfg.lr == 0.01 and cfg.batch_size == 4 and cfg.nepochs == 2


def test_hidden_forms():
    assert parse_args(["--hidden", "4096x4"]).hidden == [4096] * 4
    assert parse_args(["--hidden", "512,256"]).hidden == [512, 256]
    assert parse_args(["--hidden", "8", "--depth", "3"]).hidden == [8, 8, 8]
    assert parse_args(["--hidden", "4096x4", "--n-features", "12288"]).layer_sizes() == \
        [12288, 4096, 4096, 4096, 4096, 1]


def test_validation_errors():
    with pytest.raises(SystemExit):
        parse_args(["--scale", "bogus"])
    with pytest.raises(ValueError):
        TrainConfig(sync="nope").validate()


def test_help_mentions_real_default():
    text = build_parser().format_help()
    assert "[0.001]" in text


@pytest.mark.parametrize("env,expect", [
    ({}, (0, 1, 0, "single")),
    ({"RANK": "3", "WORLD_SIZE": "8", "LOCAL_RANK": "3"}, (3, 8, 3, "torchrun")),
    ({"OMPI_COMM_WORLD_RANK": "1", "OMPI_COMM_WORLD_SIZE": "2", "OMPI_COMM_WORLD_LOCAL_RANK": "1"},
     (1, 2, 1, "openmpi")),
    ({"PMI_RANK": "2", "PMI_SIZE": "4"}, (2, 4, 2, "pmi")),
    ({"SLURM_PROCID": "5", "SLURM_NTASKS": "6", "SLURM_LOCALID": "1"}, (5, 6, 1, "pmi")),
])
def test_launcher_discovery(env, expect):
    info = discover(env)
    assert (info.rank, info.world_size, info.local_rank, info.launcher) == expect
