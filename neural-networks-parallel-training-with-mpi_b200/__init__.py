"""b200dp - Blackwell-native data-parallel MLP trainer.

Same capabilities as btourn/Neural-Networks-parallel-training-with-MPI (a replicated MLP regressor,
row-sharded synthetic data, synchronous gradient averaging, SGD+momentum, the same CLI), built for
B200: hand-written sm_100a kernels (``csrc/``, C ABI in ``include/b200dp.h``) under a thin PyTorch
framework layer.  Import as ``import b200dp`` (alias module at the repo root).
"""
from .config import TrainConfig, build_parser, config_from_args, parse_args  # noqa: F401
from .data import RegressionDataset  # noqa: F401
from .models.mlp import MLP, FlatMLP, MLPSpec  # noqa: F401

__version__ = "0.1.0"


def dist_train(args, comm=None, device=None):
    """See ``engine.dist_train`` (the reference's entry point, ref.py:56)."""
    from .engine import dist_train as _dt

    return _dt(args, comm=comm, device=device)


def Trainer(*a, **kw):
    from .engine import Trainer as _T

    return _T(*a, **kw)
