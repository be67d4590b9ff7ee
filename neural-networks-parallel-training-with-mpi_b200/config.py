"""Configuration and command line.

The four flags of the reference (``--lr --momentum --batch_size --nepochs``, ref.py:244-252) keep
their names and defaults.  Differences, all deliberate:

* Q1: ``--lr`` / ``--momentum`` get ``type=float`` (the reference parses them as ``str`` and crashes
  when they are given on the command line).
* Q2: help text of ``--lr`` shows the real default.
* Q3: ``--batch_size`` is accepted; like the reference (ref.py:146) the batch is the whole local
  shard unless ``--use-batch-size`` is given.
* Q12: the hard-coded sizes of ref.py:72 and ref.py:41-45 become flags whose defaults are the
  reference's constants.
"""
from __future__ import annotations

import argparse
import dataclasses
from dataclasses import dataclass, field
from typing import List, Optional

SYNC_MODES = ("auto", "gather_root", "allreduce", "fused_oneshot", "fused_twoshot", "fused_nvls")
SCALE_MODES = ("per_shard", "global", "none")
GRAD_WEIGHTING = ("uniform", "samples")
DTYPES = ("fp32", "bf16")
BACKENDS = ("auto", "torch", "cuda")


@dataclass
class TrainConfig:
    # --- the reference's four options (ref.py:245-252) ---
    lr: float = 0.001
    momentum: float = 0.9
    batch_size: int = 4
    nepochs: int = 3
    # --- dataset (ref.py:72 constants as defaults) ---
    n_samples: int = 16
    n_features: int = 2
    noise: float = 1.0
    data_seed: int = 42
    data_source: str = "sklearn"          # sklearn (reference) | randn (large synthetic shapes)
    data_dist: str = "scatter"            # scatter (root generates + Scatterv, ref.py:108,138) | local
    scale: str = "per_shard"              # Q5
    use_batch_size: bool = False          # Q3
    shuffle: bool = True                  # ref.py:146; only matters when batch < shard (Q9)
    # --- model (ref.py:41-45 as default) ---
    hidden: List[int] = field(default_factory=lambda: [3])
    n_outputs: int = 1
    seed: int = 0                         # ref.py:69 seeds root with its rank == 0
    init: str = "torch_cpu"               # torch_cpu (bit-parity with nn.Linear init) | device
    # --- numerics / execution ---
    dtype: str = "fp32"
    backend: str = "auto"
    sync: str = "auto"
    grad_weighting: str = "uniform"       # Q6
    bucket_mb: float = 0.0                # 0 = one bucket per layer (large layers in <= 3 row chunks)
    calibrate_sync: bool = False          # sync=auto: time two-shot vs NVLS at start-up and keep the faster
    overlap: bool = True
    cuda_graph: bool = False
    # --- aux subsystems ---
    checkpoint: Optional[str] = None
    checkpoint_every: int = 0
    resume: Optional[str] = None
    val_frac: float = 0.0
    log_jsonl: Optional[str] = None
    check_replicas_every: int = 0
    quiet: bool = False
    global_loss: bool = False             # extra rank-0 line with the sample-weighted global loss
    watchdog_ms: int = 10000

    def layer_sizes(self) -> List[int]:
        return [self.n_features, *self.hidden, self.n_outputs]

    def validate(self) -> "TrainConfig":
        if self.scale not in SCALE_MODES:
            raise ValueError(f"scale must be one of {SCALE_MODES}")
        if self.sync not in SYNC_MODES:
            raise ValueError(f"sync must be one of {SYNC_MODES}")
        if self.grad_weighting not in GRAD_WEIGHTING:
            raise ValueError(f"grad_weighting must be one of {GRAD_WEIGHTING}")
        if self.dtype not in DTYPES:
            raise ValueError(f"dtype must be one of {DTYPES}")
        if self.backend not in BACKENDS:
            raise ValueError(f"backend must be one of {BACKENDS}")
        if self.n_samples < 1 or self.n_features < 1 or self.n_outputs < 1:
            raise ValueError("n_samples, n_features, n_outputs must be positive")
        if any(h < 1 for h in self.hidden):
            raise ValueError("hidden sizes must be positive")
        if self.nepochs < 0:
            raise ValueError("nepochs must be >= 0")
        return self

    def asdict(self):
        return dataclasses.asdict(self)


def _hidden_arg(text: str) -> List[int]:
    """'3' -> [3];  '4096x4' or '4096*4' -> four layers of 4096;  '512,256' -> [512, 256]."""
    text = text.strip()
    for sep in ("x", "*"):
        if sep in text:
            w, d = text.split(sep)
            return [int(w)] * int(d)
    return [int(t) for t in text.split(",") if t]


def build_parser() -> argparse.ArgumentParser:
    d = TrainConfig()
    p = argparse.ArgumentParser(description="Train network across multiple distributed processes.")
    # reference flags, same dest names (ref.py:245-252)
    p.add_argument("--lr", dest="lr", default=d.lr, type=float,
                   help="Learning rate for SGD optimizer. [0.001]")
    p.add_argument("--momentum", dest="momentum", default=d.momentum, type=float,
                   help="Momentum for SGD optimizer [0.9].")
    p.add_argument("--batch_size", "--batch-size", dest="batch_size", default=d.batch_size, type=int,
                   help="Batch size to use for each process (honoured only with --use-batch-size; "
                        "the reference always trains on the full local shard).")
    p.add_argument("--nepochs", dest="nepochs", default=d.nepochs, type=int,
                   help="Number of epochs (times to loop through the dataset).")
    # dataset
    p.add_argument("--n-samples", type=int, default=d.n_samples)
    p.add_argument("--n-features", type=int, default=d.n_features)
    p.add_argument("--noise", type=float, default=d.noise)
    p.add_argument("--data-seed", type=int, default=d.data_seed)
    p.add_argument("--data-source", choices=("sklearn", "randn"), default=d.data_source)
    p.add_argument("--data-dist", choices=("scatter", "local"), default=d.data_dist)
    p.add_argument("--scale", choices=SCALE_MODES, default=d.scale)
    p.add_argument("--use-batch-size", action="store_true")
    p.add_argument("--no-shuffle", dest="shuffle", action="store_false")
    # model
    p.add_argument("--hidden", type=_hidden_arg, default=d.hidden,
                   help="hidden widths: '3', '512,256' or '4096x4'")
    p.add_argument("--depth", type=int, default=None,
                   help="repeat the (single) hidden width this many times")
    p.add_argument("--n-outputs", type=int, default=d.n_outputs)
    p.add_argument("--seed", type=int, default=d.seed)
    p.add_argument("--init", choices=("torch_cpu", "device"), default=d.init)
    # execution
    p.add_argument("--dtype", choices=DTYPES, default=d.dtype)
    p.add_argument("--backend", choices=BACKENDS, default=d.backend)
    p.add_argument("--sync", choices=SYNC_MODES, default=d.sync)
    p.add_argument("--grad-weighting", choices=GRAD_WEIGHTING, default=d.grad_weighting)
    p.add_argument("--bucket-mb", type=float, default=d.bucket_mb)
    p.add_argument("--calibrate-sync", action="store_true",
                   help="with --sync auto: time the two-shot and NVLS kernels at start-up, keep the faster")
    p.add_argument("--no-overlap", dest="overlap", action="store_false")
    p.add_argument("--cuda-graph", action="store_true")
    # aux
    p.add_argument("--checkpoint", default=None)
    p.add_argument("--checkpoint-every", type=int, default=0)
    p.add_argument("--resume", default=None)
    p.add_argument("--val-frac", type=float, default=0.0)
    p.add_argument("--log-jsonl", default=None)
    p.add_argument("--check-replicas-every", type=int, default=0)
    p.add_argument("--global-loss", action="store_true")
    p.add_argument("--quiet", action="store_true")
    p.add_argument("--watchdog-ms", type=int, default=d.watchdog_ms)
    return p


def config_from_args(args: argparse.Namespace) -> TrainConfig:
    """Accepts the reference's own 4-field Namespace as well (``dist_train(args)`` contract)."""
    fields = {f.name for f in dataclasses.fields(TrainConfig)}
    kw = {k: v for k, v in vars(args).items() if k in fields and v is not None}
    # the reference hands over strings when flags are typed on its CLI (Q1); be forgiving
    for k, typ in (("lr", float), ("momentum", float), ("batch_size", int), ("nepochs", int)):
        if k in kw:
            kw[k] = typ(kw[k])
    cfg = TrainConfig(**kw)
    depth = getattr(args, "depth", None)
    if depth is not None:
        if len(cfg.hidden) != 1:
            raise ValueError("--depth needs a single --hidden width")
        cfg.hidden = cfg.hidden * depth
    return cfg.validate()


def parse_args(argv=None) -> TrainConfig:
    return config_from_args(build_parser().parse_args(argv))
