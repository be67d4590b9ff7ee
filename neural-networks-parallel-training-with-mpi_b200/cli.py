"""``python train.py [flags]`` / ``torchrun ... train.py`` / ``mpiexec -n N python train.py``
(README:12-14 launch contract; the ``__main__`` block of ref.py:242-255)."""
from __future__ import annotations

import sys


def main(argv=None) -> int:
    from .config import parse_args
    from .engine import dist_train

    cfg = parse_args(argv)
    trainer = dist_train(cfg)
    trainer.close()
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
