#!/usr/bin/env python
"""Entry script.  Single process: ``python train.py``.  Multi-GPU / multi-process:
``torchrun --nproc-per-node N --master-addr 127.0.0.1 train.py [flags]`` or
``mpiexec -n N python train.py [flags]`` (rank is read from the launcher's environment)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import b200dp  # noqa: E402,F401
from b200dp.cli import main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main())
