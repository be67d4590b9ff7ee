"""Console lines in the reference's exact format plus an optional rank-0 JSONL metrics stream."""
from __future__ import annotations

import json
import sys
import time
from typing import Any, Optional


def epoch_banner(epoch: int) -> str:
    return "[ = = = = = Epoch {} = = = = = ]".format(epoch)          # ref.py:152


def loss_line(rank: int, loss: float) -> str:
    return f"loss in worker {rank}: {loss}"                           # ref.py:224


class JsonlLogger:
    def __init__(self, path: Optional[str]):
        self._f = open(path, "a") if path else None

    def log(self, **rec: Any) -> None:
        if self._f:
            rec.setdefault("ts", time.time())
            self._f.write(json.dumps(rec) + "\n")
            self._f.flush()

    def close(self):
        if self._f:
            self._f.close()
            self._f = None


def emit(line: str, quiet: bool = False) -> None:
    if not quiet:
        print(line)
        sys.stdout.flush()
