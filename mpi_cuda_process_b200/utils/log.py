"""JSON-lines logging of run metrics (rank 0)."""
from __future__ import annotations

import json
import sys
import time
from typing import Optional


class JsonlLogger:
    def __init__(self, path: Optional[str] = None, rank: int = 0):
        self.rank = rank
        self._f = open(path, "a") if (path and rank == 0) else None

    def log(self, **kv) -> None:
        if self.rank != 0:
            return
        kv.setdefault("ts", time.time())
        line = json.dumps(kv, sort_keys=True)
        if self._f is not None:
            self._f.write(line + "\n")
            self._f.flush()
        else:
            print(line, file=sys.stderr)

    def close(self) -> None:
        if self._f is not None:
            self._f.close()
            self._f = None
