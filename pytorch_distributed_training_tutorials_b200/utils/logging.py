"""Rank-aware structured logging (JSON lines; rank 0 only unless asked otherwise).

The reference only prints (its banner lives in ``trainer.py`` unchanged); these helpers add what a
production job needs: one JSON object per line with the rank attached, and fault-injection hooks
for tests (``B200DT_FAIL_RANK`` / ``B200DT_FAIL_STEP`` make one rank raise at one step).
"""
from __future__ import annotations

import json
import os
import sys
import time

from .. import group as _group


def log_json(all_ranks: bool = False, stream=None, **fields) -> None:
    r = _group.rank()
    if r != 0 and not all_ranks:
        return
    rec = {"ts": round(time.time(), 3), "rank": r}
    rec.update(fields)
    print(json.dumps(rec), file=stream or sys.stdout, flush=True)


def maybe_inject_fault(step: int) -> None:
    """Raise on the configured (rank, step): exercises launcher teardown and error reporting."""
    fr, fs = os.environ.get("B200DT_FAIL_RANK"), os.environ.get("B200DT_FAIL_STEP")
    if fr is None or fs is None:
        return
    if int(fr) == _group.rank() and int(fs) == step:
        raise RuntimeError(f"injected fault on rank {fr} at step {fs}")
