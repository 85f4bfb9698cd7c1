"""Rank-aware structured logging (JSON lines; rank 0 only unless asked otherwise).

The reference only prints (its banner lives in ``trainer.py`` unchanged); these helpers add what a
production job needs: one JSON object per line with the rank attached.  (Fault injection for the
launcher tests lives in ``trainer.py``: ``B200DT_FAULT_INJECT="rank=R,step=S"``.)
"""
from __future__ import annotations

import json
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
