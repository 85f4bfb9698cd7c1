"""``spawn(fn, args, nprocs)``: one worker process per rank, ``fn(rank, *args)`` in each.

Mirrors the launch used by the reference (/root/reference/ddp_gpus.py:104-105, ``mp.spawn``)
including its failure contract: children die with the parent, the first failing child is
reported with its rank and traceback, and its siblings are torn down (SIGTERM, then SIGKILL)
instead of hanging in a collective.
"""
from __future__ import annotations

import ctypes
import multiprocessing as mp
import multiprocessing.connection
import os
import pickle
import signal
import sys
import tempfile
import time
import traceback
from typing import Callable, List, Optional, Sequence


class ProcessRaisedException(RuntimeError):
    """A worker raised; ``rank`` says which one, the message carries its traceback."""

    def __init__(self, msg: str, rank: int, pid: int):
        super().__init__(msg)
        self.rank = rank
        self.error_index = rank     # name used by torch.multiprocessing
        self.pid = pid


class ProcessExitedException(RuntimeError):
    """A worker died without a Python exception (signal or non-zero exit)."""

    def __init__(self, msg: str, rank: int, pid: int, exit_code: int, signal_name: Optional[str] = None):
        super().__init__(msg)
        self.rank = rank
        self.error_index = rank
        self.pid = pid
        self.exit_code = exit_code
        self.signal_name = signal_name


def _die_with_parent() -> None:
    # PR_SET_PDEATHSIG = 1: deliver SIGINT to this process when its parent exits.
    try:
        libc = ctypes.CDLL("libc.so.6", use_errno=True)
        libc.prctl(1, int(signal.SIGINT), 0, 0, 0)
    except Exception:  # non-Linux: best effort
        pass


def _worker(fn: Callable, rank: int, args: Sequence, error_path: str, env: dict) -> None:
    _die_with_parent()
    os.environ.update(env)
    try:
        fn(rank, *args)
    except KeyboardInterrupt:
        pass        # parent is tearing us down
    except Exception:
        with open(error_path, "wb") as f:
            pickle.dump(traceback.format_exc(), f)
        sys.exit(1)


class SpawnContext:
    def __init__(self, procs: List[mp.Process], error_files: List[str]):
        self.processes = procs
        self.error_files = error_files
        self._sentinels = {p.sentinel: i for i, p in enumerate(procs)}

    def pids(self) -> List[int]:
        return [p.pid for p in self.processes]

    def _terminate_others(self, failed: int, grace_s: float) -> None:
        for i, p in enumerate(self.processes):
            if i != failed and p.is_alive():
                p.terminate()
        deadline = time.monotonic() + grace_s
        for i, p in enumerate(self.processes):
            if i == failed:
                continue
            p.join(max(0.0, deadline - time.monotonic()))
            if p.is_alive():
                p.kill()
                p.join()

    def join(self, timeout: Optional[float] = None, grace_s: float = 10.0) -> bool:
        """Wait for workers. Returns True when all finished cleanly, False on timeout."""
        if not self._sentinels:
            return True
        ready = multiprocessing.connection.wait(list(self._sentinels.keys()), timeout=timeout)
        failed = None
        for s in ready:
            idx = self._sentinels.pop(s)
            p = self.processes[idx]
            p.join()
            if p.exitcode != 0 and failed is None:
                failed = idx
        if failed is None:
            return not self._sentinels
        self._terminate_others(failed, grace_s)
        p = self.processes[failed]
        path = self.error_files[failed]
        if os.path.exists(path) and os.path.getsize(path) > 0:
            with open(path, "rb") as f:
                tb = pickle.load(f)
            raise ProcessRaisedException(
                f"\n\n-- Process {failed} terminated with the following error:\n{tb}", failed, p.pid)
        code = p.exitcode
        if code is not None and code < 0:
            try:
                name = signal.Signals(-code).name
            except ValueError:
                name = f"<unknown signal {-code}>"
            raise ProcessExitedException(f"process {failed} terminated with signal {name}", failed, p.pid,
                                         code, name)
        raise ProcessExitedException(f"process {failed} terminated with exit code {code}", failed, p.pid, code)


def start_processes(fn: Callable, args: Sequence = (), nprocs: int = 1, join: bool = True,
                    daemon: bool = False, start_method: str = "spawn", env: Optional[dict] = None):
    if nprocs < 1:
        raise ValueError("nprocs must be >= 1")
    ctx = mp.get_context(start_method)
    tmpdir = tempfile.mkdtemp(prefix="b200dt_spawn_")
    procs, files = [], []
    for i in range(nprocs):
        path = os.path.join(tmpdir, f"rank{i}.err")
        p = ctx.Process(target=_worker, args=(fn, i, tuple(args), path, dict(env or {})), daemon=daemon)
        p.start()
        procs.append(p)
        files.append(path)
    sc = SpawnContext(procs, files)
    if not join:
        return sc
    try:
        while not sc.join():
            pass
    finally:
        for f in files:
            try:
                os.unlink(f)
            except OSError:
                pass
        try:
            os.rmdir(tmpdir)
        except OSError:
            pass
    return None


def spawn(fn: Callable, args: Sequence = (), nprocs: int = 1, join: bool = True, daemon: bool = False,
          env: Optional[dict] = None):
    """Start ``nprocs`` workers running ``fn(rank, *args)``; same call shape as ``mp.spawn``."""
    return start_processes(fn, args, nprocs, join, daemon, "spawn", env)
