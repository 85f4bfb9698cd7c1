"""``python -m pytorch_distributed_training_tutorials_b200.run [--nproc-per-node N] script.py ...``

A single-node worker launcher honouring the torchrun environment contract the reference relies
on (/root/reference/ddp_gpus_torchrun.py:13-14,24; launched as ``torchrun --nproc-per-node 4`` at
02.ddp_toy_example.ipynb:318): every worker gets ``LOCAL_RANK``, ``RANK``, ``GROUP_RANK``,
``ROLE_RANK``, ``LOCAL_WORLD_SIZE``, ``WORLD_SIZE``, ``MASTER_ADDR``, ``MASTER_PORT`` and the
``TORCHELASTIC_*`` restart counters.  As with torchrun, ``--nproc-per-node`` defaults to 1 (so the
tutorial's "forgot the flag" run trains on one GPU for 64 steps, 02.ddp_toy_example.ipynb:268),
``OMP_NUM_THREADS`` is pinned to 1 for multi-worker jobs, and a failing worker tears the group
down; ``--max-restarts`` re-launches the whole group.  Scripts written for torchrun run unchanged.
"""
from __future__ import annotations

import argparse
import os
import signal
import subprocess
import sys
import time
import uuid
from typing import List, Optional

from .group import find_free_port


def parse_args(argv: Optional[List[str]] = None) -> argparse.Namespace:
    p = argparse.ArgumentParser(prog="b200dt.run", description="single-node distributed launcher")
    p.add_argument("--nnodes", type=str, default="1", help="number of nodes (static)")
    p.add_argument("--nproc-per-node", "--nproc_per_node", dest="nproc_per_node", type=str, default="1",
                   help="workers per node: an int, 'gpu' or 'auto'")
    p.add_argument("--node-rank", "--node_rank", dest="node_rank", type=int, default=0)
    p.add_argument("--master-addr", "--master_addr", dest="master_addr", type=str, default="127.0.0.1")
    p.add_argument("--master-port", "--master_port", dest="master_port", type=int, default=29500)
    p.add_argument("--standalone", action="store_true", help="single node, pick a free port")
    p.add_argument("--local-addr", "--local_addr", dest="local_addr", type=str, default=None)
    p.add_argument("--max-restarts", "--max_restarts", dest="max_restarts", type=int, default=0)
    p.add_argument("--monitor-interval", "--monitor_interval", dest="monitor_interval", type=float, default=0.1)
    p.add_argument("--run-id", "--rdzv-id", "--rdzv_id", dest="run_id", type=str, default=None)
    p.add_argument("--rdzv-endpoint", "--rdzv_endpoint", dest="rdzv_endpoint", type=str, default=None,
                   help="host[:port] of rank 0, torchrun spelling of --master-addr / --master-port")
    p.add_argument("--rdzv-backend", "--rdzv_backend", dest="rdzv_backend", type=str, default="static",
                   help="accepted for torchrun command lines; the rendezvous is always the static TCP store")
    p.add_argument("-m", "--module", action="store_true", help="treat the script as a python module name")
    p.add_argument("--no-python", "--no_python", dest="no_python", action="store_true")
    p.add_argument("training_script", type=str)
    p.add_argument("training_script_args", nargs=argparse.REMAINDER)
    args = p.parse_args(argv)
    if args.rdzv_endpoint:
        host, _, port = args.rdzv_endpoint.partition(":")
        if host:
            args.master_addr = host
        if port:
            args.master_port = int(port)
    return args


def resolve_nproc(spec: str) -> int:
    if spec in ("gpu", "auto"):
        try:
            import torch
            n = torch.cuda.device_count()
        except Exception:
            n = 0
        if spec == "gpu" and n == 0:
            raise SystemExit("--nproc-per-node gpu: no CUDA device found")
        return n if n > 0 else (os.cpu_count() or 1)
    n = int(spec)
    if n < 1:
        raise SystemExit("--nproc-per-node must be >= 1")
    return n


def worker_env(args, nproc: int, local_rank: int, restart: int, port: int, run_id: str) -> dict:
    nnodes = int(str(args.nnodes).split(":")[0])
    env = dict(os.environ)
    rank = args.node_rank * nproc + local_rank
    env.update({
        "LOCAL_RANK": str(local_rank), "RANK": str(rank), "GROUP_RANK": str(args.node_rank),
        "ROLE_RANK": str(rank), "ROLE_NAME": "default", "LOCAL_WORLD_SIZE": str(nproc),
        "WORLD_SIZE": str(nproc * nnodes), "GROUP_WORLD_SIZE": str(nnodes), "ROLE_WORLD_SIZE": str(nproc * nnodes),
        "MASTER_ADDR": args.local_addr or args.master_addr, "MASTER_PORT": str(port),
        "TORCHELASTIC_RESTART_COUNT": str(restart), "TORCHELASTIC_MAX_RESTARTS": str(args.max_restarts),
        "TORCHELASTIC_RUN_ID": run_id, "TORCHELASTIC_USE_AGENT_STORE": "False",
        "TORCH_NCCL_ASYNC_ERROR_HANDLING": env.get("TORCH_NCCL_ASYNC_ERROR_HANDLING", "1"),
    })
    return env


def build_cmd(args) -> List[str]:
    cmd: List[str] = []
    if not args.no_python:
        cmd += [sys.executable, "-u"]
        if args.module:
            cmd.append("-m")
    cmd.append(args.training_script)
    cmd += list(args.training_script_args)
    return cmd


def _kill_group(procs: List[subprocess.Popen], grace_s: float = 10.0) -> None:
    for p in procs:
        if p.poll() is None:
            try:
                p.send_signal(signal.SIGTERM)
            except ProcessLookupError:
                pass
    deadline = time.monotonic() + grace_s
    for p in procs:
        try:
            p.wait(timeout=max(0.0, deadline - time.monotonic()))
        except subprocess.TimeoutExpired:
            p.kill()
            p.wait()


def run(args) -> int:
    nproc = resolve_nproc(args.nproc_per_node)
    if nproc > 1 and "OMP_NUM_THREADS" not in os.environ:
        print("*****************************************\n"
              "Setting OMP_NUM_THREADS environment variable for each process to be 1 in default, to avoid your "
              "system being overloaded, please further tune the variable for optimal performance in your "
              "application as needed. \n*****************************************", file=sys.stderr)
        os.environ["OMP_NUM_THREADS"] = "1"
    port = find_free_port() if args.standalone else args.master_port
    run_id = args.run_id or uuid.uuid4().hex[:8]
    cmd = build_cmd(args)
    restart = 0
    while True:
        procs = [subprocess.Popen(cmd, env=worker_env(args, nproc, lr, restart, port, run_id)) for lr in range(nproc)]
        failed = None
        try:
            while True:
                alive = False
                for lr, p in enumerate(procs):
                    rc = p.poll()
                    if rc is None:
                        alive = True
                    elif rc != 0 and failed is None:
                        failed = (lr, rc)
                if failed is not None or not alive:
                    break
                time.sleep(args.monitor_interval)
        except KeyboardInterrupt:
            _kill_group(procs)
            return 130
        if failed is None:
            return 0
        _kill_group(procs)
        lr, rc = failed
        print(f"[b200dt.run] worker local_rank={lr} (pid {procs[lr].pid}) failed with exit code {rc}", file=sys.stderr)
        if restart >= args.max_restarts:
            return rc if rc > 0 else 1
        restart += 1
        if args.standalone:
            port = find_free_port()
        print(f"[b200dt.run] restarting worker group ({restart}/{args.max_restarts})", file=sys.stderr)


def main(argv: Optional[List[str]] = None) -> None:
    sys.exit(run(parse_args(argv)))


if __name__ == "__main__":
    main()
