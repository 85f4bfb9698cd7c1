"""Rank-0 checkpointing that round-trips with plain ``torch.save`` / ``load_state_dict``.

The reference has no checkpointing; the target spec asks for rank-0 snapshots that
stay compatible with torch ``state_dict``.  Files written here are ordinary ``torch.save``
dictionaries ``{"model": state_dict, "optimizer": state_dict, "epoch": int, "extra": {...}}`` whose
``model`` entry has no wrapper prefix, so it loads straight into the bare model or into
torchvision's ResNet-50; ``load_checkpoint`` also accepts a bare ``state_dict`` file written by
anyone else, with or without the ``module.`` prefix that DP / DDP wrappers add.
"""
from __future__ import annotations

import os
import tempfile
from typing import Any, Dict, Optional

import torch

from . import group as _group

_WRAPPER_PREFIX = "module."


def unwrap(model: torch.nn.Module) -> torch.nn.Module:
    """The user module inside DataParallel / DistributedDataParallel wrappers (ours or torch's)."""
    while hasattr(model, "module") and isinstance(getattr(model, "module"), torch.nn.Module) \
            and type(model).__name__ in ("DistributedDataParallel", "DataParallel"):
        model = model.module
    return model


def strip_prefix(state: Dict[str, Any], prefix: str = _WRAPPER_PREFIX) -> Dict[str, Any]:
    if state and all(k.startswith(prefix) for k in state):
        return {k[len(prefix):]: v for k, v in state.items()}
    return state


def _to_cpu(obj):
    if isinstance(obj, torch.Tensor):
        return obj.detach().to("cpu")
    if isinstance(obj, dict):
        return {k: _to_cpu(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_cpu(v) for v in obj)
    return obj


def model_state(model: torch.nn.Module, dedup_aliases: bool = True) -> Dict[str, torch.Tensor]:
    """CPU ``state_dict`` of the unwrapped model.  Modules that register a sub-module twice (the
    reference's split ResNet keeps ``layer1`` *and* ``seq1.4``, 03.model_parallel.ipynb:812-825)
    are de-aliased: of several keys sharing one tensor only the shortest path is kept, which is
    the key a plain torchvision model expects.  Stage tensors on different GPUs are gathered."""
    sd = unwrap(model).state_dict()
    if dedup_aliases:
        seen: Dict[int, str] = {}
        for k, v in sd.items():
            if not isinstance(v, torch.Tensor):
                continue
            key = (v.data_ptr(), tuple(v.shape), str(v.device)) if v.numel() else (id(v),)
            prev = seen.get(key)
            if prev is None or (k.count(".") < prev.count(".")) or (k.count(".") == prev.count(".") and len(k) < len(prev)):
                seen[key] = k
        keep = set(seen.values())
        sd = {k: v for k, v in sd.items() if (not isinstance(v, torch.Tensor)) or k in keep}
    return _to_cpu(sd)


def save_checkpoint(model: torch.nn.Module, optimizer: Optional[torch.optim.Optimizer] = None, epoch: int = 0,
                    path: str = "checkpoint.pt", extra: Optional[dict] = None, barrier: bool = True) -> Optional[str]:
    """Rank 0 writes ``path`` atomically (temp file + rename); every rank leaves together."""
    written = None
    if _group.rank() == 0:
        payload = {"model": model_state(model), "epoch": int(epoch), "extra": dict(extra or {})}
        if optimizer is not None:
            payload["optimizer"] = _to_cpu(optimizer.state_dict())
        d = os.path.dirname(os.path.abspath(path))
        os.makedirs(d, exist_ok=True)
        fd, tmp = tempfile.mkstemp(prefix="tmp_ckpt_", suffix=".pt", dir=d)
        os.close(fd)
        try:
            torch.save(payload, tmp)
            os.replace(tmp, path)
        finally:
            if os.path.exists(tmp):
                os.unlink(tmp)
        written = path
    if barrier and _group.world_size() > 1:
        _group.barrier()
    return written


def load_checkpoint(path: str, model: torch.nn.Module, optimizer: Optional[torch.optim.Optimizer] = None,
                    map_location="cpu", strict: bool = True) -> dict:
    """Load on every rank (each reads the file). Returns ``{"epoch", "extra", "missing", "unexpected"}``."""
    blob = torch.load(path, map_location=map_location, weights_only=False)
    if isinstance(blob, dict) and "model" in blob and isinstance(blob["model"], dict):
        state = blob["model"]
        epoch = int(blob.get("epoch", 0))
        extra = blob.get("extra", {})
        opt_state = blob.get("optimizer")
    else:                                   # a bare state_dict written with plain torch.save
        state, epoch, extra, opt_state = blob, 0, {}, None
    state = strip_prefix(state)
    target = unwrap(model)
    own = target.state_dict()
    # aliased registrations: fill alias keys from their canonical key so strict loading still works
    if strict:
        by_ptr: Dict[tuple, list] = {}
        for k, v in own.items():
            if isinstance(v, torch.Tensor) and v.numel():
                by_ptr.setdefault((v.data_ptr(), tuple(v.shape), str(v.device)), []).append(k)
        for keys in by_ptr.values():
            src = next((k for k in keys if k in state), None)
            if src is not None:
                for k in keys:
                    state.setdefault(k, state[src])
    res = target.load_state_dict(state, strict=strict)
    if optimizer is not None and opt_state is not None:
        optimizer.load_state_dict(opt_state)
    return {"epoch": epoch, "extra": extra, "missing": list(res.missing_keys), "unexpected": list(res.unexpected_keys)}
