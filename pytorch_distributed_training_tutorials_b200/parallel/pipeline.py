"""Multi-process pipeline parallelism: one process (and GPU) per stage, micro-batches, p2p transfers.

The single-process ``model_parallel.Pipeline`` keeps both GPUs busy only while one Python thread can
issue kernels faster than a GPU retires them; on a B200 that stops being true for micro-batches of
a few dozen images.  Here every stage has its own process, so the host-side work is split the same
way the model is: stage k runs the fill-drain (GPipe) schedule over ``chunks`` micro-batches, receives
the boundary activation from stage k-1 and sends its own to stage k+1 with ``torch.distributed``
point-to-point calls (NCCL on GPUs, stream-ordered on NCCL's own stream, so transfers overlap the
micro-batch compute; gloo on CPU for the tests), and the gradients travel back the same way.

The reference only has the naive two-GPU split (/root/reference/03.model_parallel.ipynb:807-833); the
pipelined variant is required by BASELINE.json (config 4).
"""
from __future__ import annotations

from typing import Callable, List, Optional

import torch
import torch.distributed as dist
from torch import nn


class PipelineStage(nn.Module):
    """Stage ``stage`` of ``n_stages``; ranks of ``group`` are the stages in order.

    ``train_step`` runs forward and backward for one mini-batch split into ``chunks`` micro-batches:
    the first stage passes ``x``, the last stage passes ``target`` and ``loss_fn`` and gets the mean
    loss back; gradients are accumulated in the stage's parameters (call ``optimizer.step()`` after).
    """

    def __init__(self, module: nn.Module, stage: Optional[int] = None, n_stages: Optional[int] = None,
                 chunks: int = 4, group=None) -> None:
        super().__init__()
        self.module = module
        self.group = group if group is not None else dist.group.WORLD
        self.stage = dist.get_rank(self.group) if stage is None else stage
        self.n_stages = dist.get_world_size(self.group) if n_stages is None else n_stages
        self.chunks = chunks
        self.is_first = self.stage == 0
        self.is_last = self.stage == self.n_stages - 1
        self._shapes = {}            # micro-batch index -> (shape, dtype) of the incoming activation
        self._meta_sent = set()      # micro-batch indices whose shape the consumer already knows
        p = next(module.parameters(), None)
        self.device = p.device if p is not None else torch.device("cpu")

    # ------------------------------------------------------------------ p2p helpers
    def _peer(self, stage: int) -> int:
        return dist.get_global_rank(self.group, stage)

    def _send(self, t: torch.Tensor, stage: int) -> None:
        dist.send(t.contiguous(), dst=self._peer(stage), group=self.group)

    def _recv(self, shape, dtype, stage: int) -> torch.Tensor:
        t = torch.empty(shape, dtype=dtype, device=self.device)
        dist.recv(t, src=self._peer(stage), group=self.group)
        return t

    _DTYPES = [torch.float32, torch.bfloat16, torch.float16, torch.float64, torch.int64, torch.int32]

    def _send_meta(self, t: torch.Tensor, stage: int) -> None:
        meta = torch.zeros(8, dtype=torch.int64, device=self.device)
        meta[0] = t.dim()
        meta[1] = self._DTYPES.index(t.dtype)
        for i, s in enumerate(t.shape):
            meta[2 + i] = s
        self._send(meta, stage)

    def _recv_meta(self, stage: int):
        meta = self._recv((8,), torch.int64, stage).cpu().tolist()
        nd = int(meta[0])
        return tuple(int(v) for v in meta[2:2 + nd]), self._DTYPES[int(meta[1])]

    # ------------------------------------------------------------------ schedule
    def train_step(self, x: Optional[torch.Tensor] = None, target: Optional[torch.Tensor] = None,
                   loss_fn: Optional[Callable] = None, n_rows: Optional[int] = None) -> Optional[torch.Tensor]:
        """``n_rows``: mini-batch size, needed on the stages that see neither ``x`` nor ``target``."""
        m = self.chunks
        if self.is_first:
            xs = list(x.chunk(m, dim=0))
            m = len(xs)
        if self.is_last and target is not None:
            ts = list(target.chunk(m, dim=0))
            m = len(ts)
        ins: List[Optional[torch.Tensor]] = []
        outs: List[torch.Tensor] = []
        losses: List[torch.Tensor] = []
        total_rows = x.shape[0] if self.is_first else (target.shape[0] if (self.is_last and target is not None) else n_rows)
        # ---- forward, micro-batch by micro-batch (stage k works on i while k+1 works on i-1)
        for i in range(m):
            if self.is_first:
                inp = xs[i]
            else:
                if i not in self._shapes:            # first time: the producer tells us what comes
                    self._shapes[i] = self._recv_meta(self.stage - 1)
                shape, dtype = self._shapes[i]
                inp = self._recv(shape, dtype, self.stage - 1)
                if inp.is_floating_point():
                    inp.requires_grad_(True)
            out = self.module(inp)
            if self.is_last:
                if loss_fn is not None:
                    # mean over the mini-batch = sum over micro-batches of (rows_i / rows) * mean_i
                    li = loss_fn(out, ts[i]) * (ts[i].shape[0] / total_rows)
                    losses.append(li)
                    out = li
            else:
                if i not in self._meta_sent:
                    self._send_meta(out, self.stage + 1)
                    self._meta_sent.add(i)
                self._send(out.detach(), self.stage + 1)
            ins.append(inp)
            outs.append(out)
        # ---- backward in reverse order; boundary gradients flow back stage by stage
        for i in reversed(range(m)):
            if self.is_last:
                outs[i].backward()
            else:
                g = self._recv(outs[i].shape, outs[i].dtype, self.stage + 1)
                outs[i].backward(g)
            if not self.is_first:
                self._send(ins[i].grad, self.stage - 1)
            ins[i] = None
            outs[i] = None
        if self.is_last and losses:
            return torch.stack([l.detach() for l in losses]).sum()
        return None

    @torch.no_grad()
    def forward(self, x: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        """Inference through the pipeline (no micro-batching): the last stage returns the output."""
        if self.is_first:
            inp = x
        else:
            shape, dtype = self._recv_meta(self.stage - 1)
            inp = self._recv(shape, dtype, self.stage - 1)
        out = self.module(inp)
        if self.is_last:
            return out
        self._send_meta(out, self.stage + 1)
        self._send(out, self.stage + 1)
        return None
