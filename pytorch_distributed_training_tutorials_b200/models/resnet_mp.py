"""ResNet-50 split across two GPUs: naive and pipelined.

``ModelParallelResNet50`` uses the reference's split point (03.model_parallel.ipynb:807-833):
``seq1 = conv1, bn1, relu, maxpool, layer1, layer2`` on the first device, ``seq2 = layer3, layer4,
avgpool`` and ``fc`` on the second; 25 557 032 parameters either way (ibid. 857, 888).  Unlike the
reference, stages are not registered twice, so ``state_dict()`` has exactly torchvision's keys.
``PipelineParallelResNet50`` runs the same split on micro-batches of ``split_size`` images.
"""
from __future__ import annotations

import torch

from ..parallel.model_parallel import hop
from .resnet import ResNet


class ModelParallelResNet50(ResNet):
    """``split_after`` names the last unit that lives on the first device: a whole segment (``"stem"``,
    ``"layer1"`` .. ``"layer3"``) or one bottleneck block (``"layer2.1"``, see ``ResNet.unit_names``).
    ``"layer2"`` is the reference's split; on B200 the high-resolution half (stem .. layer2) takes ~70 % of the
    step time with these kernels, so a pipeline wants the cut inside layer2: ``balanced()`` measures it."""

    def __init__(self, num_classes: int = 1000, dev0="cuda:0", dev1="cuda:1", split_after: str = "layer2") -> None:
        super().__init__([3, 4, 6, 3], num_classes=num_classes)
        self.dev0, self.dev1 = torch.device(dev0), torch.device(dev1)
        names = self.unit_names()
        cut = self.unit_index(split_after)
        if not 0 <= cut < len(names) - 1:
            raise ValueError(f"split_after={split_after!r} leaves nothing for the second device")
        self.split_after, self._cut, self._last = split_after, cut, len(names) - 1
        for i, name in enumerate(names):
            for mod in self.unit_modules(name):
                mod.to(self.dev0 if i <= cut else self.dev1)
        self.fc.to(self.dev1)

    @classmethod
    def balanced(cls, example_x: torch.Tensor, chunks: int = 1, reps: int = 3, **kw):
        """Build the split whose slower stage is fastest: times forward + backward of every unit on one
        micro-batch (``example_x[: batch // chunks]``) on ``dev0`` and cuts where the two halves are closest
        (``parallel.place.profile_units`` / ``balanced_cuts``).  The measured times are kept in
        ``unit_times_ms`` on the returned model."""
        from ..parallel.place import balanced_cuts, profile_units
        dev0 = torch.device(kw.get("dev0", "cuda:0"))
        probe = ResNet([3, 4, 6, 3], num_classes=kw.get("num_classes", 1000)).to(dev0)
        names = probe.unit_names()
        units = [lambda t, n=n: probe.run_unit(n, t) for n in names]
        units[-1] = lambda t: probe.fc(probe.run_unit("pool", t))
        mb = example_x[: max(1, example_x.shape[0] // chunks)].to(dev0)
        times = profile_units(units, mb, reps=reps, module=probe)
        del probe, units
        if dev0.type == "cuda":
            torch.cuda.empty_cache()
        cut = balanced_cuts(times, 2)[0]
        cut = min(cut, len(names) - 2)
        model = cls(split_after=names[cut], **kw)
        model.unit_times_ms = dict(zip(names, times))
        return model

    def stage0(self, x):
        return self.run_units(x, 0, self._cut)

    def stage1(self, x):
        return self.fc(self.run_units(x, self._cut + 1, self._last))

    def forward(self, x):
        return self.stage1(hop(self.stage0(x.to(self.dev0)), self.dev1))

    def graphed(self, loss_fn, example_x: torch.Tensor, example_y: torch.Tensor, chunks: int = 1):
        """CUDA-graph replayed training step of this split (``parallel.graphed_pipeline``): forward + loss +
        backward of both stages are captured once and a step is a handful of graph launches, so the host stops
        being the bottleneck (the notebook's 120 x 128^2 batches are launch-bound in eager mode on a B200).
        ``chunks=1`` is the naive split; returns an object with ``train_step(x, y)`` / ``zero_grad()``."""
        from ..parallel.graphed_pipeline import GraphedPipeline
        return GraphedPipeline([self.stage0, self.stage1], [self.dev0, self.dev1], chunks, loss_fn,
                               example_x, example_y, parameters=list(self.parameters()))


class PipelineParallelResNet50(ModelParallelResNet50):
    def __init__(self, split_size: int = 20, **kw) -> None:
        super().__init__(**kw)
        self.split_size = split_size

    def forward(self, x):
        micro = list(x.to(self.dev0).split(self.split_size, dim=0))
        outs = []
        prev = None
        # fill-drain over two stages: stage 1 consumes micro-batch i-1 while stage 0 produces i
        for mb in micro:
            cur = hop(self.stage0(mb), self.dev1)
            if prev is not None:
                outs.append(self.stage1(prev))
            prev = cur
        outs.append(self.stage1(prev))
        return torch.cat(outs, dim=0)

    def graphed(self, loss_fn, example_x: torch.Tensor, example_y: torch.Tensor, chunks: int = 0):
        """Graph-replayed micro-batch pipeline; the micro-batch count defaults to ``batch // split_size``."""
        return super().graphed(loss_fn, example_x, example_y,
                               chunks or max(1, example_x.shape[0] // self.split_size))
