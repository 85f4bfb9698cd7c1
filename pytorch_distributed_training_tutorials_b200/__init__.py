"""B200-native data- and model-parallel training library.

Public surface (mirrors what the tutorial repo exercises through torch):

* ``init`` / ``destroy``                 process-group bootstrap (NCCL on GPUs, gloo on CPU)
* ``spawn``, ``python -m ...run``        launchers (mp.spawn / torchrun contracts)
* ``DistributedSampler`` + toy datasets  data sharding
* ``DistributedDataParallel``            multi-process DP, fused peer-memory gradient all-reduce
* ``DataParallel``                       single-process multi-GPU DP (scatter/replicate/apply/gather)
* ``parallel.model_parallel`` / ``pipeline`` / ``place``   naive, pipelined and auto-placed MP
* ``Trainer``, ``save_checkpoint`` / ``load_checkpoint``
* ``ops``                                sm_100a kernels: tcgen05 GEMM/conv, BN, pools, losses, optimizers
"""
from .group import destroy, init, is_initialized, rank, world_size, barrier, find_free_port  # noqa: F401
from .data import DistributedSampler, MyTrainDataset, RandomDataset, SyntheticImageDataset  # noqa: F401
from .launch.spawn import spawn, ProcessRaisedException, ProcessExitedException  # noqa: F401

__version__ = "0.1.0"

_LAZY = {
    "DistributedDataParallel": ("parallel.ddp", "DistributedDataParallel"),
    "DataParallel": ("parallel.data_parallel", "DataParallel"),
    "GraphedPipeline": ("parallel.graphed_pipeline", "GraphedPipeline"),
    "GraphedDataParallel": ("parallel.graphed_data_parallel", "GraphedDataParallel"),
    "PipelineStage": ("parallel.pipeline", "PipelineStage"),
    "Trainer": ("trainer", "Trainer"),
    "save_checkpoint": ("checkpoint", "save_checkpoint"),
    "load_checkpoint": ("checkpoint", "load_checkpoint"),
    "FusedSGD": ("ops.optim", "FusedSGD"),
    "FusedAdam": ("ops.optim", "FusedAdam"),
}


def __getattr__(name):
    if name in _LAZY:
        import importlib
        mod, attr = _LAZY[name]
        return getattr(importlib.import_module(f"{__name__}.{mod}"), attr)
    raise AttributeError(name)
