"""Multi-process DDP launched with ``spawn`` (counterpart of /root/reference/ddp_gpus.py).

    python examples/ddp_spawn.py --max_epochs 5 --batch_size 32 [--backend gloo --nprocs 2]

One worker per GPU (or ``--nprocs`` CPU workers under gloo).  Same workload as the reference:
``Linear(20, 1)``, 2048 random samples, SGD lr 1e-2, ``F.cross_entropy`` on float targets, per-epoch
banner ``[GPU: r Epoch: e, Batch size: b | Steps s]`` (16 steps on 4 ranks, 64 on one).  Unlike the
reference the rendezvous port is picked free at run time instead of the hard-coded 12345.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.utils.data import DataLoader

import pytorch_distributed_training_tutorials_b200 as b2
from pytorch_distributed_training_tutorials_b200.models.toy import ToyLinear
from pytorch_distributed_training_tutorials_b200.trainer import Trainer


def ddp_setup(rank: int, world_size: int, backend: str, port: int) -> None:
    b2.init(backend=backend, rank=rank, world_size=world_size, master_addr="127.0.0.1", master_port=port)


def main(rank: int, world_size: int, max_epochs: int, batch_size: int, backend: str, port: int):
    ddp_setup(rank, world_size, backend, port)
    train_dataset = b2.MyTrainDataset(2048)
    train_dataloader = DataLoader(train_dataset, batch_size=batch_size, pin_memory=torch.cuda.is_available(),
                                  shuffle=False, sampler=b2.DistributedSampler(train_dataset))
    model = ToyLinear()
    optimizer = torch.optim.SGD(model.parameters(), lr=1e-2)
    trainer = Trainer(model=model, gpu_id=rank, optimizer=optimizer, train_dataloader=train_dataloader)
    trainer.train(max_epochs)
    b2.destroy()


if __name__ == "__main__":
    parser = argparse.ArgumentParser(description="simple distributed training job")
    parser.add_argument("--max_epochs", type=int, default=10, help="Total epochs to train the model")
    parser.add_argument("--batch_size", type=int, default=32, help="Input batch size on each device")
    parser.add_argument("--backend", default=None, help="nccl (default on GPUs) or gloo")
    parser.add_argument("--nprocs", type=int, default=None, help="workers (default: number of GPUs)")
    args = parser.parse_args()
    backend = args.backend or ("nccl" if torch.cuda.is_available() else "gloo")
    world_size = args.nprocs or (torch.cuda.device_count() if backend == "nccl" else 2)
    b2.spawn(main, args=(world_size, args.max_epochs, args.batch_size, backend, b2.find_free_port()), nprocs=world_size)
