"""Multi-process DDP under a torchrun-style launcher (counterpart of /root/reference/ddp_gpus_torchrun.py).

    python -m pytorch_distributed_training_tutorials_b200.run --nproc-per-node 4 examples/ddp_run.py --max_epochs 5
    torchrun --nproc-per-node 4 examples/ddp_run.py --max_epochs 5          # works unchanged

Rank, world size and rendezvous come from the environment (RANK / WORLD_SIZE / LOCAL_RANK /
MASTER_ADDR / MASTER_PORT).  Launched without ``--nproc-per-node`` it runs ONE worker and 64 steps
per epoch: the lesson of the reference's notebook (02.ddp_toy_example.ipynb:268-272).
``--model mlp`` swaps in the tcgen05 MLP, ``--save ckpt.pt`` writes a rank-0 checkpoint per epoch,
``--resume ckpt.pt`` loads one (on every rank, before wrapping) and continues with the next epoch.
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


def main(max_epochs: int, batch_size: int, backend, model_name: str, save: str, resume: str = ""):
    st = b2.init(backend=backend)
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if model_name == "mlp":
        from pytorch_distributed_training_tutorials_b200.models.toy import MLP
        from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
        train_dataset = b2.MyTrainDataset(2048, in_features=1024, out_features=8)
        model = MLP(1024, 2048, 3, 8)
        loss_fn = Fn.mse_loss if st.device.type == "cuda" else torch.nn.functional.mse_loss
        optimizer = b2.FusedSGD(model.parameters(), lr=1e-3)
    else:
        train_dataset = b2.MyTrainDataset(2048)
        model = ToyLinear()
        loss_fn = None                                   # F.cross_entropy, like the reference
        optimizer = torch.optim.SGD(model.parameters(), lr=1e-2)
    train_dataloader = DataLoader(train_dataset, batch_size=batch_size, pin_memory=st.device.type == "cuda",
                                  shuffle=False, sampler=b2.DistributedSampler(train_dataset))
    start_epoch = 0
    if resume:
        info = b2.load_checkpoint(resume, model, optimizer)
        start_epoch = info["epoch"] + 1
        if st.rank == 0:
            print(f"resumed from {resume}: continuing with epoch {start_epoch}")
    trainer = Trainer(model=model, optimizer=optimizer, train_dataloader=train_dataloader, gpu_id=local_rank,
                      loss_fn=loss_fn)
    for epoch in range(start_epoch, max_epochs):
        trainer._run_epoch(epoch)
        if save:
            b2.save_checkpoint(trainer.model, optimizer, epoch, save)
    b2.destroy()


if __name__ == "__main__":
    parser = argparse.ArgumentParser(description="simple distributed training job")
    parser.add_argument("--max_epochs", type=int, default=10, help="Total epochs to train the model")
    parser.add_argument("--batch_size", type=int, default=32, help="Input batch size on each device")
    parser.add_argument("--backend", default=None)
    parser.add_argument("--model", default="linear", choices=["linear", "mlp"])
    parser.add_argument("--save", default="")
    parser.add_argument("--resume", default="")
    args = parser.parse_args()
    main(args.max_epochs, args.batch_size, args.backend, args.model, args.save, args.resume)
