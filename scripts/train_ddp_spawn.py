"""Spawn-launched DDP training - same flags and output as ``R/02 DDP基本概念/ddp_gpus.py:86-98``.

    python scripts/train_ddp_spawn.py --max_epochs 5 --batch_size 32

``--max_epochs`` has no default (as in the reference).  One process per visible GPU is started with
``torch.multiprocessing.spawn``; the dataset is built once in the parent and shipped to the workers.
Additive flags: ``--world_size`` (needed on a CPU-only box), ``--save_every``, ``--snapshot_path``,
``--loss {ce,mse}``, ``--port``.
"""
import argparse
import os
import sys

import torch
import torch.multiprocessing as mp
from torch.utils.data import DataLoader

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from pytorchdistributed_b200.comm import ddp_setup, destroy_process_group  # noqa: E402
from pytorchdistributed_b200.data import DistributedSampler, MyTrainDataset  # noqa: E402
from pytorchdistributed_b200.models import ToyLinear  # noqa: E402
from pytorchdistributed_b200.optim import SGD  # noqa: E402
from pytorchdistributed_b200.train import Trainer  # noqa: E402


def main(rank: int, world_size: int, max_epochs: int, batch_size: int, train_dataset, extra: dict):
    os.environ["MASTER_PORT"] = str(extra["port"])
    ddp_setup(rank, world_size, backend="gloo" if extra["backend"] == "gloo" else None)
    train_dataloader = DataLoader(train_dataset, batch_size=batch_size, pin_memory=torch.cuda.is_available(),
                                  shuffle=False, sampler=DistributedSampler(train_dataset))
    model = ToyLinear(20, 1)
    optimizer = SGD(model.parameters(), lr=1e-3)
    loss_fn = torch.nn.functional.mse_loss if extra["loss"] == "mse" else None
    trainer = Trainer(model=model, gpu_id=rank, optimizer=optimizer, train_dataloader=train_dataloader,
                      save_every=extra["save_every"], snapshot_path=extra["snapshot_path"], loss_fn=loss_fn,
                      ddp_kwargs=extra["ddp_kwargs"])
    trainer.train(max_epochs)
    destroy_process_group()


if __name__ == "__main__":
    train_dataset = MyTrainDataset(2048)
    parser = argparse.ArgumentParser(description="simple distributed training job")
    parser.add_argument("--max_epochs", type=int, help="Total epochs to train the model")
    parser.add_argument("--batch_size", default=32, type=int, help="Input batch size on each device (default: 32)")
    parser.add_argument("--world_size", type=int, default=None, help="processes to start (default: visible GPUs)")
    parser.add_argument("--save_every", type=int, default=0)
    parser.add_argument("--snapshot_path", default="snapshot.pt")
    parser.add_argument("--loss", choices=["ce", "mse"], default="ce")
    parser.add_argument("--port", type=int, default=12355)
    # communication knobs (additive; see scripts/train_ddp_torchrun.py)
    parser.add_argument("--backend", choices=["auto", "p2p", "nccl", "gloo"], default="auto")
    parser.add_argument("--dtype", choices=["fp32", "bf16"], default="fp32", help="gradient format on the wire")
    parser.add_argument("--bucket_cap_mb", type=float, default=25)
    args = parser.parse_args()

    world_size = args.world_size or torch.cuda.device_count() or 2
    kw = {"bucket_cap_mb": args.bucket_cap_mb}
    if args.backend != "auto":
        kw["comm"] = "p2p" if args.backend == "p2p" else "torch"
    if args.dtype == "bf16":
        kw["wire_dtype"] = torch.bfloat16
    extra = dict(save_every=args.save_every, snapshot_path=args.snapshot_path, loss=args.loss, port=args.port,
                 backend=args.backend, ddp_kwargs=kw)
    mp.spawn(main, args=(world_size, args.max_epochs, args.batch_size, train_dataset, extra), nprocs=world_size)
