"""torchrun-launched DDP training - same flags and output as ``R/02 DDP基本概念/ddp_gpus_torchrun.py:86-102``.

    torchrun --nproc-per-node=2 --master-addr 127.0.0.1 --master-port=12355 \
        scripts/train_ddp_torchrun.py --max_epochs 5 --batch_size 32

Rank, world size and rendezvous come from the environment torchrun sets.  Additive flags (defaults keep the
reference's behaviour): ``--model {linear,resnet50}``, ``--save_every``, ``--snapshot_path``, ``--resume``,
``--loss {ce,mse}``, ``--img_size``, ``--dataset_size``, and the communication knobs
``--backend {auto,p2p,nccl,gloo}`` (p2p = our NVLink / NVSwitch kernels average the gradients, nccl = torch's NCCL
all-reduce on the same buckets, gloo = CPU plumbing), ``--dtype {fp32,bf16}`` (format of the gradients on the wire),
``--bucket_cap_mb``, ``--cuda_graph``.
"""
import argparse
import os
import sys

import torch
from torch.utils.data import DataLoader

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from pytorchdistributed_b200.comm import ddp_setup, destroy_process_group  # noqa: E402
from pytorchdistributed_b200.data import DistributedSampler, MyTrainDataset, SyntheticImageDataset  # noqa: E402
from pytorchdistributed_b200.models import ToyLinear, resnet50  # noqa: E402
from pytorchdistributed_b200.optim import SGD  # noqa: E402
from pytorchdistributed_b200.train import Trainer  # noqa: E402


def ddp_kwargs(args) -> dict:
    """``DistributedDataParallel`` keywords for the communication flags."""
    kw = {"bucket_cap_mb": args.bucket_cap_mb}
    if args.backend == "p2p":
        kw["comm"] = "p2p"
    elif args.backend in ("nccl", "gloo"):
        kw["comm"] = "torch"
    if args.dtype == "bf16":
        kw["wire_dtype"] = torch.bfloat16
    return kw


def main(max_epochs: int, batch_size: int, args):
    ddp_setup(backend="gloo" if args.backend == "gloo" else None)
    if args.model == "resnet50":
        train_dataset = SyntheticImageDataset(args.dataset_size, (args.img_size, args.img_size))
        model = resnet50()
    else:
        train_dataset = MyTrainDataset(2048)
        model = ToyLinear(20, 1)
    sampler = DistributedSampler(train_dataset, shuffle=True)
    train_dataloader = DataLoader(train_dataset, batch_size=batch_size, pin_memory=torch.cuda.is_available(),
                                  sampler=sampler)
    optimizer = SGD(model.parameters(), lr=1e-3)
    loss_fn = torch.nn.functional.mse_loss if args.loss == "mse" else None
    trainer = Trainer(model=model, optimizer=optimizer, train_dataloader=train_dataloader,
                      save_every=args.save_every, snapshot_path=args.snapshot_path, resume=args.resume, loss_fn=loss_fn,
                      ddp_kwargs=ddp_kwargs(args), cuda_graph=args.cuda_graph)
    trainer.train(max_epochs)
    destroy_process_group()


if __name__ == "__main__":
    parser = argparse.ArgumentParser(description="DDP Example")
    parser.add_argument("--max_epochs", type=int, default=10, help="number of epochs to train")
    parser.add_argument("--batch_size", type=int, default=32, help="batch size")
    parser.add_argument("--model", choices=["linear", "resnet50"], default="linear")
    parser.add_argument("--save_every", type=int, default=0)
    parser.add_argument("--snapshot_path", default="snapshot.pt")
    parser.add_argument("--resume", action="store_true")
    parser.add_argument("--loss", choices=["ce", "mse"], default="ce")
    parser.add_argument("--img_size", type=int, default=224)
    parser.add_argument("--dataset_size", type=int, default=2048)
    parser.add_argument("--backend", choices=["auto", "p2p", "nccl", "gloo"], default="auto")
    parser.add_argument("--dtype", choices=["fp32", "bf16"], default="fp32", help="gradient format on the wire")
    parser.add_argument("--bucket_cap_mb", type=float, default=25)
    parser.add_argument("--cuda_graph", action="store_true", help="replay the whole step from a CUDA graph")
    args = parser.parse_args()
    print(f"local_rank: {int(os.environ['LOCAL_RANK'])}, world_size: {torch.cuda.device_count()}")
    main(args.max_epochs, args.batch_size, args)
