"""The notebook-01 walk-through as a script: toy dataset, the 10-20-20-20-5 MLP, ``DataParallel`` over every visible
GPU, a forward-only loop that prints what the notebook prints
(``R/01 数据并行/01_multi_gpus_data_parallelism.ipynb#cell1-6``: json L33-42 hyper-parameters, L52-63 dataset,
L81-84 shape check, L94-107 model, L117-121 wrapper, L139-145 loop).

    python scripts/demo_data_parallel.py [--batches 1] [--train]

Differences from the notebook, on purpose: the model is moved to ``cuda:0`` before wrapping (the notebook's cell only
works because ``nn.DataParallel`` moves a one-GPU model for you), and on a machine without a GPU the wrapper runs the
module as it is, so the script also runs (and is tested) on CPU.  ``--train`` adds what the notebook never does under
DataParallel: a loss, ``backward()`` (gradients reduced onto the first GPU) and an SGD step.
"""
import argparse
import os
import sys

import torch
from torch.utils.data import DataLoader

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorchdistributed_b200.data import SimpleDataset  # noqa: E402
from pytorchdistributed_b200.models import MLP  # noqa: E402
from pytorchdistributed_b200.parallel import DataParallel  # noqa: E402

input_size = 10
hidden_size = 20
output_size = 5
batch_size = 32
data_size = 1000


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", type=int, default=1, help="batches to run (the notebook breaks after the first)")
    ap.add_argument("--train", action="store_true", help="also run loss / backward / SGD step under DataParallel")
    args = ap.parse_args()
    random_loader = DataLoader(SimpleDataset(data_size, input_size, output_size), batch_size=batch_size, shuffle=True)
    for data, labels in random_loader:
        print(f"Data shape: {data.shape}, Labels shape: {labels.shape}")
        break
    device = torch.device("cuda:0") if torch.cuda.is_available() else torch.device("cpu")
    model = MLP(input_size, hidden_size, output_size).to(device)
    # device_ids=None: every visible GPU takes part; output_device=None: results are gathered on device_ids[0]
    model = DataParallel(model)
    opt = torch.optim.SGD(model.parameters(), lr=1e-3) if args.train else None
    for i, (inputs, labels) in enumerate(random_loader):
        if i >= args.batches:
            break
        input = inputs.to(device)
        if args.train:
            opt.zero_grad()
            output = model(input)
            loss = torch.nn.functional.mse_loss(output, labels.to(device).float())
            loss.backward()
            opt.step()
            print("Outside: input size", input.size(), "output_size", output.size(), f"loss {loss.item():.4f}")
        else:
            with torch.no_grad():
                output = model(input)
            print("Outside: input size", input.size(), "output_size", output.size())


if __name__ == "__main__":
    main()
