"""pytorchdistributed_b200 - Blackwell-native data-/model-parallel training framework.

Public surface (mirrors what JoeyOL/PytorchDistributed exercises through torch):

* ``parallel.DataParallel``            single-process replicate / scatter / gather wrapper
* ``parallel.DistributedDataParallel`` one process per GPU, bucketed gradient all-reduce
* ``parallel.ModelParallelResNet50`` / ``PipelineParallelResNet50(split_size=...)``
* ``comm.ddp_setup`` / ``destroy_process_group``, ``data.DistributedSampler``, ``train.Trainer``
* ``models.resnet50``, ``models.ToyLinear``, ``models.MLP``; ``ops`` for the individual kernels
"""
__version__ = "0.1.0"

from . import _lib  # noqa: F401

__all__ = ["comm", "data", "models", "ops", "parallel", "train", "utils", "optim"]
