from .data_parallel import DataParallel
from .ddp import DistributedDataParallel, plan_buckets
from .pipeline import ModelParallelResNet50, PipelineParallelResNet50

__all__ = ["DataParallel", "DistributedDataParallel", "plan_buckets", "ModelParallelResNet50",
           "PipelineParallelResNet50"]
