from .data_parallel import DataParallel
from .ddp import DistributedDataParallel, plan_buckets
from .pipeline import ModelParallelResNet50, PipelineParallelResNet50
from .pipeline_mp import PipelineStage, max_live_activations, schedule_ops

__all__ = ["DataParallel", "DistributedDataParallel", "plan_buckets", "ModelParallelResNet50",
           "PipelineParallelResNet50", "PipelineStage", "schedule_ops", "max_live_activations"]
