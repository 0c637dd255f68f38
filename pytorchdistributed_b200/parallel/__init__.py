from .ddp import DistributedDataParallel, plan_buckets

__all__ = ["DistributedDataParallel", "plan_buckets"]
