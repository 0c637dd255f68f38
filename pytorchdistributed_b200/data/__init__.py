from .datasets import (DeviceSyntheticBatches, MyTrainDataset, SimpleDataset, SyntheticImageDataset,
                       generate_random_data)
from .sampler import DistributedSampler, ShardedSampler

__all__ = ["DeviceSyntheticBatches", "MyTrainDataset", "SimpleDataset", "SyntheticImageDataset", "generate_random_data",
           "DistributedSampler", "ShardedSampler"]
