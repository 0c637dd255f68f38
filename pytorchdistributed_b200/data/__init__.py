from .datasets import MyTrainDataset, SimpleDataset, SyntheticImageDataset, generate_random_data
from .sampler import DistributedSampler, ShardedSampler

__all__ = ["MyTrainDataset", "SimpleDataset", "SyntheticImageDataset", "generate_random_data",
           "DistributedSampler", "ShardedSampler"]
