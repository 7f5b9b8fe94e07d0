"""Data plane: symmetric heap, bucket plan, sharded parameter-server engine."""
from .buckets import Bucket, BucketPlan, ParamSlot, plan_buckets
from .engine import ShardedParameterServer
from .sparse import ShardedEmbedding

__all__ = ["Bucket", "BucketPlan", "ParamSlot", "plan_buckets", "ShardedParameterServer",
           "ShardedEmbedding"]
