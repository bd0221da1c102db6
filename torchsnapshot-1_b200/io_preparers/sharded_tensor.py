"""Import path of the reference's sharded-tensor preparer (torchsnapshot/io_preparers/sharded_tensor.py)."""

from ..io_preparer import (  # noqa: F401
    _OverlappingRegion,
    ShardedTensorBufferConsumer,
    ShardedTensorIOPreparer,
)
