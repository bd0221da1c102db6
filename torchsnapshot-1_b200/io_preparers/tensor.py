"""Import path of the reference's dense-tensor preparer (torchsnapshot/io_preparers/tensor.py)."""

from ..io_preparer import (  # noqa: F401
    tensor_copy,
    TensorBufferConsumer,
    TensorBufferStager,
    TensorIOPreparer,
)
