"""Import path of the reference's chunked-tensor preparer (torchsnapshot/io_preparers/chunked_tensor.py)."""

from ..io_preparer import Chunk, ChunkedTensorIOPreparer  # noqa: F401
