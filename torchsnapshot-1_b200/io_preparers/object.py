"""Import path of the reference's object preparer (torchsnapshot/io_preparers/object.py)."""

from ..io_preparer import ObjectBufferConsumer, ObjectBufferStager, ObjectIOPreparer  # noqa: F401
