"""The reference splits its I/O preparers over `torchsnapshot/io_preparers/{tensor,chunked_tensor,
sharded_tensor,object}.py`; here they live in one module (`io_preparer.py`) because they share the region
builders of the CUDA path.  These modules keep the reference's import paths working."""
