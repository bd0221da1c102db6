"""B200-native checkpoint hot path behind torchsnapshot's BufferStager / BufferConsumer /
StoragePlugin interfaces.

The directory name carries a hyphen, so the package is loaded under the module name
`torchsnapshot_b200` (see `tsb200_loader.py` at the repository root):

    import tsb200_loader; tsb = tsb200_loader.load()
    from torchsnapshot_b200.batcher import batch_write_requests
"""

__version__ = "0.1.0"
