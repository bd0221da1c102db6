"""B200-native checkpoint hot path behind torchsnapshot's BufferStager / BufferConsumer /
StoragePlugin interfaces.

The directory name carries a hyphen, so the package is loaded under the module name
`torchsnapshot_b200` (see `tsb200_loader.py` at the repository root):

    import tsb200_loader; tsb = tsb200_loader.load()
    from torchsnapshot_b200.batcher import batch_write_requests
"""

from .version import __version__

__all__ = ["__version__", "Snapshot", "Stateful", "StateDict", "RNGState"]

_LAZY = {"Snapshot": "snapshot", "StateDict": "state_dict", "Stateful": "stateful", "RNGState": "rng_state"}


def __getattr__(name: str):
    # the reference's top-level names (torchsnapshot/__init__.py:10-36), imported on first use so that
    # `import torchsnapshot_b200._native` alone does not pull torch.distributed in
    if name in _LAZY:
        import importlib

        value = getattr(importlib.import_module(f"{__name__}.{_LAZY[name]}"), name)
        globals()[name] = value
        return value
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
