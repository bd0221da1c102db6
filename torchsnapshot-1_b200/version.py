"""Version of the snapshot metadata schema this build writes: the reference's (torchsnapshot/version.py:17),
so its `Snapshot` opens these snapshots and vice versa."""

__version__: str = "0.1.0"
