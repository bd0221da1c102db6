"""URL -> StoragePlugin dispatch (reference torchsnapshot/storage_plugin.py:18-67).  Only the local
filesystem plugin is part of this build; other schemes resolve through the same `storage_plugins`
entry-point group the reference uses, so third-party plugins keep working."""

from typing import Any, Dict, Optional

from .io_types import StoragePlugin
from .storage_plugins.fs import FSStoragePlugin


def url_to_storage_plugin(url_path: str, storage_options: Optional[Dict[str, Any]] = None) -> StoragePlugin:
    if "://" in url_path:
        protocol, path = url_path.split("://", 1)
        if len(protocol) == 0:
            protocol = "fs"
    else:
        protocol, path = "fs", url_path
    if storage_options is None:
        storage_options = {}
    if protocol == "fs":
        return FSStoragePlugin(root=path, storage_options=storage_options)
    from importlib.metadata import entry_points

    eps = entry_points(group="storage_plugins")
    registered = {ep.name: ep for ep in eps}
    if protocol in registered:
        factory = registered[protocol].load()
        plugin = factory(path, storage_options)
        return plugin
    raise RuntimeError(f"Unsupported protocol: {protocol}.")
