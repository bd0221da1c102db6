"""Flatten nested state dicts into {logical_path: leaf} and back.

Path grammar and container entries are the reference's (torchsnapshot/flatten.py:18-224): "/" joins
levels, "%" and "/" inside keys are percent-encoded, lists and (ordered) dicts become container
entries, dicts with colliding or non str/int keys stay opaque leaves.  Host-only bookkeeping; it is
here because the manifest format needs it, not because it is on the hot path."""

from collections import OrderedDict
from typing import Any, Dict, Tuple

from .manifest import DictEntry, ListEntry, Manifest, OrderedDictEntry


def _encode(s: str) -> str:
    if "%" in s or "/" in s:
        return s.replace("%", "%25").replace("/", "%2F")
    return s


def _decode(s: str) -> str:
    return s.replace("%2F", "/").replace("%25", "%")


_STR_ONLY = {str}


def _flattenable_dict(d: Dict[Any, Any]) -> bool:
    kinds = set(map(type, d))
    if not kinds or kinds == _STR_ONLY:
        return True  # distinct str keys cannot collide
    if not all(issubclass(k, (str, int)) for k in kinds):
        return False
    # mixed int/str keys may collide ("1" vs 1)
    return len({str(k) for k in d}) == len(d)


_CONTAINER_TYPES = frozenset((list, dict, OrderedDict))

# Child paths of the wide dicts flattened last (container path -> (keys, child paths)): a job that checkpoints the
# same state dict again gets the same 10-20 k path strings back instead of formatting them again (and everything
# downstream that compares or hashes paths meets the same objects).  A handful of entries, replaced as they come.
_WIDE = 1024
_wide_paths: Dict[str, Tuple[Any, Any]] = {}


def _child_paths(path: str, keys: Any) -> Any:
    wide = len(keys) >= _WIDE
    if wide:
        hit = _wide_paths.get(path)
        if hit is not None and hit[0] == keys:
            return hit[1]
    if set(map(type, keys)) == _STR_ONLY and "%" not in (joined := "".join(keys)) and "/" not in joined:
        out = [f"{path}/{k}" for k in keys]
    else:
        out = [f"{path}/{_encode(str(k))}" for k in keys]
    if wide:
        if len(_wide_paths) >= 8 and path not in _wide_paths:
            try:
                _wide_paths.pop(next(iter(_wide_paths)), None)
            except (StopIteration, RuntimeError):  # another thread emptied or resized it meanwhile
                pass
        _wide_paths[path] = (list(keys), out)
    return out


def flatten(obj: Any, prefix: str) -> Tuple[Manifest, Dict[str, Any]]:
    manifest: Manifest = {}
    flattened: Dict[str, Any] = {}
    stack = [(obj, _encode(prefix))]
    # iterative pre-order walk; children are pushed in reverse so dict order is preserved
    while stack:
        node, path = stack.pop()
        kind = type(node)
        if kind is list:
            manifest[path] = ListEntry()
            stack.extend((elem, f"{path}/{i}") for i, elem in reversed(list(enumerate(node))))
        elif (kind is dict or kind is OrderedDict) and _flattenable_dict(node):
            keys = list(node.keys())
            manifest[path] = DictEntry(keys=keys) if kind is dict else OrderedDictEntry(keys=keys)
            if len(keys) > 8 and _CONTAINER_TYPES.isdisjoint(map(type, node.values())):
                # a dict of leaves (a model's state dict, the bulk of a 20 k-tensor checkpoint): its leaves are
                # next in pre-order whatever is still waiting on the stack, so they go in as they are
                flattened.update(zip(_child_paths(path, keys), node.values()))
                continue
            stack.extend((node[k], f"{path}/{_encode(str(k))}") for k in reversed(keys))
        else:
            flattened[path] = node
    return manifest, flattened


def inflate(manifest: Manifest, flattened: Dict[str, Any], prefix: str) -> Any:
    prefix = _encode(prefix)
    below = prefix + "/"
    manifest = {k: v for k, v in manifest.items() if k == prefix or k.startswith(below)}
    flattened = {k: v for k, v in flattened.items() if k == prefix or k.startswith(below)}
    if prefix in flattened:
        return flattened[prefix]
    if prefix not in manifest:
        raise AssertionError(f"{prefix} is absent in both manifest and flattened.")
    containers: Dict[str, Any] = {}
    lists = set()
    for path, entry in manifest.items():
        kind = entry.type
        if kind == "list":
            containers[path] = []
            lists.add(path)
        else:
            containers[path] = OrderedDict() if kind == "OrderedDict" else {}
    # elements of lists: whatever is there, in index order (a list entry does not say how long the list was)
    elements: Dict[str, Dict[int, Any]] = {}
    if lists:
        for source in (containers, flattened):
            for path, val in source.items():
                parent, _, key = path.rpartition("/")
                if parent in lists:
                    elements.setdefault(parent, {})[int(key)] = val
    missing = object()
    for path, entry in manifest.items():
        container = containers[path]
        if path in lists:
            found = elements.get(path)
            if found:
                container.extend(found[i] for i in sorted(found))
            continue
        # original key objects (ints stay ints), original order; children are looked up by their paths
        keys = entry.keys
        if len(keys) > 8 and set(map(type, keys)) == _STR_ONLY and "%" not in (joined := "".join(keys)) and "/" not in joined:
            child_paths = [f"{path}/{k}" for k in keys]
        else:
            child_paths = [f"{path}/{_encode(str(k))}" for k in keys]
        for key, child in zip(keys, child_paths):
            val = flattened.get(child, missing)
            if val is missing:
                val = containers.get(child, missing)
                if val is missing:
                    continue
            container[key] = val
    return containers[prefix]
