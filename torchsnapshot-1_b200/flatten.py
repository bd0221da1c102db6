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


def _flattenable_dict(d: Dict[Any, Any]) -> bool:
    all_str = True
    for k in d:
        if type(k) is not str:
            all_str = False
            if not isinstance(k, (str, int)):
                return False
    # distinct str keys cannot collide; mixed int/str keys may ("1" vs 1)
    return all_str or len({str(k) for k in d}) == len(d)


_CONTAINER_TYPES = frozenset((list, dict, OrderedDict))


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
                if all(type(k) is str for k in keys) and "%" not in (joined := "".join(keys)) and "/" not in joined:
                    flattened.update(zip([f"{path}/{k}" for k in keys], node.values()))
                else:
                    flattened.update(zip([f"{path}/{_encode(str(k))}" for k in keys], node.values()))
                continue
            stack.extend((node[k], f"{path}/{_encode(str(k))}") for k in reversed(keys))
        else:
            flattened[path] = node
    return manifest, flattened


def inflate(manifest: Manifest, flattened: Dict[str, Any], prefix: str) -> Any:
    prefix = _encode(prefix)
    manifest = {k: v for k, v in manifest.items() if k.split("/")[0] == prefix}
    flattened = {k: v for k, v in flattened.items() if k.split("/")[0] == prefix}
    if prefix in flattened:
        return flattened[prefix]
    if prefix not in manifest:
        raise AssertionError(f"{prefix} is absent in both manifest and flattened.")
    containers: Dict[str, Any] = {}
    for path, entry in manifest.items():
        kind = entry.type
        containers[path] = [] if kind == "list" else (OrderedDict() if kind == "OrderedDict" else {})
    children: Dict[str, Dict[str, Any]] = {}
    for path, val in list(containers.items()) + list(flattened.items()):
        if path == prefix:
            continue
        parent, _, key = path.rpartition("/")
        children.setdefault(parent, {})[key] = val
    for path, vals in children.items():
        container = containers[path]
        entry = manifest[path]
        if isinstance(container, list):
            for i in sorted(int(k) for k in vals):
                container.append(vals[str(i)])
        else:
            for key in entry.keys:  # original key objects (ints stay ints), original order
                enc = _encode(str(key))
                if enc in vals:
                    container[key] = vals[enc]
    return containers[prefix]
