"""Manifest entries and the `.snapshot_metadata` document.

The on-disk format is the reference's and is not changed by this build: a JSON document
(json is a yaml subset) `{"version", "world_size", "manifest": {path: entry}}` where every entry
is a tagged union keyed by "type" (reference torchsnapshot/manifest.py:28-314).  Only the
vocabulary is restated here; the field names and their order in the serialised form are what the
reference writes, so either implementation loads the other's snapshots.
"""

import base64
import json
import struct
from dataclasses import dataclass, field, fields
from typing import Any, Dict, List, Optional, Tuple, Union


@dataclass
class Entry:
    type: str = field(init=False)

    def to_obj(self) -> Dict[str, Any]:
        return {f.name: _to_obj(getattr(self, f.name)) for f in fields(self)}


def _to_obj(v: Any) -> Any:
    if isinstance(v, (Entry, Shard)):
        return v.to_obj()
    if isinstance(v, list):
        return [_to_obj(x) for x in v]
    return v


@dataclass
class TensorEntry(Entry):
    location: str
    serializer: str
    dtype: str
    shape: List[int]
    replicated: bool
    byte_range: Optional[List[int]] = None

    def __post_init__(self) -> None:
        self.type = "Tensor"

    @property
    def byte_range_tuple(self) -> Optional[Tuple[int, int]]:
        br = self.byte_range
        return None if br is None else (br[0], br[1])


@dataclass
class Shard:
    offsets: List[int]
    sizes: List[int]
    tensor: TensorEntry

    def to_obj(self) -> Dict[str, Any]:
        return {"offsets": list(self.offsets), "sizes": list(self.sizes), "tensor": self.tensor.to_obj()}


@dataclass
class ShardedTensorEntry(Entry):
    shards: List[Shard]

    def __post_init__(self) -> None:
        self.type = "ShardedTensor"


@dataclass
class ChunkedTensorEntry(Entry):
    dtype: str
    shape: List[int]
    chunks: List[Shard]
    replicated: bool

    def __post_init__(self) -> None:
        self.type = "ChunkedTensor"


@dataclass
class ObjectEntry(Entry):
    location: str
    serializer: str
    obj_type: str
    replicated: bool

    def __post_init__(self) -> None:
        self.type = "object"


@dataclass
class ListEntry(Entry):
    def __post_init__(self) -> None:
        self.type = "list"


@dataclass
class DictEntry(Entry):
    keys: List[Union[str, int]]

    def __post_init__(self) -> None:
        self.type = "dict"


@dataclass
class OrderedDictEntry(Entry):
    keys: List[Union[str, int]]

    def __post_init__(self) -> None:
        self.type = "OrderedDict"


_PRIMITIVE_TYPES = ("int", "str", "bool", "bytes", "float")


@dataclass
class PrimitiveEntry(Entry):
    """int/str/bool/bytes/float stored inline in the metadata (reference manifest.py:186-270)."""

    serialized_value: str
    replicated: bool
    readable: Optional[str] = None
    type_name: str = field(default="int", repr=False)

    supported_types = list(_PRIMITIVE_TYPES)

    def __post_init__(self) -> None:
        self.type = self.type_name

    def to_obj(self) -> Dict[str, Any]:
        return {
            "type": self.type,
            "serialized_value": self.serialized_value,
            "replicated": self.replicated,
            "readable": self.readable,
        }

    def get_value(self) -> Union[int, str, bool, bytes, float]:
        t, s = self.type, self.serialized_value
        if t == "int":
            return int(s)
        if t == "str":
            return s
        if t == "bool":
            if s not in ("True", "False"):
                raise RuntimeError(f"Unexpected serialized_value for bool type: {s}")
            return s == "True"
        if t == "bytes":
            return base64.b64decode(s.encode("utf-8"))
        if t == "float":
            return struct.unpack("d", base64.b64decode(s.encode("utf-8")))[0]
        raise ValueError(f"Unable to get deserialized value for {s}")

    @classmethod
    def from_object(cls, obj: Any) -> "PrimitiveEntry":
        t = type(obj).__name__
        if t not in _PRIMITIVE_TYPES:
            raise TypeError(f"Unsupported primitive obj of type {t}")
        if t == "bytes":
            s = base64.b64encode(obj).decode("utf-8")
        elif t == "float":
            s = base64.b64encode(struct.pack("d", float(obj))).decode("utf-8")
        else:
            s = str(obj)
        return cls(serialized_value=s, replicated=False, readable=str(obj) if t == "float" else None, type_name=t)


Manifest = Dict[str, Entry]

try:  # loader for metadata files written as real yaml (reference manifest.py:20-25)
    from yaml import CSafeLoader as Loader
except ImportError:  # pragma: no cover
    try:
        from yaml import SafeLoader as Loader
    except ImportError:
        Loader = None

_esc = json.encoder.encode_basestring_ascii  # C implementation; same escaping as json.dumps


def _scalar_list_text(vals: Any, pad: str) -> str:
    """A JSON array of scalars laid out the way json.dumps(indent=2) does: one element per line."""
    if vals is None:
        return "null"
    if not vals:
        return "[]"
    inner = pad + "  "
    return "[\n" + ",\n".join(inner + (str(v) if type(v) is int else json.dumps(v)) for v in vals) + "\n" + pad + "]"


def _tensor_text(e: "TensorEntry", pad: str) -> str:
    p = pad + "  "
    return (
        "{\n"
        f'{p}"type": "Tensor",\n'
        f'{p}"location": {_esc(e.location)},\n'
        f'{p}"serializer": {_esc(e.serializer)},\n'
        f'{p}"dtype": {_esc(e.dtype)},\n'
        f'{p}"shape": {_scalar_list_text(e.shape, p)},\n'
        f'{p}"replicated": {"true" if e.replicated else "false"},\n'
        f'{p}"byte_range": {_scalar_list_text(e.byte_range, p)}\n'
        f"{pad}}}"
    )


def _shards_text(shards: List["Shard"], pad: str) -> str:
    if not shards:
        return "[]"
    p, q = pad + "  ", pad + "    "
    items = [
        "{\n"
        f'{q}"offsets": {_scalar_list_text(s.offsets, q)},\n'
        f'{q}"sizes": {_scalar_list_text(s.sizes, q)},\n'
        f'{q}"tensor": {_tensor_text(s.tensor, q)}\n'
        f"{p}}}"
        for s in shards
    ]
    return "[\n" + ",\n".join(p + it for it in items) + "\n" + pad + "]"


# Entries that are known not to change any more (batcher.TakePlanCache registers the entries of a cached plan,
# which successive snapshots of one training job share): their text is produced once.  id -> (entry, text);
# the entry is held so that the id cannot be reused by another object.
_stable_text: Dict[int, Any] = {}


def register_stable_entries(entries: Any) -> None:
    """Replace the set of entries whose serialised text may be memoised (top-level manifest entries only)."""
    _stable_text.clear()
    for e in entries:
        _stable_text[id(e)] = [e, None]


def _entry_text(entry: Entry, pad: str) -> str:
    kind = type(entry)
    if kind is TensorEntry:
        memo = _stable_text.get(id(entry)) if pad == "    " else None
        if memo is None:
            return _tensor_text(entry, pad)
        if memo[1] is None:
            memo[1] = _tensor_text(entry, pad)
        return memo[1]
    p = pad + "  "
    if kind is ShardedTensorEntry:
        return "{\n" f'{p}"type": "ShardedTensor",\n{p}"shards": {_shards_text(entry.shards, p)}\n{pad}}}'
    if kind is ChunkedTensorEntry:
        return (
            "{\n"
            f'{p}"type": "ChunkedTensor",\n'
            f'{p}"dtype": {_esc(entry.dtype)},\n'
            f'{p}"shape": {_scalar_list_text(entry.shape, p)},\n'
            f'{p}"chunks": {_shards_text(entry.chunks, p)},\n'
            f'{p}"replicated": {"true" if entry.replicated else "false"}\n'
            f"{pad}}}"
        )
    # containers, objects, primitives: few per snapshot; the generic encoder, re-indented
    return json.dumps(entry.to_obj(), sort_keys=False, indent=2).replace("\n", "\n" + pad)


_TENSOR_KEYS = ("type", "location", "serializer", "dtype", "shape", "replicated", "byte_range")


def _tensor_from_obj(o: Dict[str, Any]) -> TensorEntry:
    if len(o) == 7 and tuple(o) == _TENSOR_KEYS:
        # the document already is the entry's attribute dict (restore of 20 k entries: 3x faster)
        e = TensorEntry.__new__(TensorEntry)
        e.__dict__ = o
        return e
    return TensorEntry(
        location=o["location"],
        serializer=o["serializer"],
        dtype=o["dtype"],
        shape=list(o["shape"]),
        replicated=o["replicated"],
        byte_range=o.get("byte_range"),
    )


def _shard_from_obj(o: Dict[str, Any]) -> Shard:
    return Shard(offsets=list(o["offsets"]), sizes=list(o["sizes"]), tensor=_tensor_from_obj(o["tensor"]))


def entry_from_obj(o: Dict[str, Any]) -> Entry:
    t = o["type"]
    if t == "Tensor":
        return _tensor_from_obj(o)
    if t == "ShardedTensor":
        return ShardedTensorEntry(shards=[_shard_from_obj(s) for s in o["shards"]])
    if t == "ChunkedTensor":
        return ChunkedTensorEntry(
            dtype=o["dtype"], shape=list(o["shape"]), chunks=[_shard_from_obj(s) for s in o["chunks"]],
            replicated=o["replicated"],
        )
    if t == "object":
        return ObjectEntry(location=o["location"], serializer=o["serializer"], obj_type=o["obj_type"],
                           replicated=o["replicated"])
    if t == "list":
        return ListEntry()
    if t == "dict":
        return DictEntry(keys=list(o["keys"]))
    if t == "OrderedDict":
        return OrderedDictEntry(keys=list(o["keys"]))
    if t in _PRIMITIVE_TYPES:
        return PrimitiveEntry(serialized_value=o["serialized_value"], replicated=o["replicated"],
                              readable=o.get("readable"), type_name=t)
    raise ValueError(f"Unsupported entry type: {t}")


@dataclass
class SnapshotMetadata:
    version: str
    world_size: int
    manifest: Manifest

    def to_yaml(self) -> str:
        """The `.snapshot_metadata` text: character for character what the reference's
        `json.dumps(asdict(self), sort_keys=False, indent=2)` gives (manifest.py:283-289), produced by a
        schema-aware writer because the generic encoder with `indent` runs in pure Python (0.45 s for
        13 k entries, on the critical path of a take)."""
        pad = "    "
        parts = [f"{pad}{_esc(path)}: {_entry_text(entry, pad)}" for path, entry in self.manifest.items()]
        manifest = "{\n" + ",\n".join(parts) + "\n  }" if parts else "{}"
        return (
            "{\n"
            f'  "version": {_esc(self.version)},\n'
            f'  "world_size": {json.dumps(self.world_size)},\n'
            f'  "manifest": {manifest}\n'
            "}"
        )

    @classmethod
    def from_yaml(cls, yaml_str: str) -> "SnapshotMetadata":
        try:
            doc = json.loads(yaml_str)  # what to_yaml() has written since the reference switched to json
        except json.JSONDecodeError:
            import yaml  # legacy snapshots written as real yaml

            doc = yaml.load(yaml_str, Loader=Loader if Loader is not None else yaml.SafeLoader)
        manifest: Manifest = {}
        new, cls_tensor, keys = object.__new__, TensorEntry, _TENSOR_KEYS
        for path, o in doc["manifest"].items():
            if len(o) == 7 and o.get("type") == "Tensor" and tuple(o) == keys:
                # the document already is the entry's attribute dict (practically every entry of a big snapshot)
                e = new(cls_tensor)
                e.__dict__ = o
                manifest[path] = e
            else:
                manifest[path] = entry_from_obj(o)
        return cls(version=doc["version"], world_size=doc["world_size"], manifest=manifest)


# ------------------------------------------------------------------------------------------------
# compact wire form of a rank's manifest (the take-side all-gather, reference snapshot.py:352-358)
# ------------------------------------------------------------------------------------------------
# The reference gathers every rank's manifest as pickled entry objects; with 13 k tensors per rank that is most
# of the time an 8-rank job is blocked in async_take (each rank unpickles 100 k dataclass instances).  Here a
# rank sends its TensorEntry objects -- practically all of a manifest -- as columns of strings and ints, which
# pickle an order of magnitude faster, and a receiver only rebuilds entries when somebody asks for them
# (`SnapshotMetadata.lazy`): rank 0 in the thread that writes the metadata file, the others not at all unless
# the snapshot object is used to restore.
def encode_manifest(manifest: Manifest, tensor_columns: Optional[Any] = None, rest: Optional[Manifest] = None,
                    parts: bool = False) -> Any:
    """`manifest` as bytes (`parts`: as the two byte strings the wire form is the pickle of, for
    `manifest_digest`).  `tensor_columns` = (entries dict, pickled columns) from an earlier call with the SAME
    TensorEntry objects (see `tensor_columns_of`) skips re-encoding them; they then come last in the decoded
    manifest.  `rest`, with `tensor_columns`, is the caller's statement that `manifest` is exactly `rest` plus
    the entries of `tensor_columns`: only `rest` is walked then."""
    import pickle

    cached: Dict[str, Any] = tensor_columns[0] if tensor_columns is not None else {}
    if rest is not None and tensor_columns is not None:
        if len(rest) + len(cached) != len(manifest):
            raise AssertionError("encode_manifest: `rest` and the cached entries do not add up to the manifest")
        manifest = rest
    paths, locations, serializers, dtypes, shapes, replicated, ranges, others = [], [], [], [], [], [], [], []
    for path, e in manifest.items():
        if type(e) is TensorEntry:
            if cached and cached.get(path) is e:
                continue
            paths.append(path)
            locations.append(e.location)
            serializers.append(e.serializer)
            dtypes.append(e.dtype)
            shapes.append(e.shape)
            replicated.append(e.replicated)
            ranges.append(e.byte_range)
        else:
            others.append((len(paths), path, e))  # position among the tensor entries, to restore the order
    head = pickle.dumps((paths, locations, serializers, dtypes, shapes, replicated, ranges, others), protocol=5)
    if parts:
        return head, tensor_columns[1] if tensor_columns is not None else None
    return pickle.dumps((head, tensor_columns[1] if tensor_columns is not None else None), protocol=5)


def manifest_digest(head: bytes, tensor_columns: Optional[Any]) -> bytes:
    """16 + 16 bytes that change when the encoded manifest (`head`, plus the cached columns when there are any)
    changes: the columns' digest was made once, with the columns."""
    import hashlib

    tail = tensor_columns[2] if tensor_columns is not None else b"\0" * 16
    return hashlib.blake2b(head, digest_size=16).digest() + tail


def tensor_columns_of(entries: Dict[str, Any]) -> Any:
    """Pre-encoded columns of a dict of TensorEntry objects that will be sent unchanged take after take."""
    import pickle

    vals = list(entries.values())
    blob = pickle.dumps((list(entries.keys()), [e.location for e in vals], [e.serializer for e in vals],
                         [e.dtype for e in vals], [e.shape for e in vals], [e.replicated for e in vals],
                         [e.byte_range for e in vals], []), protocol=5)
    import hashlib

    return (entries, blob, hashlib.blake2b(blob, digest_size=16).digest())


def _decode_columns(blob: bytes, out: Manifest) -> None:
    import pickle

    paths, locations, serializers, dtypes, shapes, replicated, ranges, others = pickle.loads(blob)
    new, cls = object.__new__, TensorEntry
    k = 0
    for i, path in enumerate(paths):
        while k < len(others) and others[k][0] <= i:
            out[others[k][1]] = others[k][2]
            k += 1
        e = new(cls)
        e.__dict__ = {"type": "Tensor", "location": locations[i], "serializer": serializers[i], "dtype": dtypes[i],
                      "shape": shapes[i], "replicated": replicated[i], "byte_range": ranges[i]}
        out[path] = e
    for _pos, path, e in others[k:]:
        out[path] = e


def decode_manifest(blob: bytes) -> Manifest:
    import pickle

    head, tail = pickle.loads(blob)
    out: Manifest = {}
    _decode_columns(head, out)
    if tail is not None:
        _decode_columns(tail, out)
    return out


class LazySnapshotMetadata(SnapshotMetadata):
    """SnapshotMetadata whose manifest is rebuilt from the ranks' encoded manifests on first use."""

    def __init__(self, version: str, world_size: int, parts: List[Any]) -> None:
        self.version = version
        self.world_size = world_size
        self._parts = parts  # per rank: its manifest as a dict, or encoded (`encode_manifest`)

    def __getattr__(self, name: str) -> Any:
        if name == "manifest":
            manifest: Manifest = {}
            for rank, part in enumerate(self.__dict__["_parts"]):
                prefix = f"{rank}/"
                for path, entry in (decode_manifest(part) if isinstance(part, bytes) else part).items():
                    manifest[prefix + path] = entry
            self.manifest = manifest
            return manifest
        raise AttributeError(name)


def is_dict_entry(entry: Any) -> bool:
    return getattr(entry, "type", None) in ("dict", "OrderedDict")


def is_container_entry(entry: Any) -> bool:
    return getattr(entry, "type", None) in ("list", "dict", "OrderedDict")


def is_replicated(entry: Any) -> bool:
    return bool(getattr(entry, "replicated", False))
