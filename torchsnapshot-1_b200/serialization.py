"""dtype tables and the payload byte layout contract.

The payload of a buffer-protocol tensor is its elements in row-major order, native
little-endian, no header, no padding: what tensor_as_memoryview yields in the reference
(torchsnapshot/serialization.py:162-188, after .contiguous() at 182-185).  On the B200 path
those bytes are produced by the gather-pack kernel, never by a host copy; the string tables
below are the manifest vocabulary (serialization.py:32-138).
"""

from enum import Enum
from typing import Dict, List

import torch


class Serializer(Enum):
    TORCH_SAVE = "torch_save"
    BUFFER_PROTOCOL = "buffer_protocol"


_ELEMENT_SIZE: Dict[torch.dtype, int] = {
    torch.float64: 8,
    torch.float32: 4,
    torch.float16: 2,
    torch.bfloat16: 2,
    torch.complex128: 16,
    torch.complex64: 8,
    torch.int64: 8,
    torch.int32: 4,
    torch.int16: 2,
    torch.int8: 1,
    torch.uint8: 1,
    torch.bool: 1,
    torch.qint32: 4,
    torch.qint8: 1,
    torch.quint8: 1,
}
ALL_SUPPORTED_DTYPES: List[torch.dtype] = list(_ELEMENT_SIZE)
SUPPORTED_QUANTIZED_DTYPES: List[torch.dtype] = [torch.qint32, torch.qint8, torch.quint8]
_DTYPE_TO_STRING: Dict[torch.dtype, str] = {dt: str(dt) for dt in ALL_SUPPORTED_DTYPES}
_STRING_TO_DTYPE: Dict[str, torch.dtype] = {s: dt for dt, s in _DTYPE_TO_STRING.items()}

# dtypes whose payload is the raw element bytes (reference serialization.py:148-159)
BUFFER_PROTOCOL_SUPPORTED_DTYPES: List[torch.dtype] = [
    torch.float64,
    torch.float32,
    torch.float16,
    torch.bfloat16,
    torch.int64,
    torch.int32,
    torch.int16,
    torch.int8,
    torch.uint8,
    torch.bool,
]
_BUFFER_PROTOCOL_SET = frozenset(BUFFER_PROTOCOL_SUPPORTED_DTYPES)


def is_buffer_protocol_dtype(dtype: torch.dtype) -> bool:
    return dtype in _BUFFER_PROTOCOL_SET


def dtype_to_string(dtype: torch.dtype) -> str:
    try:
        return _DTYPE_TO_STRING[dtype]
    except KeyError:
        raise ValueError(f"Unsupported dtype {dtype}. (Supported dtypes are: {ALL_SUPPORTED_DTYPES})") from None


def string_to_dtype(s: str) -> torch.dtype:
    try:
        return _STRING_TO_DTYPE[s]
    except KeyError:
        raise ValueError(f"Unsupported dtype {s}. (Supported dtypes are: {ALL_SUPPORTED_DTYPES})") from None


def dtype_to_element_size(dtype: torch.dtype) -> int:
    try:
        return _ELEMENT_SIZE[dtype]
    except KeyError:
        raise ValueError(f"Unsupported dtype {dtype}. (Supported dtypes are: {ALL_SUPPORTED_DTYPES})") from None
