"""dtype tables and the payload byte layout contract.

The payload of a buffer-protocol tensor is its elements in row-major order, native
little-endian, no header, no padding: what tensor_as_memoryview yields in the reference
(torchsnapshot/serialization.py:162-188, after .contiguous() at 182-185).  On the B200 path
those bytes are produced by the gather-pack kernel, never by a host copy; the string tables
below are the manifest vocabulary (serialization.py:32-138).
"""

import io
import struct
import warnings
from enum import Enum
from functools import reduce
from operator import mul
from typing import Dict, List, Sequence, Union

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


# ----------------------------------------------------------------------------------------------
# host-side codecs (CPU tensors; the B200 path produces the same bytes with the pack kernel)
# ----------------------------------------------------------------------------------------------
def tensor_as_memoryview(tensor: torch.Tensor) -> memoryview:
    """Zero-copy byte view of a CPU tensor in the payload layout (reference serialization.py:162-188).

    Unlike the reference, bf16 tensors with an odd element count keep their last element (the
    reference drops two bytes there).
    """
    if not is_buffer_protocol_dtype(tensor.dtype):
        raise ValueError(f"tensor_as_memoryview() doesn't support the dtype {tensor.dtype}.")
    if tensor.device != torch.device("cpu"):
        raise ValueError("tensor_as_memoryview() only accepts CPU tensors.")
    flat = tensor.detach().contiguous().reshape(-1)
    return memoryview(flat.view(torch.uint8).numpy()).cast("b")


# tensors made from read-only buffers are intermediate, read-only views (reference serialization.py:236-244)
warnings.filterwarnings("ignore", message="The given buffer is not writable")


def tensor_from_memoryview(mv: Union[bytes, bytearray, memoryview], dtype: torch.dtype, shape: Sequence[int]) -> torch.Tensor:
    if len(mv) == 0:
        return torch.empty(list(shape), dtype=dtype)
    return torch.frombuffer(mv, dtype=dtype).reshape(list(shape))



def torch_save_as_bytes(tensor: torch.Tensor) -> bytes:
    buf = io.BytesIO()
    torch.save(tensor, buf)
    return buf.getvalue()


def torch_load_from_bytes(buf: bytes) -> torch.Tensor:
    return torch.load(io.BytesIO(buf), weights_only=False)


# Quantized-tensor codecs (reference serialization.py:257-456).  The reference never routes a snapshot
# through them (quantized dtypes are pickled by torch.save, tensor.py:68-71); they are kept so that code
# and tests written against the reference's module find the same functions with the same byte formats:
#   per-tensor : int representation | q_scale (C double) | q_zero_point (C long long)
#   per-channel: axis (C long long) | int representation | scales (float64[C]) | zero points (int64[C])
def _int_repr_bytes(tensor: torch.Tensor) -> bytes:
    return tensor.int_repr().contiguous().reshape(-1).view(torch.uint8).numpy().tobytes()


_QUANTIZED_INT_DTYPE = {torch.qint8: torch.int8, torch.quint8: torch.uint8, torch.qint32: torch.int32}


def _ints_from_bytes(data: bytes, dtype: torch.dtype, shape: Sequence[int]) -> torch.Tensor:
    if dtype not in _QUANTIZED_INT_DTYPE:
        raise RuntimeError(f"{dtype} is not a supported quantized dtype ({SUPPORTED_QUANTIZED_DTYPES}).")
    return torch.frombuffer(bytearray(data), dtype=_QUANTIZED_INT_DTYPE[dtype]).reshape(list(shape))


def per_tensor_qtensor_as_bytes(tensor: torch.Tensor) -> bytes:
    if not tensor.is_quantized or tensor.qscheme() != torch.per_tensor_affine:
        raise RuntimeError("per_tensor_qtensor_as_bytes() only supports per_tensor_affine quantized tensor.")
    return _int_repr_bytes(tensor) + struct.pack("d", tensor.q_scale()) + struct.pack("q", tensor.q_zero_point())


def per_tensor_qtensor_from_bytes(buf: bytes, dtype: torch.dtype, shape: List[int]) -> torch.Tensor:
    nbytes = reduce(mul, shape, 1) * dtype_to_element_size(dtype)
    if len(buf) != nbytes + 16:
        raise RuntimeError(
            f"The expected buffer size for the per-tensor quantized tensor (dtype: {dtype}, shape: {shape}) "
            f"is {nbytes + 16}. The size of the input buffer is {len(buf)}."
        )
    scale, zero_point = struct.unpack("dq", bytes(buf[nbytes:]))
    return torch._make_per_tensor_quantized_tensor(_ints_from_bytes(buf[:nbytes], dtype, shape), scale, zero_point)


def per_channel_qtensor_as_bytes(tensor: torch.Tensor) -> bytes:
    if not tensor.is_quantized or tensor.qscheme() not in (torch.per_channel_affine, torch.per_channel_affine_float_qparams):
        raise RuntimeError("per_channel_qtensor_as_bytes() only supports per_channel_affine quantized tensor.")
    return b"".join((
        struct.pack("q", tensor.q_per_channel_axis()),
        _int_repr_bytes(tensor),
        tensor.q_per_channel_scales().to(torch.float64).contiguous().numpy().tobytes(),
        tensor.q_per_channel_zero_points().to(torch.int64).contiguous().numpy().tobytes(),
    ))


def per_channel_qtensor_from_bytes(buf: bytes, dtype: torch.dtype, shape: List[int]) -> torch.Tensor:
    nbytes = reduce(mul, shape, 1) * dtype_to_element_size(dtype)
    (axis,) = struct.unpack("q", bytes(buf[:8]))
    if axis < 0 or axis >= len(shape):
        raise RuntimeError(
            f"Read invalid axis ({axis}) from the input buffer when deserializing the per-channel quantized "
            f"tensor (dtype: {dtype}, shape: {shape})."
        )
    channels = shape[axis]
    if len(buf) != 8 + nbytes + 16 * channels:
        raise RuntimeError(
            f"The expected buffer size for the per-channel quantized tensor (dtype: {dtype}, shape: {shape}, "
            f"axis: {axis}) is {8 + nbytes + 16 * channels}. The size of the input buffer is {len(buf)}."
        )
    body = 8 + nbytes
    scales = torch.frombuffer(bytearray(buf[body : body + 8 * channels]), dtype=torch.float64)
    zero_points = torch.frombuffer(bytearray(buf[body + 8 * channels :]), dtype=torch.int64)
    return torch._make_per_channel_quantized_tensor(_ints_from_bytes(buf[8:body], dtype, shape), scales, zero_points, axis)
