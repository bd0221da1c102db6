"""TEST INFRASTRUCTURE -- the CPU oracle for the checkpoint hot path.

`oracle/` restates, in plain C (slab_oracle.c) and pure-Python integer math (planner.py), what the
reference's hot path computes, each function citing the reference file:line it follows.  It exists
so the CUDA path can be checked bit-for-bit on a GPU box where /root/reference is absent.

Only tests/, __graft_entry__.smoke() and bench.py's `cpu_baseline` / `--impl reference` legs may
import it.  The product (torchsnapshot-1_b200/) never does.

Parity pinning: tests/test_oracle_golden.py checks this oracle against fixtures written by the
unmodified reference (oracle/gen_golden.py -> tests/golden/) and against the golden vectors in the
reference's own tests.  The reference is pure Python, so there is nothing to compile into
oracle/_ref; the timed CPU baseline is this port (`kind: "port"`).
"""

import ctypes
import os
import subprocess
from typing import List, Sequence, Tuple

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None


def build() -> None:
    subprocess.run(["make", "-C", _HERE, "liboracle.so"], check=True, capture_output=True)


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = ctypes.CDLL(_LIB_PATH)
        i64p = ctypes.POINTER(ctypes.c_int64)
        L.orc_serialize.restype = ctypes.c_size_t
        L.orc_serialize.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, i64p, i64p]
        L.orc_deserialize.restype = ctypes.c_size_t
        L.orc_deserialize.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, i64p, i64p]
        L.orc_slab_put.restype = None
        L.orc_slab_put.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint64]
        L.orc_copy_box.restype = None
        L.orc_copy_box.argtypes = [ctypes.c_void_p, i64p, ctypes.c_void_p, i64p, ctypes.c_size_t, ctypes.c_int,
                                   i64p, i64p, i64p]
        _lib = L
    return _lib


def _i64(vals: Sequence[int]):
    return (ctypes.c_int64 * max(len(vals), 1))(*vals)


def serialize(t: torch.Tensor) -> bytes:
    """Payload bytes of a CPU tensor (reference serialization.py:162-188)."""
    assert t.device.type == "cpu"
    t = t.detach()
    n = t.numel() * t.element_size()
    out = ctypes.create_string_buffer(max(n, 1))
    if n:
        wrote = lib().orc_serialize(out, t.data_ptr(), t.element_size(), t.ndim, _i64(list(t.shape)),
                                    _i64(list(t.stride())))
        assert wrote == n
    return out.raw[:n]


def deserialize_into(t_out: torch.Tensor, payload: bytes) -> None:
    """payload -> CPU tensor of the same dtype (reference tensor.py:317-338, 356-358)."""
    assert t_out.device.type == "cpu"
    n = t_out.numel() * t_out.element_size()
    assert len(payload) == n, (len(payload), n)
    if n:
        lib().orc_deserialize(t_out.data_ptr(), payload, t_out.element_size(), t_out.ndim, _i64(list(t_out.shape)),
                              _i64(list(t_out.stride())))


def pack_slab(members: Sequence[Tuple[torch.Tensor, int, int]], size: int) -> bytes:
    """Slab bytes from (tensor, lo, hi) members (reference batcher.py:63-90 / 151-155)."""
    slab = ctypes.create_string_buffer(max(size, 1))
    for t, lo, hi in members:
        payload = serialize(t)
        assert len(payload) == hi - lo
        lib().orc_slab_put(slab, lo, payload, len(payload))
    return slab.raw[:size]


def copy_box(dst: torch.Tensor, piece_payload: bytes, piece_sizes: Sequence[int],
             region: Sequence[Tuple[int, int, int, int]]) -> None:
    """One overlapping region of a persisted piece -> local shard (reference sharded_tensor.py:271-309)."""
    assert dst.device.type == "cpu"
    nd = len(piece_sizes)
    src_off = [0] * nd
    dst_off = [0] * nd
    length = list(piece_sizes)
    for dim, so, do, ln in region:
        src_off[dim], dst_off[dim], length[dim] = so, do, ln
    lib().orc_copy_box(dst.data_ptr(), _i64(list(dst.stride())), piece_payload, _i64(list(piece_sizes)),
                       dst.element_size(), nd, _i64(src_off), _i64(dst_off), _i64(length))
