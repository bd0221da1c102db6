"""Geometry of many tensors as one int64 table (one row per object), for the planners and the plan caches.

`scan(objects)` returns an array of shape [len(objects), WIDTH]; the columns are named below.  The rows come
from the in-tree CPython module csrc/_tsb200_meta.so (csrc/tensor_table.cpp: one pass over the list in C++),
or, when that module is not there (TSB200_TENSOR_TABLE=python, or the build is missing: logged once), from
`scan_py`, which reads the same attributes from Python and gives the same rows (tests/test_tensor_table.py).
Host-side bookkeeping only: no tensor data is touched here and nothing is launched.

What the reference does tensor by tensor in Python (io_preparer.py:71-147, io_preparers/tensor.py:57-112,
batcher.py:262-316) the planners here do on these columns with numpy.
"""

import importlib.util
import logging
import os
from typing import Any, Dict, Optional, Sequence

import numpy as np
import torch

logger = logging.getLogger(__name__)

WIDTH = 24
MAX_DIMS = 8
KIND, PTR, DEVICE, DTYPE, ITEMSIZE, NUMEL, NDIM, CONTIGUOUS = range(8)
SIZES, STRIDES = 8, 16

NOT_A_TENSOR, PLAIN, UNDESCRIBED, SUBCLASS = 0, 1, 2, 3

# c10::ScalarType codes of the dtypes a snapshot can hold (c10/core/ScalarType.h); checked against the native
# module when it loads
_CODES: Dict[torch.dtype, int] = {
    torch.uint8: 0, torch.int8: 1, torch.int16: 2, torch.int32: 3, torch.int64: 4, torch.float16: 5,
    torch.float32: 6, torch.float64: 7, torch.complex64: 9, torch.complex128: 10, torch.bool: 11,
    torch.bfloat16: 15,
}
for _name, _code in (("complex32", 8), ("float8_e5m2", 23), ("float8_e4m3fn", 24), ("float8_e5m2fnuz", 25),
                     ("float8_e4m3fnuz", 26), ("uint16", 27), ("uint32", 28), ("uint64", 29)):
    if hasattr(torch, _name):
        _CODES[getattr(torch, _name)] = _code
_DTYPES: Dict[int, torch.dtype] = {c: d for d, c in _CODES.items()}
_KNOWN = np.zeros(64, dtype=np.bool_)  # indexed by code
_KNOWN[list(_DTYPES)] = True


def dtype_code(dtype: torch.dtype) -> int:
    """The table's code of `dtype`, or -1 for a dtype the table does not name (quantised ones)."""
    return _CODES.get(dtype, -1)


def dtype_of(code: int) -> torch.dtype:
    return _DTYPES[int(code)]


def scan_py(objects: Sequence[Any]) -> np.ndarray:
    """The table, read attribute by attribute from Python (the fallback, and the check of the native module)."""
    plain, param, tensor = torch.Tensor, torch.nn.Parameter, torch.Tensor
    out = np.zeros((len(objects), WIDTH), dtype=np.int64)
    for i, o in enumerate(objects):
        kind = type(o)
        if kind is not plain and kind is not param:
            out[i, KIND] = SUBCLASS if isinstance(o, tensor) else NOT_A_TENSOR
            continue
        out[i, KIND] = UNDESCRIBED
        try:
            if (o.layout is not torch.strided or o.is_quantized or o.is_nested or o.is_conj() or o.is_neg()
                    or o.dim() > MAX_DIMS or o.device.type not in ("cpu", "cuda") or o.dtype not in _CODES):
                continue
            ptr = o.data_ptr()
        except Exception:  # fake / functional tensors have no address
            continue
        nd = o.dim()
        row = out[i]
        row[KIND] = PLAIN
        row[PTR] = np.uint64(ptr).astype(np.int64)
        row[DEVICE] = o.get_device() if o.is_cuda else -1
        row[DTYPE] = _CODES[o.dtype]
        row[ITEMSIZE] = o.element_size()
        row[NUMEL] = o.numel()
        row[NDIM] = nd
        row[CONTIGUOUS] = 1 if o.is_contiguous() else 0
        row[SIZES:SIZES + nd] = o.size()
        row[STRIDES:STRIDES + nd] = o.stride()
    return out


_native_scan: Optional[Any] = None
_tried = False


def _load_native() -> Optional[Any]:
    global _native_scan, _tried
    if _tried:
        return _native_scan
    _tried = True
    if os.environ.get("TSB200_TENSOR_TABLE", "").lower() == "python":
        return None
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "_tsb200_meta.so")
    try:
        spec = importlib.util.spec_from_file_location("_tsb200_meta", path)
        if spec is None or spec.loader is None or not os.path.exists(path):
            raise ImportError(f"{path} has not been built (make -C csrc _tsb200_meta.so)")
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        if mod.WIDTH != WIDTH or mod.MAX_DIMS != MAX_DIMS:
            raise ImportError(f"{path}: row layout {mod.WIDTH}/{mod.MAX_DIMS}, this package expects {WIDTH}/{MAX_DIMS}")
        # the codes above are the module's: one empty tensor per dtype, both ways
        import warnings

        with warnings.catch_warnings():
            warnings.simplefilter("ignore")  # "ComplexHalf support is experimental"
            probes = [torch.empty(0, dtype=d) for d in _CODES]
        got = np.frombuffer(mod.scan(probes), dtype=np.int64).reshape(len(probes), WIDTH)
        if got[:, DTYPE].tolist() != list(_CODES.values()) or not np.array_equal(got, scan_py(probes)):
            raise ImportError(f"{path}: dtype codes differ from this package's table")
        _native_scan = mod.scan
    except Exception as e:  # planners fall back to reading the attributes from Python: slower, same tables
        logger.warning(f"tensor tables come from Python, not from csrc/_tsb200_meta.so: {e}")
        _native_scan = None
    return _native_scan


def native_loaded() -> bool:
    return _load_native() is not None


def scan(objects: Sequence[Any]) -> np.ndarray:
    """[len(objects), WIDTH] int64 rows; see the column names at the top of this module."""
    native = _load_native()
    if native is None:
        return scan_py(objects)
    if not isinstance(objects, (list, tuple)):
        objects = list(objects)
    table = np.frombuffer(native(objects), dtype=np.int64).reshape(len(objects), WIDTH)
    unnamed = (table[:, KIND] == PLAIN) & ~_KNOWN[np.clip(table[:, DTYPE], 0, 63)]
    if unnamed.any():  # a dtype without a code here (sub-byte ints, float4 ...): left to the per-leaf code
        table[unnamed] = 0
        table[unnamed, KIND] = UNDESCRIBED
    return table


def dtype_mask(dtypes: Any) -> np.ndarray:
    """Boolean look-up table over dtype codes: True for the codes of `dtypes` (index it with a DTYPE column)."""
    lut = np.zeros(64, dtype=np.bool_)
    lut[[_CODES[d] for d in dtypes if d in _CODES]] = True
    return lut


def shapes_of(table: np.ndarray) -> Any:
    """Per row, the sizes as a list of ints (what TensorEntry.shape holds)."""
    return [s[:nd] for s, nd in zip(table[:, SIZES:SIZES + MAX_DIMS].tolist(), table[:, NDIM].tolist())]


def nbytes_of(table: np.ndarray) -> np.ndarray:
    return table[:, NUMEL] * table[:, ITEMSIZE]


def ptrs_of(table: np.ndarray) -> np.ndarray:
    return table[:, PTR].astype(np.uint64)
