"""tensor_table.scan: the native module (csrc/_tsb200_meta.so) and the Python reader give the same rows, and the
rows say what the tensors say of themselves."""
import numpy as np
import pytest
import torch

import tsb200_loader

tsb200_loader.load()
from torchsnapshot_b200 import tensor_table as tt  # noqa: E402


class _Sub(torch.Tensor):
    pass


def _zoo(device="cpu"):
    base = torch.arange(6 * 8, dtype=torch.float32, device=device).reshape(6, 8)
    objs = [
        base, base[:, ::2], base.t(), base[1:4, 2:5], torch.nn.Parameter(base.clone()), torch.tensor(1.5, device=device),
        torch.empty(0, 3, device=device), torch.empty(0, device=device), torch.zeros([2] * 8, device=device),
        torch.zeros(3, 1, 4, device=device).expand(3, 5, 4),
        3, 2.5, "text", None, [1, 2], {"a": 1}, np.zeros(3),
        torch.zeros([1] * 9, device=device),  # too many dims for a row
        torch.zeros(2, dtype=torch.complex64, device=device).conj(),
        torch.zeros(2, device=device).as_subclass(_Sub),
    ]
    for dt in (torch.uint8, torch.int8, torch.int16, torch.int32, torch.int64, torch.float16, torch.bfloat16,
               torch.float64, torch.bool, torch.complex64, torch.complex128):
        objs.append(torch.zeros(5, dtype=dt, device=device))
    if device == "cpu":
        objs += [torch.eye(3).to_sparse(), torch.quantize_per_tensor(torch.randn(4), 0.1, 0, torch.qint8),
                 torch.empty(4, device="meta")]
    return objs


def test_native_module_is_built_and_loaded():
    assert tt.native_loaded(), "csrc/_tsb200_meta.so is missing: make -C torchsnapshot-1_b200/csrc _tsb200_meta.so"


def _check_rows(objs):
    table = tt.scan(objs)
    assert table.shape == (len(objs), tt.WIDTH) and table.dtype == np.int64
    assert np.array_equal(table, tt.scan_py(objs))
    for row, o in zip(table, objs):
        if not isinstance(o, torch.Tensor):
            assert row[tt.KIND] == tt.NOT_A_TENSOR and not row[1:].any()
        elif type(o) not in (torch.Tensor, torch.nn.Parameter):
            assert row[tt.KIND] == tt.SUBCLASS
        elif (o.layout is not torch.strided or o.is_quantized or o.dim() > tt.MAX_DIMS or o.is_conj()
              or o.device.type == "meta"):
            assert row[tt.KIND] == tt.UNDESCRIBED and not row[1:].any()
        else:
            assert row[tt.KIND] == tt.PLAIN
            assert int(np.uint64(o.data_ptr()).astype(np.int64)) == row[tt.PTR]
            assert row[tt.DEVICE] == (o.get_device() if o.is_cuda else -1)
            assert tt.dtype_of(row[tt.DTYPE]) == o.dtype and row[tt.ITEMSIZE] == o.element_size()
            assert row[tt.NUMEL] == o.numel() and row[tt.NDIM] == o.dim()
            assert bool(row[tt.CONTIGUOUS]) == o.is_contiguous()
            assert row[tt.SIZES:tt.SIZES + o.dim()].tolist() == list(o.shape)
            assert row[tt.STRIDES:tt.STRIDES + o.dim()].tolist() == list(o.stride())
    return table


def test_rows_on_cpu():
    objs = _zoo()
    table = _check_rows(objs)
    assert tt.shapes_of(table[:3]) == [[6, 8], [6, 4], [8, 6]]
    assert tt.nbytes_of(table[:2]).tolist() == [192, 96]
    # any sequence will do
    assert np.array_equal(tt.scan(tuple(objs)), table)
    assert np.array_equal(tt.scan(iter(objs)) if not tt.native_loaded() else tt.scan(list(iter(objs))), table)
    assert tt.scan([]).shape == (0, tt.WIDTH)


def test_dtype_mask_and_codes():
    mask = tt.dtype_mask([torch.float32, torch.bool])
    table = tt.scan([torch.zeros(1), torch.zeros(1, dtype=torch.bool), torch.zeros(1, dtype=torch.int64)])
    assert mask[table[:, tt.DTYPE]].tolist() == [True, True, False]
    assert tt.dtype_code(torch.bfloat16) == table.dtype.type(15)
    assert tt.dtype_code(torch.qint8) == -1


def test_a_moved_or_reshaped_tensor_changes_its_row():
    t = torch.zeros(4, 6)
    before = tt.scan([t])
    assert np.array_equal(before, tt.scan([t]))
    t.data = torch.zeros(4, 6)  # same shape, other memory
    moved = tt.scan([t])
    assert not np.array_equal(before, moved) and moved[0, tt.PTR] == t.data_ptr()
    t.resize_(3, 8)
    assert tt.scan([t])[0, tt.SIZES:tt.SIZES + 2].tolist() == [3, 8]
    v = t.t()
    assert tt.scan([v])[0, tt.CONTIGUOUS] == 0


@pytest.mark.gpu
def test_rows_on_cuda():
    table = _check_rows(_zoo("cuda"))
    assert (table[table[:, tt.KIND] == tt.PLAIN, tt.DEVICE] == torch.cuda.current_device()).all()
