// _tsb200_meta: the geometry of many torch tensors in one call.
//
// Planning a take (or a restore) of a state dict with 10-20 k tensors is host work per tensor: address, device,
// dtype, sizes, strides, contiguity -- five or six attribute reads each from Python, about 0.7 us per tensor and
// attribute.  scan() reads them all from the TensorImpl for a whole list and hands back one int64 table, which
// the planners (batcher.plan_cuda_tensors / plan_cuda_reads) and the plan caches work on with numpy.  It touches
// no tensor data and launches nothing; the copies themselves go through libtsb200.so's C-ABI (include/tsb200.h),
// which knows nothing of torch.  This file is the one place where the package reads torch objects from C++.
//
// The reference does this walk in Python, tensor by tensor (torchsnapshot/io_preparer.py:71-147,
// io_preparers/tensor.py:57-112, batcher.py:262-316).
//
// Row layout (TSB_META_WIDTH int64 columns), mirrored by tensor_table.py:
//   0 kind   0 = not a tensor, 1 = plain dense tensor described by the rest of the row,
//            2 = torch.Tensor / Parameter this table does not describe (sparse, quantised, conj/neg view,
//                more than 8 dims, no storage, neither CPU nor CUDA), 3 = instance of a Tensor subclass
//   1 data_ptr   2 device index (-1 = CPU)   3 c10::ScalarType code   4 element size   5 numel
//   6 ndim       7 is_contiguous            8..15 sizes               16..23 strides (elements)
#include <torch/csrc/python_headers.h>

#include <ATen/ATen.h>
#include <torch/csrc/autograd/python_variable.h>

#include <cstdint>
#include <cstring>

namespace {

constexpr int kWidth = 24;
constexpr int kMaxDims = 8;

void describe(PyObject* obj, int64_t* row) {
  if (!THPVariable_CheckExact(obj)) {
    int is_tensor = THPVariableClass ? PyObject_IsInstance(obj, THPVariableClass) : 0;
    if (is_tensor < 0) {
      PyErr_Clear();
      is_tensor = 0;
    }
    row[0] = is_tensor ? 3 : 0;
    return;
  }
  const at::Tensor& t = THPVariable_Unpack(obj);
  row[0] = 2;
  if (!t.defined() || t.layout() != at::kStrided || t.is_quantized() || t.is_nested() || t.is_conj() || t.is_neg() ||
      t.dim() > kMaxDims || !(t.is_cpu() || t.is_cuda()) || !t.has_storage()) {
    return;
  }
  const void* ptr = nullptr;
  try {
    ptr = t.const_data_ptr();
  } catch (...) {  // fake / functional tensors have no address
    return;
  }
  row[0] = 1;
  row[1] = static_cast<int64_t>(reinterpret_cast<uintptr_t>(ptr));
  row[2] = t.is_cuda() ? static_cast<int64_t>(t.get_device()) : -1;
  row[3] = static_cast<int64_t>(t.scalar_type());
  row[4] = static_cast<int64_t>(t.element_size());
  row[5] = t.numel();
  const int64_t nd = t.dim();
  row[6] = nd;
  row[7] = t.is_contiguous() ? 1 : 0;
  const auto sizes = t.sizes();
  const auto strides = t.strides();
  for (int64_t d = 0; d < nd; ++d) {
    row[8 + d] = sizes[d];
    row[16 + d] = strides[d];
  }
}

PyObject* scan(PyObject* /*self*/, PyObject* arg) {
  PyObject* seq = PySequence_Fast(arg, "scan() takes a sequence of objects");
  if (seq == nullptr) {
    return nullptr;
  }
  const Py_ssize_t n = PySequence_Fast_GET_SIZE(seq);
  PyObject* out = PyByteArray_FromStringAndSize(nullptr, n * kWidth * static_cast<Py_ssize_t>(sizeof(int64_t)));
  if (out == nullptr) {
    Py_DECREF(seq);
    return nullptr;
  }
  int64_t* rows = reinterpret_cast<int64_t*>(PyByteArray_AS_STRING(out));
  if (n > 0) {
    std::memset(rows, 0, static_cast<size_t>(n) * kWidth * sizeof(int64_t));
  }
  PyObject** items = PySequence_Fast_ITEMS(seq);
  for (Py_ssize_t i = 0; i < n; ++i) {
    try {
      describe(items[i], rows + i * kWidth);
    } catch (...) {
      std::memset(rows + i * kWidth, 0, kWidth * sizeof(int64_t));
      rows[i * kWidth] = 2;
    }
  }
  Py_DECREF(seq);
  return out;
}

PyMethodDef methods[] = {
    {"scan", scan, METH_O,
     "scan(objects) -> bytearray holding len(objects) rows of 24 int64 (see tensor_table.py for the columns)"},
    {nullptr, nullptr, 0, nullptr},
};

PyModuleDef module = {PyModuleDef_HEAD_INIT, "_tsb200_meta", "tensor geometry tables for the checkpoint planners", -1,
                      methods, nullptr, nullptr, nullptr, nullptr};

}  // namespace

PyMODINIT_FUNC PyInit__tsb200_meta(void) {
  PyObject* m = PyModule_Create(&module);
  if (m != nullptr) {
    PyModule_AddIntConstant(m, "WIDTH", kWidth);
    PyModule_AddIntConstant(m, "MAX_DIMS", kMaxDims);
  }
  return m;
}
