/*
 * slab_oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C restatement of the byte movement on torchsnapshot's checkpoint hot path, used only by
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs as the
 * checker the CUDA path is compared with.  Nothing under torchsnapshot-1_b200/ links or calls it.
 *
 * Pinned against the reference by tests/test_oracle_golden.py: fixtures produced by running the
 * unmodified reference (oracle/gen_golden.py, outputs in tests/golden/).
 *
 * Citations are file:line in the reference repository.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define ORC_MAX_DIMS 16

/*
 * Serialised payload of a strided tensor: elements in row-major order of the logical shape,
 * native byte order, no header, no padding.
 *   torchsnapshot/serialization.py:162-188  tensor_as_memoryview: `.contiguous()` (182-185) then
 *                                           memoryview(tensor.numpy()).cast("b")
 *   torchsnapshot/serialization.py:191-212  bf16 goes through untyped storage: same bytes
 * `strides` are in elements (as torch reports them), `src` points at element [0,...,0].
 * Returns the number of bytes written.
 */
size_t orc_serialize(uint8_t* dst, const uint8_t* src, size_t elsize, int ndim, const int64_t* sizes,
                     const int64_t* strides) {
  int64_t idx[ORC_MAX_DIMS] = {0};
  size_t numel = 1;
  for (int d = 0; d < ndim; ++d) {
    if (sizes[d] == 0) return 0;
    numel *= (size_t)sizes[d];
  }
  size_t out = 0;
  for (size_t n = 0; n < numel; ++n) {
    int64_t off = 0;
    for (int d = 0; d < ndim; ++d) off += idx[d] * strides[d];
    memcpy(dst + out, src + off * (int64_t)elsize, elsize);
    out += elsize;
    for (int d = ndim - 1; d >= 0; --d) {
      if (++idx[d] < sizes[d]) break;
      idx[d] = 0;
    }
  }
  return out;
}

/*
 * The inverse: payload bytes -> a strided destination tensor of the same dtype.
 *   torchsnapshot/io_preparers/tensor.py:317-338  deserialize_tensor (frombuffer + reshape) then
 *                                                 tensor_copy -> dst.detach().copy_(src) (356-358)
 */
size_t orc_deserialize(uint8_t* dst, const uint8_t* src, size_t elsize, int ndim, const int64_t* sizes,
                       const int64_t* strides) {
  int64_t idx[ORC_MAX_DIMS] = {0};
  size_t numel = 1;
  for (int d = 0; d < ndim; ++d) {
    if (sizes[d] == 0) return 0;
    numel *= (size_t)sizes[d];
  }
  size_t in = 0;
  for (size_t n = 0; n < numel; ++n) {
    int64_t off = 0;
    for (int d = 0; d < ndim; ++d) off += idx[d] * strides[d];
    memcpy(dst + off * (int64_t)elsize, src + in, elsize);
    in += elsize;
    for (int d = ndim - 1; d >= 0; --d) {
      if (++idx[d] < sizes[d]) break;
      idx[d] = 0;
    }
  }
  return in;
}

/*
 * Slab fill: member buffer -> slab[lo:hi].
 *   torchsnapshot/batcher.py:63-90    BatchedBufferStager.stage_buffer: slab[byte_range] = buf
 *   torchsnapshot/batcher.py:151-155  GPUBatchedBufferStager: buf_storage[lo:hi].copy_(tensor_storage)
 */
void orc_slab_put(uint8_t* slab, uint64_t lo, const uint8_t* buf, uint64_t nbytes) { memcpy(slab + lo, buf, nbytes); }

/*
 * Reshard copy of one overlapping region: narrow() both tensors per dim, then copy_.
 *   torchsnapshot/io_preparers/sharded_tensor.py:271-284  _OverlappingRegion.get_views
 *   torchsnapshot/io_preparers/sharded_tensor.py:296-309  ShardedTensorBufferConsumer.consume_buffer
 * `src` is the persisted piece (dense, row-major, shape src_sizes); `dst` the local shard with
 * element strides dst_strides.  Per dim: src_off, dst_off, len.
 */
void orc_copy_box(uint8_t* dst, const int64_t* dst_strides, const uint8_t* src, const int64_t* src_sizes, size_t elsize,
                  int ndim, const int64_t* src_off, const int64_t* dst_off, const int64_t* len) {
  int64_t src_strides[ORC_MAX_DIMS];
  int64_t acc = 1;
  for (int d = ndim - 1; d >= 0; --d) {
    src_strides[d] = acc;
    acc *= src_sizes[d];
  }
  int64_t idx[ORC_MAX_DIMS] = {0};
  size_t numel = 1;
  for (int d = 0; d < ndim; ++d) {
    if (len[d] <= 0) return;
    numel *= (size_t)len[d];
  }
  for (size_t n = 0; n < numel; ++n) {
    int64_t so = 0, dofs = 0;
    for (int d = 0; d < ndim; ++d) {
      so += (src_off[d] + idx[d]) * src_strides[d];
      dofs += (dst_off[d] + idx[d]) * dst_strides[d];
    }
    memcpy(dst + dofs * (int64_t)elsize, src + so * (int64_t)elsize, elsize);
    for (int d = ndim - 1; d >= 0; --d) {
      if (++idx[d] < len[d]) break;
      idx[d] = 0;
    }
  }
}
