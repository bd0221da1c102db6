// Multi-region strided copy kernel for sm_100a: the gather-pack / scatter-unpack / reshard
// hot op of the checkpoint path (replaces the per-tensor copy_ launches of the reference,
// torchsnapshot/batcher.py:152-155, io_preparers/tensor.py:329-338, sharded_tensor.py:296-309).
//
// Work decomposition: the host splits the flat payload of every region into tiles of at most
// TSB_TILE_BYTES; a persistent grid (SMs x resident CTAs) walks the tile table with a stride of
// gridDim.x.  Tile and region descriptors of the next iteration are prefetched into registers
// while the current tile streams, so the dependent descriptor->data chain never sits on the
// critical path.  All payload traffic is 16-byte vector loads/stores with the DESTINATION
// aligned; a source whose alignment differs is read with aligned vectors and re-aligned in
// registers with funnel shifts, so slab offsets may be arbitrary (the reference packs with no
// padding, batcher.py:300-316).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tsb200.h"

#define TSB_THREADS 256
#define TSB_CTAS_PER_SM 4
#define TSB_UNROLL 4
// A run at least this long is copied by the whole CTA; shorter runs go one per warp.
#define TSB_CTA_RUN_MIN (16u * 1024u)
// A run shorter than this is copied element-wise, one unit per thread.
#define TSB_WARP_RUN_MIN 64u

struct __align__(16) TsbTile {
  uint64_t off;     // flat payload byte offset inside the region
  uint32_t len;     // payload bytes in this tile
  uint32_t region;  // index into the region table
};

struct RegionHot {  // the fields every tile needs, prefetched into registers
  uint64_t src, dst, run_bytes;
  uint32_t ndim;
};

__device__ __forceinline__ TsbTile ld_tile(const TsbTile* p) {
  uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
  TsbTile t;
  t.off = (uint64_t)v.x | ((uint64_t)v.y << 32);
  t.len = v.z;
  t.region = v.w;
  return t;
}

__device__ __forceinline__ RegionHot ld_hot(const TsbRegion* r) {
  const uint4* p = reinterpret_cast<const uint4*>(r);
  uint4 a = __ldg(p);
  uint4 b = __ldg(p + 1);
  RegionHot h;
  h.src = (uint64_t)a.x | ((uint64_t)a.y << 32);
  h.dst = (uint64_t)a.z | ((uint64_t)a.w << 32);
  h.run_bytes = (uint64_t)b.x | ((uint64_t)b.y << 32);
  h.ndim = b.z;
  return h;
}

// streaming 16-byte accessors: payload is touched exactly once
__device__ __forceinline__ uint4 ld_stream(const uint4* p) { return __ldcs(p); }
__device__ __forceinline__ void st_stream(uint4* p, const uint4& v) { __stcs(p, v); }

// out = bytes [S, S+16) of the 32-byte little-endian stream (a, b); SW = S / 4, sb = 8 * (S % 4)
template <int SW>
__device__ __forceinline__ uint4 realign(const uint4& a, const uint4& b, uint32_t sb) {
  const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  uint4 o;
  o.x = __funnelshift_r(w[SW + 0], w[SW + 1], sb);
  o.y = __funnelshift_r(w[SW + 1], w[SW + 2], sb);
  o.z = __funnelshift_r(w[SW + 2], w[SW + 3], sb);
  o.w = __funnelshift_r(w[SW + 3], w[SW + 4], sb);
  return o;
}

template <int SW, int NL>
__device__ __forceinline__ void copy_vec_shifted(uint4* __restrict__ d4, const uint4* __restrict__ a4,
                                                 uint64_t nvec, uint32_t sb, uint32_t lane) {
  uint64_t j = lane;
  // a4[j] and a4[j+1] both lie in aligned 16-byte granules that contain payload bytes, so the
  // extra vector never leaves the allocation that holds the run.
  for (; j + (uint64_t)(TSB_UNROLL - 1) * NL < nvec; j += (uint64_t)TSB_UNROLL * NL) {
    uint4 a[TSB_UNROLL], b[TSB_UNROLL];
#pragma unroll
    for (int u = 0; u < TSB_UNROLL; ++u) {
      a[u] = __ldg(a4 + j + (uint64_t)u * NL);
      b[u] = __ldg(a4 + j + (uint64_t)u * NL + 1);
    }
#pragma unroll
    for (int u = 0; u < TSB_UNROLL; ++u) st_stream(d4 + j + (uint64_t)u * NL, realign<SW>(a[u], b[u], sb));
  }
  for (; j < nvec; j += NL) {
    uint4 a = __ldg(a4 + j);
    uint4 b = __ldg(a4 + j + 1);
    st_stream(d4 + j, realign<SW>(a, b, sb));
  }
}

// Cooperative byte-exact copy of n bytes by NL threads (lane in [0, NL)); NL >= 32.
template <int NL>
__device__ __forceinline__ void copy_span(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint64_t n,
                                          uint32_t lane) {
  if (n == 0) return;
  uint32_t head = (uint32_t)((16u - ((uint32_t)(uintptr_t)dst & 15u)) & 15u);
  if ((uint64_t)head > n) head = (uint32_t)n;
  if (lane < head) dst[lane] = src[lane];
  dst += head;
  src += head;
  n -= head;
  const uint64_t nvec = n >> 4;
  const uint32_t tail = (uint32_t)(n & 15u);
  // tail bytes by the last lanes, so head and tail do not serialise on the same threads
  if (lane >= NL - tail) {
    const uint64_t k = (nvec << 4) + (lane - (NL - tail));
    dst[k] = src[k];
  }
  if (nvec == 0) return;
  uint4* d4 = reinterpret_cast<uint4*>(dst);
  const uint32_t s = (uint32_t)(uintptr_t)src & 15u;
  if (s == 0) {
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    uint64_t j = lane;
    for (; j + (uint64_t)(TSB_UNROLL - 1) * NL < nvec; j += (uint64_t)TSB_UNROLL * NL) {
      uint4 v[TSB_UNROLL];
#pragma unroll
      for (int u = 0; u < TSB_UNROLL; ++u) v[u] = ld_stream(s4 + j + (uint64_t)u * NL);
#pragma unroll
      for (int u = 0; u < TSB_UNROLL; ++u) st_stream(d4 + j + (uint64_t)u * NL, v[u]);
    }
    for (; j < nvec; j += NL) st_stream(d4 + j, ld_stream(s4 + j));
  } else {
    const uint4* a4 = reinterpret_cast<const uint4*>(src - s);
    const uint32_t sb = (s & 3u) * 8u;
    switch (s >> 2) {
      case 0: copy_vec_shifted<0, NL>(d4, a4, nvec, sb, lane); break;
      case 1: copy_vec_shifted<1, NL>(d4, a4, nvec, sb, lane); break;
      case 2: copy_vec_shifted<2, NL>(d4, a4, nvec, sb, lane); break;
      default: copy_vec_shifted<3, NL>(d4, a4, nvec, sb, lane); break;
    }
  }
}

// Byte offsets of run q of a strided region on the source and destination side.
__device__ __forceinline__ void run_offsets(const TsbRegion* __restrict__ r, uint32_t ndim, uint64_t q, int64_t& so,
                                            int64_t& dofs) {
  so = 0;
  dofs = 0;
  for (int d = (int)ndim - 1; d >= 0; --d) {
    const uint64_t sz = __ldg(&r->size[d]);
    uint64_t c;
    if ((q | sz) >> 32) {
      c = q % sz;
      q = q / sz;
    } else {
      c = (uint32_t)q % (uint32_t)sz;
      q = (uint32_t)q / (uint32_t)sz;
    }
    so += (int64_t)c * __ldg(&r->src_stride[d]);
    dofs += (int64_t)c * __ldg(&r->dst_stride[d]);
  }
}

template <typename T>
__device__ __forceinline__ void copy_units(const TsbRegion* __restrict__ r, const RegionHot& h, uint64_t first_unit,
                                           uint64_t nunits, uint32_t units_per_run, uint32_t tid) {
  for (uint64_t u = tid; u < nunits; u += TSB_THREADS) {
    const uint64_t f = first_unit + u;
    uint64_t q, k;
    if ((f >> 32) == 0) {
      q = (uint32_t)f / units_per_run;
      k = (uint32_t)f % units_per_run;
    } else {
      q = f / units_per_run;
      k = f % units_per_run;
    }
    int64_t so, dofs;
    run_offsets(r, h.ndim, q, so, dofs);
    const T* sp = reinterpret_cast<const T*>(h.src + so) + k;
    T* dp = reinterpret_cast<T*>(h.dst + dofs) + k;
    *dp = *sp;
  }
}

__device__ __forceinline__ void process_tile(const TsbTile& t, const RegionHot& h, const TsbRegion* __restrict__ r,
                                             uint32_t tid) {
  if (h.ndim == 0) {
    copy_span<TSB_THREADS>(reinterpret_cast<uint8_t*>(h.dst + t.off), reinterpret_cast<const uint8_t*>(h.src + t.off),
                           t.len, tid);
    return;
  }
  if (h.run_bytes >= TSB_CTA_RUN_MIN) {
    uint64_t q = t.off / h.run_bytes;
    uint64_t inner = t.off - q * h.run_bytes;
    uint64_t remaining = t.len;
    while (remaining) {
      uint64_t piece = h.run_bytes - inner;
      if (piece > remaining) piece = remaining;
      int64_t so, dofs;
      run_offsets(r, h.ndim, q, so, dofs);
      copy_span<TSB_THREADS>(reinterpret_cast<uint8_t*>(h.dst + dofs + inner),
                             reinterpret_cast<const uint8_t*>(h.src + so + inner), piece, tid);
      remaining -= piece;
      inner = 0;
      ++q;
    }
    return;
  }
  if (h.run_bytes >= TSB_WARP_RUN_MIN) {
    // the host emits whole runs per tile for this class
    const uint64_t q0 = t.off / h.run_bytes;
    const uint32_t nruns = t.len / (uint32_t)h.run_bytes;
    const uint32_t warp = tid >> 5, lane = tid & 31u;
    for (uint32_t i = warp; i < nruns; i += TSB_THREADS / 32) {
      int64_t so, dofs;
      run_offsets(r, h.ndim, q0 + i, so, dofs);
      copy_span<32>(reinterpret_cast<uint8_t*>(h.dst + dofs), reinterpret_cast<const uint8_t*>(h.src + so),
                    h.run_bytes, lane);
    }
    return;
  }
  // element mode: the widest power-of-two unit every address of the region is aligned to
  uint64_t m = h.src | h.dst | h.run_bytes | 16u;
  for (uint32_t d = 0; d < h.ndim; ++d)
    m |= (uint64_t)__ldg(&r->src_stride[d]) | (uint64_t)__ldg(&r->dst_stride[d]);
  const uint32_t unit = (uint32_t)(m & (~m + 1));  // lowest set bit: 1, 2, 4, 8 or 16
  const uint32_t upr = (uint32_t)h.run_bytes / unit;
  const uint64_t first = t.off / unit;
  const uint64_t nunits = t.len / unit;
  switch (unit) {
    case 1: copy_units<uint8_t>(r, h, first, nunits, upr, tid); break;
    case 2: copy_units<uint16_t>(r, h, first, nunits, upr, tid); break;
    case 4: copy_units<uint32_t>(r, h, first, nunits, upr, tid); break;
    case 8: copy_units<uint2>(r, h, first, nunits, upr, tid); break;
    default: copy_units<uint4>(r, h, first, nunits, upr, tid); break;
  }
}

__global__ void __launch_bounds__(TSB_THREADS, TSB_CTAS_PER_SM)
tsb_copy_regions_kernel(const TsbRegion* __restrict__ regions, const TsbTile* __restrict__ tiles, uint32_t n_tiles) {
  const uint32_t tid = threadIdx.x;
  const uint64_t stride = gridDim.x;
  uint64_t t = blockIdx.x;
  if (t >= n_tiles) return;
  TsbTile cur = ld_tile(tiles + t);
  RegionHot hot = ld_hot(regions + cur.region);
  bool has_next = t + stride < n_tiles;
  TsbTile nxt = cur;
  if (has_next) nxt = ld_tile(tiles + t + stride);
  for (;;) {
    // prefetch: tile entry two iterations ahead, region header one iteration ahead
    const bool has_nn = has_next && (t + 2 * stride < n_tiles);
    TsbTile nn = nxt;
    RegionHot nhot = hot;
    if (has_nn) nn = ld_tile(tiles + t + 2 * stride);
    if (has_next) nhot = ld_hot(regions + nxt.region);
    process_tile(cur, hot, regions + cur.region, tid);
    if (!has_next) break;
    cur = nxt;
    hot = nhot;
    nxt = nn;
    has_next = has_nn;
    t += stride;
  }
}
