// Multi-region strided copy kernels for sm_100a: the gather-pack / scatter-unpack / reshard
// hot op of the checkpoint path (replaces the per-tensor copy_ launches of the reference,
// torchsnapshot/batcher.py:152-155, io_preparers/tensor.py:329-338, sharded_tensor.py:296-309).
//
// Work decomposition (measured on B200, scripts/copy_variants.cu, profiles/r01_copy_variants.txt):
// the host splits the flat payload of every region into tiles of at most TSB_TILE_BYTES and
// launches ONE CTA PER TILE in table order, so the CTAs resident at any instant sweep a dense
// window of the source and of the slab (6.9 TB/s r+w; a persistent grid that strides over the table
// reaches only 6.0 TB/s, and cudaMemcpy D2D 6.55 TB/s).  A tile entry carries absolute source and
// destination addresses, so a CTA has exactly one dependent descriptor load before it streams.
//
// All payload traffic is 16-byte vector loads/stores with the DESTINATION aligned; a source whose
// alignment differs is read with aligned vectors and re-aligned in registers with funnel shifts, so
// slab offsets may be arbitrary (the reference packs with no padding, batcher.py:300-316).
//
// Two kernels share the tile table format:
//   tsb_copy_flat_kernel     contiguous runs (dense tensors, the >99% case): every thread issues
//                            all its loads before its first store, no loop, <= 32 registers so
//                            2048 threads stay resident per SM;
//   tsb_copy_strided_kernel  strided boxes with short runs (non-contiguous views, reshard
//                            overlaps): one run per warp, or one element per thread when runs are
//                            shorter than 64 bytes.  Boxes with runs of 8 KB or more are expanded
//                            into flat tiles by the host and go through the flat kernel.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tsb200.h"

#define TSB_FLAT_THREADS 512
#define TSB_FLAT_UNROLL 4
// payload bytes per tile: every thread of the flat kernel moves TSB_FLAT_UNROLL 16-byte vectors
#define TSB_TILE_BYTES (TSB_FLAT_THREADS * TSB_FLAT_UNROLL * 16)

#define TSB_STRIDED_THREADS 256
#define TSB_UNROLL 4
// Strided regions whose runs are at least this long are expanded by the host into flat tiles (one
// or more per run); shorter runs go to the strided kernel, one per warp.
#define TSB_FLAT_RUN_MIN (8u * 1024u)
// Misaligned runs shorter than this many UNITS (the widest power of two every address of the region is
// aligned to) are copied unit by unit, longer ones one run per warp with realignment.
#define TSB_WARP_RUN_MIN 128u

// One unit of work.  Flat tiles: src/dst are the absolute addresses of the tile's first byte.
// Strided tiles: `region` indexes the region table and `off` is the flat payload offset.
struct __align__(32) TsbTile {
  uint64_t src;
  uint64_t dst;
  uint32_t len;
  uint32_t region;
  uint64_t off;
};

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

// 16-byte load that asks L2 for the 64-byte half line only (see ld_unit below): for the short runs of the strided
// kernel, where a miss promoted to the whole 128-byte line doubles the DRAM reads.
template <bool HINT>
__device__ __forceinline__ uint4 ld_vec16(const uint4* p) {
  if (!HINT) return __ldg(p);
  uint4 v;
  asm volatile("ld.global.L2::64B.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// ---------------------------------------------------------------------------------------------
// tile-table expansion (device side)
// ---------------------------------------------------------------------------------------------
// One contiguous run to copy; `first_tile` is the index of its first tile in the flat tile table
// (an exclusive prefix sum the host computes), so the table can be expanded on the device and the
// host uploads O(#runs) bytes instead of O(payload / tile) bytes.
struct __align__(32) TsbRun {
  uint64_t src;
  uint64_t dst;
  uint64_t nbytes;
  uint64_t first_tile;
};

// Tiles of a run: the first one ends on a 16-byte boundary of the destination, the others are
// TSB_TILE_BYTES long, so only the first tile has a head and only the last one a tail.
__host__ __device__ __forceinline__ uint64_t tsb_first_tile_len(uint64_t dst) { return TSB_TILE_BYTES - (dst & 15u); }
__host__ __device__ __forceinline__ uint64_t tsb_run_tiles(uint64_t dst, uint64_t nbytes) {
  const uint64_t l0 = tsb_first_tile_len(dst);
  if (nbytes <= l0) return nbytes ? 1 : 0;
  return 1 + (nbytes - l0 + TSB_TILE_BYTES - 1) / TSB_TILE_BYTES;
}

__global__ void __launch_bounds__(256)
tsb_expand_flat_kernel(const TsbRun* __restrict__ runs, uint32_t n_runs, TsbTile* __restrict__ tiles,
                       uint32_t n_tiles) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tiles) return;
  // largest r with runs[r].first_tile <= t
  uint32_t lo = 0, hi = n_runs - 1;
  while (lo < hi) {
    const uint32_t mid = (lo + hi + 1) >> 1;
    if (__ldg(&runs[mid].first_tile) <= t) lo = mid; else hi = mid - 1;
  }
  const uint4* rp = reinterpret_cast<const uint4*>(runs + lo);
  const uint4 a = __ldg(rp), b = __ldg(rp + 1);
  const uint64_t src = ((uint64_t)a.y << 32) | a.x, dst = ((uint64_t)a.w << 32) | a.z;
  const uint64_t nbytes = ((uint64_t)b.y << 32) | b.x, first = ((uint64_t)b.w << 32) | b.z;
  const uint64_t k = t - first;
  const uint64_t l0 = tsb_first_tile_len(dst);
  const uint64_t off = k == 0 ? 0 : l0 + (k - 1) * (uint64_t)TSB_TILE_BYTES;
  uint64_t len = k == 0 ? l0 : (uint64_t)TSB_TILE_BYTES;
  if (len > nbytes - off) len = nbytes - off;
  const uint64_t s = src + off, d = dst + off;
  uint4* out = reinterpret_cast<uint4*>(tiles + t);
  out[0] = make_uint4((uint32_t)s, (uint32_t)(s >> 32), (uint32_t)d, (uint32_t)(d >> 32));
  out[1] = make_uint4((uint32_t)len, 0xffffffffu, 0u, 0u);
}

// ---------------------------------------------------------------------------------------------
// flat kernel
// ---------------------------------------------------------------------------------------------
template <int SW>
__device__ __forceinline__ void flat_shifted(uint4* __restrict__ d4, const uint4* __restrict__ a4, uint32_t nvec,
                                             uint32_t sb, uint32_t tid) {
  // a4[j] and a4[j+1] both lie in aligned 16-byte granules that contain payload bytes, so the
  // extra vector never leaves the allocation that holds the run.
  uint4 a[TSB_FLAT_UNROLL], b[TSB_FLAT_UNROLL];
#pragma unroll
  for (int u = 0; u < TSB_FLAT_UNROLL; ++u) {
    const uint32_t j = tid + u * TSB_FLAT_THREADS;
    if (j < nvec) {
      a[u] = __ldg(a4 + j);
      b[u] = __ldg(a4 + j + 1);
    }
  }
#pragma unroll
  for (int u = 0; u < TSB_FLAT_UNROLL; ++u) {
    const uint32_t j = tid + u * TSB_FLAT_THREADS;
    if (j < nvec) st_stream(d4 + j, realign<SW>(a[u], b[u], sb));
  }
}

// ---- TMA (bulk async copy) helpers: global -> shared -> global without register staging ----------
__device__ __forceinline__ uint32_t tsb_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// One tile whose source, destination and length are all multiples of 16 bytes: a single thread moves
// it with cp.async.bulk (SASS: UBLKCP) through the CTA's shared-memory buffer.
__device__ __forceinline__ void flat_tile_tma(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint32_t len,
                                              uint8_t* smem, uint64_t* bar) {
  const uint32_t bar_a = tsb_smem_u32(bar), buf_a = tsb_smem_u32(smem);
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(len) : "memory");
  asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(buf_a),
               "l"(src), "r"(len), "r"(bar_a)
               : "memory");
  asm volatile(
      "{\n.reg .pred p;\nTSB_WAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra TSB_DONE_%=;\nbra "
      "TSB_WAIT_%=;\nTSB_DONE_%=:\n}" ::"r"(bar_a)
      : "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(buf_a), "r"(len) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// USE_TMA = false: the pure LSU kernel (no shared memory, 32 registers, 2048 threads per SM).
// USE_TMA = true : tiles that are 16-byte aligned on both sides go through cp.async.bulk; the rest (first
//                  / last tile of a run, shifted sources) take the LSU path of the same kernel.  Launched
//                  with TSB_TILE_BYTES of dynamic shared memory.  On B200 it is 0.8-1.7 % faster than the
//                  LSU path on aligned tiles and identical elsewhere (profiles/r01_kernel_bench_tma.txt),
//                  so it is the default; TSB200_USE_TMA=0 selects the LSU kernel.
template <bool USE_TMA>
__device__ __forceinline__ void flat_tile(const TsbTile* __restrict__ tiles, uint8_t* smem, uint64_t* bar) {
  const uint32_t tid = threadIdx.x;
  const uint4* tp = reinterpret_cast<const uint4*>(tiles + blockIdx.x);
  const uint4 t0 = __ldg(tp);
  const uint32_t len = __ldg(reinterpret_cast<const uint32_t*>(tp + 1));
  const uint8_t* src = reinterpret_cast<const uint8_t*>(((uint64_t)t0.y << 32) | t0.x);
  uint8_t* dst = reinterpret_cast<uint8_t*>(((uint64_t)t0.w << 32) | t0.z);
  if (USE_TMA) {
    // full tiles only: a CTA has one bulk copy in flight, so short tiles (small tensors, 8-16 KiB rows) would
    // leave too few bytes in flight per SM (measured: 16 KiB rows 6.43 TB/s by TMA vs 6.85 TB/s by LSU)
    if ((((uint32_t)(uintptr_t)src | (uint32_t)(uintptr_t)dst) & 15u) == 0 && len == TSB_TILE_BYTES) {
      if (tid == 0) flat_tile_tma(dst, src, len, smem, bar);
      return;
    }
  }
  // the host puts interior tile boundaries on 16-byte boundaries of the destination, so only the
  // first tile of a region has a head and only the last one a tail
  uint32_t head = (16u - ((uint32_t)(uintptr_t)dst & 15u)) & 15u;
  if (head > len) head = len;
  if (tid < head) dst[tid] = src[tid];
  src += head;
  dst += head;
  const uint32_t n = len - head;
  const uint32_t nvec = n >> 4;
  const uint32_t tail = n & 15u;
  if (tid >= TSB_FLAT_THREADS - tail) {
    const uint32_t k = (nvec << 4) + (tid - (TSB_FLAT_THREADS - tail));
    dst[k] = src[k];
  }
  uint4* d4 = reinterpret_cast<uint4*>(dst);
  const uint32_t s = (uint32_t)(uintptr_t)src & 15u;
  if (s == 0) {
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    uint4 v[TSB_FLAT_UNROLL];
#pragma unroll
    for (int u = 0; u < TSB_FLAT_UNROLL; ++u) {
      const uint32_t j = tid + u * TSB_FLAT_THREADS;
      if (j < nvec) v[u] = ld_stream(s4 + j);
    }
#pragma unroll
    for (int u = 0; u < TSB_FLAT_UNROLL; ++u) {
      const uint32_t j = tid + u * TSB_FLAT_THREADS;
      if (j < nvec) st_stream(d4 + j, v[u]);
    }
  } else {
    const uint4* a4 = reinterpret_cast<const uint4*>(src - s);
    const uint32_t sb = (s & 3u) * 8u;
    switch (s >> 2) {
      case 0: flat_shifted<0>(d4, a4, nvec, sb, tid); break;
      case 1: flat_shifted<1>(d4, a4, nvec, sb, tid); break;
      case 2: flat_shifted<2>(d4, a4, nvec, sb, tid); break;
      default: flat_shifted<3>(d4, a4, nvec, sb, tid); break;
    }
  }
}

__global__ void __launch_bounds__(TSB_FLAT_THREADS)
tsb_copy_flat_kernel(const TsbTile* __restrict__ tiles) {
  flat_tile<false>(tiles, nullptr, nullptr);
}

__global__ void __launch_bounds__(TSB_FLAT_THREADS)
tsb_copy_flat_tma_kernel(const TsbTile* __restrict__ tiles) {
  extern __shared__ __align__(128) uint8_t tsb_tma_smem[];
  __shared__ uint64_t tsb_tma_bar;
  flat_tile<true>(tiles, tsb_tma_smem, &tsb_tma_bar);
}

// ---------------------------------------------------------------------------------------------
// strided kernel
// ---------------------------------------------------------------------------------------------
template <int SW, int NL>
__device__ __forceinline__ void copy_vec_shifted(uint4* __restrict__ d4, const uint4* __restrict__ a4,
                                                 uint64_t nvec, uint32_t sb, uint32_t lane) {
  uint64_t j = lane;
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
    uint4 a = ld_vec16<(NL <= 8)>(a4 + j);
    uint4 b = ld_vec16<(NL <= 8)>(a4 + j + 1);
    st_stream(d4 + j, realign<SW>(a, b, sb));
  }
}

// Cooperative byte-exact copy of n bytes by a group of NL threads (lane in [0, NL)); NL = 4, 8, 16 or 32.
template <int NL>
__device__ __forceinline__ void copy_span(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint64_t n,
                                          uint32_t lane) {
  if (n == 0) return;
  uint32_t head = (uint32_t)((16u - ((uint32_t)(uintptr_t)dst & 15u)) & 15u);
  if ((uint64_t)head > n) head = (uint32_t)n;
  if (NL >= 16) {
    if (lane < head) dst[lane] = src[lane];
  } else {
    for (uint32_t b = lane; b < head; b += NL) dst[b] = src[b];
  }
  dst += head;
  src += head;
  n -= head;
  const uint64_t nvec = n >> 4;
  const uint32_t tail = (uint32_t)(n & 15u);
  // tail bytes by the last lanes, so head and tail do not serialise on the same threads
  if (NL >= 16) {
    if (lane >= NL - tail) {
      const uint64_t k = (nvec << 4) + (lane - (NL - tail));
      dst[k] = src[k];
    }
  } else {
    for (uint32_t b = NL - 1 - lane; b < tail; b += NL) dst[(nvec << 4) + b] = src[(nvec << 4) + b];
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
  if (ndim == 1) {  // pitched rows, the common case: no division at all
    so = (int64_t)q * __ldg(&r->src_stride[0]);
    dofs = (int64_t)q * __ldg(&r->dst_stride[0]);
    return;
  }
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

// 16-byte unit load that asks L2 to fetch only the 64-byte half line on a miss.  The default promotes a miss
// to the whole 128-byte line, which doubles the DRAM reads of a gather of 64-byte runs: pitched 64-byte rows
// (a column shard of a 128-column fp32 table over 8 ranks) move 4.25 TB/s without the hint and 5.19 TB/s with
// it; longer runs are unaffected, and on 1/2/4-byte units the hint costs 2-4 %, so only this width carries it
// (scripts/strided_bench.py, scripts/strided_sweep.py; TSB200_L2_64B=0 turns it off).
template <typename T>
__device__ __forceinline__ T ld_unit(const T* p, bool hint64) {
  return *p;
}
template <>
__device__ __forceinline__ uint4 ld_unit<uint4>(const uint4* p, bool hint64) {
  if (!hint64) return *p;
  uint4 v;
  asm volatile("ld.global.L2::64B.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// Unit-wise copy: unit u of the tile is element k of run q; consecutive threads take consecutive units,
// so accesses stay coalesced within a run and across adjacent runs.  Four units per thread are loaded
// before the first is stored.
template <typename T>
__device__ __forceinline__ void copy_units(const TsbRegion* __restrict__ r, uint64_t src, uint64_t dst, uint32_t ndim,
                                           uint64_t first_unit, uint64_t nunits, uint32_t units_per_run,
                                           uint32_t tid, bool hint64) {
  constexpr int U = 4;
  for (uint64_t base = 0; base < nunits; base += (uint64_t)U * TSB_STRIDED_THREADS) {
    T v[U];
    T* dp[U];
    bool ok[U];
#pragma unroll
    for (int i = 0; i < U; ++i) {
      const uint64_t u = base + tid + (uint64_t)i * TSB_STRIDED_THREADS;
      ok[i] = u < nunits;
      if (ok[i]) {
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
        run_offsets(r, ndim, q, so, dofs);
        v[i] = ld_unit<T>(reinterpret_cast<const T*>(src + so) + k, hint64);
        dp[i] = reinterpret_cast<T*>(dst + dofs) + k;
      }
    }
#pragma unroll
    for (int i = 0; i < U; ++i)
      if (ok[i]) *dp[i] = v[i];
  }
}

// One run per group of NL lanes (the groups of a CTA take the runs of the tile round-robin).
template <int NL>
__device__ __forceinline__ void strided_runs(const TsbRegion* __restrict__ r, uint64_t src, uint64_t dst, uint32_t ndim,
                                             uint64_t q0, uint32_t nruns, uint64_t run_bytes, uint32_t tid) {
  const uint32_t group = tid / NL, lane = tid % NL;
  for (uint32_t i = group; i < nruns; i += TSB_STRIDED_THREADS / NL) {
    int64_t so, dofs;
    run_offsets(r, ndim, q0 + i, so, dofs);
    copy_span<NL>(reinterpret_cast<uint8_t*>(dst + dofs), reinterpret_cast<const uint8_t*>(src + so), run_bytes, lane);
  }
}

__global__ void __launch_bounds__(TSB_STRIDED_THREADS)
tsb_copy_strided_kernel(const TsbRegion* __restrict__ regions, const TsbTile* __restrict__ tiles,
                        uint32_t warp_run_min_and_flags) {
  const uint32_t tid = threadIdx.x;
  const uint4* tp = reinterpret_cast<const uint4*>(tiles + blockIdx.x);
  const uint4 t1 = __ldg(tp + 1);
  const uint32_t len = t1.x;
  const uint64_t off = ((uint64_t)t1.w << 32) | t1.z;
  const TsbRegion* r = regions + t1.y;
  const uint4* rp = reinterpret_cast<const uint4*>(r);
  const uint4 ra = __ldg(rp);
  const uint4 rb = __ldg(rp + 1);
  const uint64_t src = ((uint64_t)ra.y << 32) | ra.x;
  const uint64_t dst = ((uint64_t)ra.w << 32) | ra.z;
  const uint64_t run_bytes = ((uint64_t)rb.y << 32) | rb.x;
  const uint32_t ndim = rb.z;

  // runs of TSB_FLAT_RUN_MIN bytes or more never get here: the host expands them into flat tiles
  // the widest power-of-two unit (1..16 bytes) every address of the region is aligned to
  uint64_t m = src | dst | run_bytes | 16u;
  for (uint32_t d = 0; d < ndim; ++d) m |= (uint64_t)__ldg(&r->src_stride[d]) | (uint64_t)__ldg(&r->dst_stride[d]);
  const uint32_t unit = (uint32_t)(m & (~m + 1));
  // Runs whose addresses are NOT all 16-byte aligned go one per warp (the span copy realigns them to
  // 16-byte stores) once they are long enough to keep a warp busy; aligned runs and shorter runs are copied
  // unit by unit.  Measured crossover on B200 (scripts/strided_sweep.py): unit-wise copies move 0.79 / 1.45 /
  // 2.65 / 4.0-4.4 TB/s with 1 / 2 / 4 / 8-byte units whatever the run length, one-run-per-warp moves 0.75 /
  // 1.42 / 2.48 / 3.99 / 4.24 / 6.34 TB/s for runs of 64 / 128 / 256 / 512 / 1024 / 4096 bytes -- so the
  // warp form wins from 128 x unit bytes on.
  // (r02) The group that takes a run is as wide as the run is long: 4 lanes up to 64 bytes, 8 up to 128, 16 up to 256, a whole
  // warp above -- a warp per 64-byte run left 28 lanes idle (0.75 TB/s).  With groups, runs misaligned by 1, 2
  // or 4 bytes take this form from 48 bytes on instead of crawling unit by unit (0.79 / 1.45 / 2.65 TB/s);
  // 8-byte-aligned runs keep the unit form (4.0-4.4 TB/s) until the run form overtakes it (`warp_run_min` units).
  const uint32_t warp_run_min = warp_run_min_and_flags & 0x7fffffffu;
  const bool run_form = unit < 16u && (unit >= 8u ? run_bytes >= (uint64_t)warp_run_min * unit : run_bytes >= 48u);
  if (run_form) {
    // the host emits whole runs per tile for this class
    const uint64_t q0 = off / run_bytes;
    const uint32_t nruns = len / (uint32_t)run_bytes;
    if (run_bytes <= 64u) {
      strided_runs<4>(r, src, dst, ndim, q0, nruns, run_bytes, tid);
    } else if (run_bytes <= 128u) {
      strided_runs<8>(r, src, dst, ndim, q0, nruns, run_bytes, tid);
    } else if (run_bytes <= 256u) {
      strided_runs<16>(r, src, dst, ndim, q0, nruns, run_bytes, tid);
    } else {
      strided_runs<32>(r, src, dst, ndim, q0, nruns, run_bytes, tid);
    }
    return;
  }
  const bool hint64 = (warp_run_min_and_flags >> 31) != 0;
  const uint32_t upr = (uint32_t)run_bytes / unit;
  const uint64_t first = off / unit;
  const uint64_t nunits = len / unit;
  switch (unit) {
    case 1: copy_units<uint8_t>(r, src, dst, ndim, first, nunits, upr, tid, hint64); break;
    case 2: copy_units<uint16_t>(r, src, dst, ndim, first, nunits, upr, tid, hint64); break;
    case 4: copy_units<uint32_t>(r, src, dst, ndim, first, nunits, upr, tid, hint64); break;
    case 8: copy_units<uint2>(r, src, dst, ndim, first, nunits, upr, tid, hint64); break;
    default: copy_units<uint4>(r, src, dst, ndim, first, nunits, upr, tid, hint64); break;
  }
}

// ---------------------------------------------------------------------------------------------
// transpose kernel
// ---------------------------------------------------------------------------------------------
// Boxes of single elements whose source is contiguous along one dimension and whose destination is contiguous
// along ANOTHER one: a transposed view packed into a slab (x.t().contiguous() in the reference,
// serialization.py:182-185), or a reshard between layouts whose fastest dimensions differ.  Element by element
// one side of such a copy is always a gather of single elements (1.16 TB/s for fp32).  Here a CTA moves a square
// tile through shared memory: loads run along the source's contiguous dimension, stores along the destination's.
// The host permutes the region's dimensions so that the source-contiguous one is second to last and the
// destination-contiguous one last (the order of a region's dimensions does not change what it copies) and sets
// `reserved` to TSB_REGION_TRANSPOSE; tile `off` counts tiles: (batch index, tile along J, tile along L).
#define TSB_REGION_TRANSPOSE 1u
#define TSB_TRANSPOSE_THREADS 256
#define TSB_TRANSPOSE_SMEM (64 * 65 * 8)

__host__ __device__ __forceinline__ uint32_t tsb_transpose_tile(uint64_t elem_bytes) { return elem_bytes >= 16 ? 32u : 64u; }

template <typename T, int TILE>
__device__ __forceinline__ void transpose_tile(const TsbRegion* __restrict__ r, uint64_t src, uint64_t dst, uint32_t ndim,
                                               uint64_t tile_id, unsigned char* smem_raw, uint32_t tid) {
  T(*tile)[TILE + 1] = reinterpret_cast<T(*)[TILE + 1]>(smem_raw);
  const uint32_t J = ndim - 2, L = ndim - 1;
  const uint64_t nJ = __ldg(&r->size[J]), nL = __ldg(&r->size[L]);
  const uint64_t tJ = (nJ + TILE - 1) / TILE, tL = (nL + TILE - 1) / TILE;
  const uint64_t tl = tile_id % tL;
  tile_id /= tL;
  const uint64_t tj = tile_id % tJ;
  uint64_t b = tile_id / tJ;
  int64_t so = 0, dofs = 0;
  for (int d = (int)ndim - 3; d >= 0; --d) {
    const uint64_t sz = __ldg(&r->size[d]);
    const uint64_t c = b % sz;
    b /= sz;
    so += (int64_t)c * __ldg(&r->src_stride[d]);
    dofs += (int64_t)c * __ldg(&r->dst_stride[d]);
  }
  const uint64_t j0 = tj * TILE, l0 = tl * TILE;
  const int64_t sL = __ldg(&r->src_stride[L]), dJ = __ldg(&r->dst_stride[J]);
  const uint32_t wJ = (uint32_t)(nJ - j0 < TILE ? nJ - j0 : TILE), wL = (uint32_t)(nL - l0 < TILE ? nL - l0 : TILE);
  const unsigned char* sp = reinterpret_cast<const unsigned char*>(src + so) + j0 * sizeof(T) + (int64_t)l0 * sL;
  unsigned char* dp = reinterpret_cast<unsigned char*>(dst + dofs) + (int64_t)j0 * dJ + l0 * sizeof(T);
  constexpr int ROWS = TSB_TRANSPOSE_THREADS / TILE;  // tile rows one pass of the CTA covers
  constexpr int PASSES = TILE / ROWS;
  const uint32_t c = tid % TILE, r0 = tid / TILE;  // this thread's column, and its row in pass 0
  if (sizeof(T) == 2 && wJ == TILE && wL == TILE &&
      ((((uint64_t)(uintptr_t)sp) | ((uint64_t)(uintptr_t)dp) | (uint64_t)sL | (uint64_t)dJ) & 3u) == 0) {
    // 2-byte elements, full tile, rows 4-byte aligned on both sides: move PAIRS.  A warp then reads and writes
    // 128 contiguous bytes per row instead of 64, and issues half the instructions (2.9 -> see DESIGN section 3).
    constexpr int PITCH = TILE + 2;  // in elements: rows start 4-byte aligned, (PITCH / 2) odd -> conflict-free stores
    uint16_t(*t16)[PITCH] = reinterpret_cast<uint16_t(*)[PITCH]>(smem_raw);
    constexpr int PAIRS = TILE / 2;                      // u32 per row
    constexpr int RPP = TSB_TRANSPOSE_THREADS / PAIRS;   // rows per pass
    constexpr int NP = TILE / RPP;
    const uint32_t q = tid % PAIRS, row0 = tid / PAIRS;
    uint32_t w[NP];
#pragma unroll
    for (int p = 0; p < NP; ++p)
      w[p] = *reinterpret_cast<const uint32_t*>(sp + (uint64_t)q * 4 + (int64_t)(row0 + p * RPP) * sL);
#pragma unroll
    for (int p = 0; p < NP; ++p) *reinterpret_cast<uint32_t*>(&t16[row0 + p * RPP][2 * q]) = w[p];
    __syncthreads();
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      const uint32_t j = row0 + p * RPP;
      w[p] = (uint32_t)t16[2 * q][j] | ((uint32_t)t16[2 * q + 1][j] << 16);
    }
#pragma unroll
    for (int p = 0; p < NP; ++p)
      *reinterpret_cast<uint32_t*>(dp + (int64_t)(row0 + p * RPP) * dJ + (uint64_t)q * 4) = w[p];
    return;
  }
  if (wJ == TILE && wL == TILE) {
    // full tile: every load of the thread is in flight before its first shared-memory store
    T v[PASSES];
#pragma unroll
    for (int p = 0; p < PASSES; ++p)
      v[p] = *reinterpret_cast<const T*>(sp + (uint64_t)c * sizeof(T) + (int64_t)(r0 + p * ROWS) * sL);
#pragma unroll
    for (int p = 0; p < PASSES; ++p) tile[r0 + p * ROWS][c] = v[p];
    __syncthreads();
#pragma unroll
    for (int p = 0; p < PASSES; ++p) v[p] = tile[c][r0 + p * ROWS];
#pragma unroll
    for (int p = 0; p < PASSES; ++p)
      *reinterpret_cast<T*>(dp + (int64_t)(r0 + p * ROWS) * dJ + (uint64_t)c * sizeof(T)) = v[p];
    return;
  }
  // edge tile.  In: consecutive threads along J (the source's contiguous dimension) ...
  for (int p = 0; p < PASSES; ++p) {
    const uint32_t l = r0 + p * ROWS;
    if (c < wJ && l < wL) tile[l][c] = *reinterpret_cast<const T*>(sp + (uint64_t)c * sizeof(T) + (int64_t)l * sL);
  }
  __syncthreads();
  // ... out: consecutive threads along L (the destination's contiguous dimension)
  for (int p = 0; p < PASSES; ++p) {
    const uint32_t j = r0 + p * ROWS;
    if (j < wJ && c < wL) *reinterpret_cast<T*>(dp + (int64_t)j * dJ + (uint64_t)c * sizeof(T)) = tile[c][j];
  }
}

__global__ void __launch_bounds__(TSB_TRANSPOSE_THREADS)
tsb_copy_transpose_kernel(const TsbRegion* __restrict__ regions, const TsbTile* __restrict__ tiles) {
  __shared__ __align__(16) unsigned char smem[TSB_TRANSPOSE_SMEM];
  const uint32_t tid = threadIdx.x;
  const uint4 t1 = __ldg(reinterpret_cast<const uint4*>(tiles + blockIdx.x) + 1);
  const uint64_t tile_id = ((uint64_t)t1.w << 32) | t1.z;
  const TsbRegion* r = regions + t1.y;
  const uint4* rp = reinterpret_cast<const uint4*>(r);
  const uint4 ra = __ldg(rp);
  const uint4 rb = __ldg(rp + 1);
  const uint64_t src = ((uint64_t)ra.y << 32) | ra.x;
  const uint64_t dst = ((uint64_t)ra.w << 32) | ra.z;
  const uint32_t elem = rb.x;  // run_bytes: one element
  const uint32_t ndim = rb.z;
  switch (elem) {
    case 1: transpose_tile<uint8_t, 64>(r, src, dst, ndim, tile_id, smem, tid); break;
    case 2: transpose_tile<uint16_t, 64>(r, src, dst, ndim, tile_id, smem, tid); break;
    case 4: transpose_tile<uint32_t, 64>(r, src, dst, ndim, tile_id, smem, tid); break;
    case 8: transpose_tile<uint2, 64>(r, src, dst, ndim, tile_id, smem, tid); break;
    default: transpose_tile<uint4, 32>(r, src, dst, ndim, tile_id, smem, tid); break;
  }
}
