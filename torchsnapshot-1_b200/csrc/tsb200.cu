// libtsb200.so -- C-ABI implementation (see include/tsb200.h for the contract and the
// reference call sites each entry point replaces).  CUDA runtime only; no torch types.
#include <cuda_runtime.h>
#include <errno.h>
#include <fcntl.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "copy_kernel.cuh"

static_assert(sizeof(TsbRegion) == 128, "TsbRegion must be 128 bytes");
static_assert(sizeof(TsbTile) == 32, "TsbTile must be 32 bytes");
static_assert(sizeof(TsbRun) == 32, "TsbRun must be 32 bytes");

// ------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int set_err(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define CUDA_TRY(expr)                                                                             \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      return set_err(TSB_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                     __LINE__);                                                                    \
  } while (0)

// ------------------------------------------------------------------------------------------
// arena: segments obtained from the driver on demand, first-fit free lists inside
// ------------------------------------------------------------------------------------------
struct Segment {
  uint64_t base = 0;
  uint64_t size = 0;
  std::map<uint64_t, uint64_t> free_blocks;  // offset -> size
  std::map<uint64_t, uint64_t> used_blocks;  // offset -> size
};

class Arena {
 public:
  Arena(bool host, uint64_t bound, uint64_t seg_default, uint64_t align)
      : host_(host), bound_(bound), seg_default_(seg_default), align_(align) {}

  int alloc(uint64_t nbytes, uint64_t* out, bool may_grow = true) {
    if (nbytes == 0) nbytes = 1;
    const uint64_t need = (nbytes + align_ - 1) / align_ * align_;
    std::lock_guard<std::mutex> g(mu_);
    if (try_alloc(need, out)) return TSB_OK;
    if (!may_grow) return set_err(TSB_EAGAIN, "%s arena: no free block of %llu bytes in the existing segments",
                                  host_ ? "host" : "device", (unsigned long long)need);
    if (need > bound_)
      return set_err(TSB_ENOMEM, "%s arena: request of %llu bytes exceeds the arena bound of %llu bytes",
                     host_ ? "host" : "device", (unsigned long long)need, (unsigned long long)bound_);
    // grow: a new segment if the bound allows it, else drop empty segments first
    if (reserved_ + need > bound_) release_empty_segments();
    if (reserved_ + need > bound_) return set_err(TSB_EAGAIN, "%s arena full", host_ ? "host" : "device");
    uint64_t seg = std::max(need, seg_default_);
    if (!host_ && need > seg_default_) {
      // a staging block for a whole wave: an eighth of headroom, so that the next take, whose share of the
      // replicated writes may be a few percent larger, fits the same segment (a fresh cudaMalloc of gigabytes
      // with peer mappings present was seen to hold async_take up for tens of milliseconds)
      const uint64_t g = 64ull << 20;
      seg = (need + need / 8 + g - 1) / g * g;
    }
    if (reserved_ + seg > bound_) seg = std::max(need, bound_ - reserved_);
    if (reserved_ + seg > bound_) seg = bound_ - reserved_;
    uint64_t base = 0;
    cudaError_t e = raw_alloc(seg, &base);
    if (e != cudaSuccess && seg > need) {
      (void)cudaGetLastError();
      seg = need;
      e = raw_alloc(seg, &base);
    }
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      if (in_use_ > 0) return set_err(TSB_EAGAIN, "%s arena: driver allocation failed (%s), retry after a free",
                                      host_ ? "host" : "device", cudaGetErrorString(e));
      return set_err(TSB_ENOMEM, "%s arena: driver allocation of %llu bytes failed: %s", host_ ? "host" : "device",
                     (unsigned long long)seg, cudaGetErrorString(e));
    }
    Segment s;
    s.base = base;
    s.size = seg;
    s.free_blocks[0] = seg;
    segs_.push_back(std::move(s));
    reserved_ += seg;
    if (!try_alloc(need, out)) return set_err(TSB_ENOMEM, "arena: internal error");
    return TSB_OK;
  }

  int free(uint64_t addr) {
    std::lock_guard<std::mutex> g(mu_);
    for (auto& s : segs_) {
      if (addr < s.base || addr >= s.base + s.size) continue;
      const uint64_t off = addr - s.base;
      auto it = s.used_blocks.find(off);
      if (it == s.used_blocks.end()) break;
      uint64_t sz = it->second;
      s.used_blocks.erase(it);
      in_use_ -= sz;
      // coalesce with neighbours
      uint64_t lo = off, hi = off + sz;
      auto nxt = s.free_blocks.lower_bound(off);
      if (nxt != s.free_blocks.end() && nxt->first == hi) {
        hi += nxt->second;
        nxt = s.free_blocks.erase(nxt);
      }
      if (nxt != s.free_blocks.begin()) {
        auto prv = std::prev(nxt);
        if (prv->first + prv->second == lo) {
          lo = prv->first;
          s.free_blocks.erase(prv);
        }
      }
      s.free_blocks[lo] = hi - lo;
      return TSB_OK;
    }
    return set_err(TSB_EINVAL, "arena: address 0x%llx was not allocated here", (unsigned long long)addr);
  }

  // segment that contains addr (for IPC export)
  bool find_segment(uint64_t addr, uint64_t* base, uint64_t* size) {
    std::lock_guard<std::mutex> g(mu_);
    for (auto& s : segs_)
      if (addr >= s.base && addr < s.base + s.size) {
        *base = s.base;
        *size = s.size;
        return true;
      }
    return false;
  }

  void stats(uint64_t* in_use, uint64_t* reserved, uint64_t* bound) {
    std::lock_guard<std::mutex> g(mu_);
    *in_use = in_use_;
    *reserved = reserved_;
    *bound = bound_;
  }

  void trim() {
    std::lock_guard<std::mutex> g(mu_);
    release_empty_segments();
  }

  void destroy() {
    std::lock_guard<std::mutex> g(mu_);
    for (auto& s : segs_) raw_free(s.base);
    segs_.clear();
    reserved_ = in_use_ = 0;
  }

 private:
  bool try_alloc(uint64_t need, uint64_t* out) {
    for (auto& s : segs_) {
      for (auto it = s.free_blocks.begin(); it != s.free_blocks.end(); ++it) {
        if (it->second < need) continue;
        const uint64_t off = it->first, sz = it->second;
        s.free_blocks.erase(it);
        if (sz > need) s.free_blocks[off + need] = sz - need;
        s.used_blocks[off] = need;
        in_use_ += need;
        *out = s.base + off;
        return true;
      }
    }
    return false;
  }
  void release_empty_segments() {
    for (size_t i = 0; i < segs_.size();) {
      if (segs_[i].used_blocks.empty()) {
        raw_free(segs_[i].base);
        reserved_ -= segs_[i].size;
        segs_.erase(segs_.begin() + i);
      } else {
        ++i;
      }
    }
  }
  // Pinned host segments: an anonymous mapping with transparent-huge-page advice, first-touched by
  // several threads, then registered with the driver.  On the B200 box this costs 39 ms per GiB against
  // 390-490 ms per GiB for cudaHostAlloc (which faults 4 KiB pages in on one thread), with the same D2H
  // rate (scripts/pinned_alloc_probe.cu) -- the first take of a 9 GB state otherwise spends 3.7 s here.
  // TSB200_PINNED_ALLOC=hostalloc selects cudaHostAlloc.
  cudaError_t raw_alloc_host(uint64_t n, uint64_t* base) {
    static const bool use_hostalloc = [] {
      const char* v = getenv("TSB200_PINNED_ALLOC");
      return v != nullptr && strcmp(v, "hostalloc") == 0;
    }();
    if (!use_hostalloc) {
      const uint64_t huge = 2ull << 20;
      const uint64_t len = n + huge;
      void* m = mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
      if (m != MAP_FAILED) {
        char* q = (char*)(((uintptr_t)m + huge - 1) & ~(uintptr_t)(huge - 1));
        (void)madvise(q, n, MADV_HUGEPAGE);
        // like driver-allocated pinned memory, the segment is not inherited by forked children (data-loader
        // workers): no copy-on-write can ever move a page the DMA engines are writing to
        (void)madvise(q, n, MADV_DONTFORK);
        unsigned hw = std::thread::hardware_concurrency();
        int nthreads = (int)std::min<uint64_t>(std::min<unsigned>(hw ? hw : 4, 16), std::max<uint64_t>(n >> 26, 1));
        const uint64_t per = ((n + nthreads - 1) / nthreads + 4095) & ~4095ull;
        auto touch = [q, n, per](int t) {
          const uint64_t lo = (uint64_t)t * per, hi = std::min(n, lo + per);
          for (uint64_t o = lo; o < hi; o += 4096) ((volatile char*)q)[o] = 0;
        };
        std::vector<std::thread> th;
        for (int t = 1; t < nthreads; ++t) th.emplace_back(touch, t);
        touch(0);
        for (auto& t : th) t.join();
        // Mapped: SMs may store straight into the segment (the pack kernel writes host memory when no device
        // staging is available); under unified addressing the device uses the host address as it is
        cudaError_t e = cudaHostRegister(q, n, cudaHostRegisterPortable | cudaHostRegisterMapped);
        if (e == cudaSuccess) {
          mapped_[(uint64_t)(uintptr_t)q] = {(uint64_t)(uintptr_t)m, len};
          *base = (uint64_t)(uintptr_t)q;
          return cudaSuccess;
        }
        (void)cudaGetLastError();
        munmap(m, len);
      }
    }
    void* p = nullptr;
    cudaError_t e = cudaHostAlloc(&p, n, cudaHostAllocPortable | cudaHostAllocMapped);
    *base = (uint64_t)(uintptr_t)p;
    return e;
  }
  cudaError_t raw_alloc(uint64_t n, uint64_t* base) {
    if (host_) return raw_alloc_host(n, base);
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, n);
    *base = (uint64_t)(uintptr_t)p;
    return e;
  }
  void raw_free(uint64_t base) {
    if (!host_) {
      cudaFree((void*)(uintptr_t)base);
      return;
    }
    auto it = mapped_.find(base);
    if (it == mapped_.end()) {
      cudaFreeHost((void*)(uintptr_t)base);
      return;
    }
    cudaHostUnregister((void*)(uintptr_t)base);
    munmap((void*)(uintptr_t)it->second.first, it->second.second);
    mapped_.erase(it);
  }

  bool host_;
  uint64_t bound_, seg_default_, align_;
  uint64_t reserved_ = 0, in_use_ = 0;
  std::vector<Segment> segs_;
  std::map<uint64_t, std::pair<uint64_t, uint64_t>> mapped_;  // registered base -> (mmap address, length)
  std::mutex mu_;
};

// ------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------
struct PendingScratch {
  cudaEvent_t done;
  uint64_t hbuf;
  uint64_t dbuf;
};

struct TsbCtx {
  int device = 0;
  int sm_count = 0;
  cudaStream_t streams[TSB_NUM_STREAMS] = {};
  Arena* host_arena = nullptr;
  Arena* dev_arena = nullptr;
  // descriptor / tile-table scratch of the launches: separate arenas, so that payload buffers filling the
  // pinned or staging arena to their bound can never starve a launch of its (small) tables
  Arena* scratch_host = nullptr;
  Arena* scratch_dev = nullptr;
  std::mutex mu;  // events + scratch list
  std::vector<cudaEvent_t> events;
  std::vector<uint32_t> event_gen;
  std::vector<uint32_t> free_events;
  std::vector<PendingScratch> pending;
  std::vector<cudaEvent_t> spare_scratch_events;
  // device blocks that are driver allocations of their own (tsb_dev_alloc_exportable): base -> size
  std::map<uint64_t, uint64_t> standalone;
  bool use_tma = true;  // aligned tiles through cp.async.bulk; TSB200_USE_TMA=0 turns it off
  // counters (tsb_ctx_counters)
  uint64_t n_kernel_launches = 0;  // expand + flat + strided kernels
  uint64_t n_copy_calls = 0;       // tsb_copy_regions calls that launched something
  uint64_t desc_upload_bytes = 0;  // descriptor bytes copied host -> device
  uint64_t kernel_payload_bytes = 0;
};

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
  }
  ~DeviceGuard() {
    int cur = -1;
    if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
  }
};

#define CHECK_CTX(ctx)                                     \
  if ((ctx) == nullptr) return set_err(TSB_EINVAL, "null context"); \
  DeviceGuard _guard((ctx)->device);                       \
  if (!_guard.ok) return set_err(TSB_ECUDA, "cudaSetDevice(%d) failed", (ctx)->device)

#define CHECK_STREAM(s) \
  if ((s) < 0 || (s) >= TSB_NUM_STREAMS) return set_err(TSB_EINVAL, "bad stream index %d", (s))

static int event_lookup(TsbCtx* ctx, TsbEvent h, cudaEvent_t* out) {
  const uint32_t idx = (uint32_t)(h & 0xffffffffu), gen = (uint32_t)(h >> 32);
  std::lock_guard<std::mutex> g(ctx->mu);
  if (idx == 0 || idx > ctx->events.size() || ctx->event_gen[idx - 1] != gen)
    return set_err(TSB_EINVAL, "stale or invalid event handle 0x%llx", (unsigned long long)h);
  *out = ctx->events[idx - 1];
  return TSB_OK;
}

// free scratch of launches that have completed (caller holds no lock).  With `wait_all` the launches still
// in flight are waited for -- outside ctx->mu, so event records / queries of other threads never stall behind
// a device wait.
static void reap_scratch(TsbCtx* ctx, bool wait_all) {
  std::vector<PendingScratch> done, waiting;
  {
    std::lock_guard<std::mutex> g(ctx->mu);
    for (size_t i = 0; i < ctx->pending.size();) {
      cudaError_t q = cudaEventQuery(ctx->pending[i].done);
      if (q == cudaErrorNotReady) {
        (void)cudaGetLastError();
        if (wait_all) {
          waiting.push_back(ctx->pending[i]);
          ctx->pending.erase(ctx->pending.begin() + i);
        } else {
          ++i;
        }
        continue;
      }
      if (q != cudaSuccess) (void)cudaGetLastError();
      done.push_back(ctx->pending[i]);
      ctx->pending.erase(ctx->pending.begin() + i);
    }
  }
  for (auto& p : waiting) {
    if (cudaEventSynchronize(p.done) != cudaSuccess) (void)cudaGetLastError();
    done.push_back(p);
  }
  if (!done.empty()) {
    std::lock_guard<std::mutex> g(ctx->mu);
    for (auto& p : done) ctx->spare_scratch_events.push_back(p.done);
  }
  for (auto& p : done) {
    if (p.hbuf) ctx->scratch_host->free(p.hbuf);
    if (p.dbuf) ctx->scratch_dev->free(p.dbuf);
  }
}

// ------------------------------------------------------------------------------------------
// tile tables
// ------------------------------------------------------------------------------------------
static int validate_region(const TsbRegion& r, uint32_t i) {
  if (r.ndim > TSB_MAX_DIMS) return set_err(TSB_EINVAL, "region %u: ndim %u > %d", i, r.ndim, TSB_MAX_DIMS);
  if (r.run_bytes == 0) return TSB_OK;
  if ((r.src == 0 || r.dst == 0)) return set_err(TSB_EINVAL, "region %u: null address", i);
  return TSB_OK;
}

// Payload bytes of a region; false on overflow.
static bool region_payload(const TsbRegion& r, uint64_t* total) {
  unsigned __int128 t = r.run_bytes;
  for (uint32_t d = 0; d < r.ndim; ++d) {
    t *= r.size[d];
    if (t >> 63) return false;
  }
  *total = (uint64_t)t;
  return true;
}

// Queue one contiguous run; its tiles are expanded on the device (tsb_expand_flat_kernel).
static void flat_run(uint64_t src, uint64_t dst, uint64_t n, std::vector<TsbRun>& runs, uint64_t* n_tiles) {
  if (n == 0) return;
  TsbRun r;
  r.src = src;
  r.dst = dst;
  r.nbytes = n;
  r.first_tile = *n_tiles;
  runs.push_back(r);
  *n_tiles += tsb_run_tiles(dst, n);
}

// Launch parameter of the strided kernel.  Low 31 bits: shortest run, in alignment units, that is copied
// one-per-warp when its addresses are not 16-byte aligned (shorter ones go unit by unit); TSB200_WARP_RUN_MIN
// overrides the default for measurements.
static uint32_t warp_run_min() {
  static const uint32_t v = [] {
    const char* e = getenv("TSB200_WARP_RUN_MIN");
    long x = e ? atol(e) : 0;
    uint32_t r = x > 0 ? (uint32_t)x : (uint32_t)TSB_WARP_RUN_MIN;
    const char* h = getenv("TSB200_L2_64B");  // bit 31: 64-byte L2 fetch hint on 16-byte unit loads (default on)
    if (!(h && h[0] == '0')) r |= 0x80000000u;
    return r;
  }();
  return v;
}

struct TilePlan {
  std::vector<TsbRun> flat_runs;
  uint64_t flat_tiles = 0;
  uint64_t payload = 0;
  std::vector<TsbTile> strided;
  std::vector<TsbTile> transposed;           // tiles of tsb_copy_transpose_kernel (`off` = tile number in its region)
  std::vector<TsbRegion> strided_regions;    // regions of both tile lists
};

// A box of single elements that is contiguous along one dimension on the source side and along another on the
// destination side (a transposed view being packed, a reshard between layouts with different fastest
// dimensions): returns the region with its dimensions reordered to [the others..., J, L], J the source's
// contiguous dimension and L the destination's, for tsb_copy_transpose_kernel.  TSB200_TRANSPOSE=0 turns the
// class off (such boxes are then copied element by element by the strided kernel).
static bool as_transpose(const TsbRegion& r, TsbRegion* out) {
  static const bool enabled = [] {
    const char* v = getenv("TSB200_TRANSPOSE");
    return !(v && v[0] == '0');
  }();
  if (!enabled || r.ndim < 2) return false;
  const uint64_t es = r.run_bytes;
  if (es != 1 && es != 2 && es != 4 && es != 8 && es != 16) return false;
  int J = -1, L = -1;
  uint64_t misalign = r.src | r.dst;
  for (uint32_t d = 0; d < r.ndim; ++d) {
    misalign |= (uint64_t)r.src_stride[d] | (uint64_t)r.dst_stride[d];
    if (r.src_stride[d] == (int64_t)es && r.dst_stride[d] != (int64_t)es && J < 0) J = (int)d;
    if (r.dst_stride[d] == (int64_t)es && r.src_stride[d] != (int64_t)es) L = (int)d;
  }
  if (J < 0 || L < 0 || J == L || (misalign & (es - 1))) return false;
  // worth a tile only when both extents fill a good part of it
  if (r.size[J] < 16 || r.size[L] < 16) return false;
  TsbRegion t = r;
  uint32_t k = 0;
  for (uint32_t d = 0; d < r.ndim; ++d) {
    if ((int)d == J || (int)d == L) continue;
    t.size[k] = r.size[d];
    t.src_stride[k] = r.src_stride[d];
    t.dst_stride[k] = r.dst_stride[d];
    ++k;
  }
  t.size[k] = r.size[J];
  t.src_stride[k] = r.src_stride[J];
  t.dst_stride[k] = r.dst_stride[J];
  ++k;
  t.size[k] = r.size[L];
  t.src_stride[k] = r.src_stride[L];
  t.dst_stride[k] = r.dst_stride[L];
  t.reserved = TSB_REGION_TRANSPOSE;
  *out = t;
  return true;
}

static int build_tiles(const TsbRegion* regions, uint32_t n, TilePlan& plan) {
  for (uint32_t i = 0; i < n; ++i) {
    const TsbRegion& r = regions[i];
    int rc = validate_region(r, i);
    if (rc != TSB_OK) return rc;
    uint64_t total;
    if (!region_payload(r, &total)) return set_err(TSB_EINVAL, "region %u: size overflow", i);
    if (total == 0) continue;
    plan.payload += total;
    if (r.ndim == 0) {
      flat_run(r.src, r.dst, r.run_bytes, plan.flat_runs, &plan.flat_tiles);
      continue;
    }
    if (r.run_bytes >= TSB_FLAT_RUN_MIN) {
      // long runs: every run is a contiguous copy of its own
      uint64_t idx[TSB_MAX_DIMS] = {0, 0, 0, 0};
      const uint64_t runs = total / r.run_bytes;
      for (uint64_t q = 0; q < runs; ++q) {
        int64_t so = 0, dofs = 0;
        for (uint32_t d = 0; d < r.ndim; ++d) {
          so += (int64_t)idx[d] * r.src_stride[d];
          dofs += (int64_t)idx[d] * r.dst_stride[d];
        }
        flat_run(r.src + so, r.dst + dofs, r.run_bytes, plan.flat_runs, &plan.flat_tiles);
        for (int d = (int)r.ndim - 1; d >= 0; --d) {
          if (++idx[d] < r.size[d]) break;
          idx[d] = 0;
        }
      }
      continue;
    }
    const uint32_t ridx = (uint32_t)plan.strided_regions.size();
    TsbRegion tr;
    if (as_transpose(r, &tr)) {
      plan.strided_regions.push_back(tr);
      const uint64_t tile = tsb_transpose_tile(r.run_bytes);
      uint64_t count = ((tr.size[tr.ndim - 2] + tile - 1) / tile) * ((tr.size[tr.ndim - 1] + tile - 1) / tile);
      for (uint32_t d = 0; d + 2 < tr.ndim; ++d) count *= tr.size[d];
      for (uint64_t id = 0; id < count; ++id) {
        TsbTile t;
        t.src = 0;
        t.dst = 0;
        t.off = id;
        t.len = 0;
        t.region = ridx;
        plan.transposed.push_back(t);
      }
      continue;
    }
    // short runs: tiles hold whole runs; the strided kernel walks them
    plan.strided_regions.push_back(r);
    plan.strided_regions.back().reserved = 0;
    uint64_t runs_per_tile = TSB_TILE_BYTES / r.run_bytes;
    if (runs_per_tile == 0) runs_per_tile = 1;
    const uint64_t step = runs_per_tile * r.run_bytes;
    for (uint64_t off = 0; off < total; off += step) {
      TsbTile t;
      t.src = 0;
      t.dst = 0;
      t.off = off;
      t.len = (uint32_t)std::min<uint64_t>(step, total - off);
      t.region = ridx;
      plan.strided.push_back(t);
    }
  }
  return TSB_OK;
}

// ------------------------------------------------------------------------------------------
// C-ABI
// ------------------------------------------------------------------------------------------
extern "C" {

int tsb_version(void) { return 1003; }

const char* tsb_last_error(void) { return g_err; }

int tsb_device_count(int* count) {
  if (!count) return set_err(TSB_EINVAL, "null argument");
  CUDA_TRY(cudaGetDeviceCount(count));
  return TSB_OK;
}

int tsb_ctx_create(int device, uint64_t host_arena_max, uint64_t dev_arena_max, TsbCtx** out) {
  if (!out) return set_err(TSB_EINVAL, "null argument");
  int count = 0;
  CUDA_TRY(cudaGetDeviceCount(&count));
  if (device < 0 || device >= count) return set_err(TSB_EINVAL, "device %d out of range (%d visible)", device, count);
  DeviceGuard guard(device);
  if (!guard.ok) return set_err(TSB_ECUDA, "cudaSetDevice(%d) failed", device);
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10)
    return set_err(TSB_ECUDA, "device %d is sm_%d%d; libtsb200 carries sm_100a code only", device, prop.major,
                   prop.minor);
  TsbCtx* ctx = new TsbCtx();
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  for (int i = 0; i < TSB_NUM_STREAMS; ++i) {
    cudaError_t e = cudaStreamCreateWithFlags(&ctx->streams[i], cudaStreamNonBlocking);
    if (e != cudaSuccess) {
      delete ctx;
      return set_err(TSB_ECUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(e));
    }
  }
  {
    // default on: measured 0.8-1.7 % faster than the LSU path on aligned tiles and equal elsewhere
    // (profiles/r01_kernel_bench_tma.txt); TSB200_USE_TMA=0 selects the pure LSU kernel
    const char* tma = getenv("TSB200_USE_TMA");
    ctx->use_tma = !(tma && tma[0] == '0');
    if (ctx->use_tma &&
        cudaFuncSetAttribute(tsb_copy_flat_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TSB_TILE_BYTES) !=
            cudaSuccess) {
      (void)cudaGetLastError();
      ctx->use_tma = false;
    }
  }
  if (host_arena_max == 0) host_arena_max = 4ull << 30;
  if (dev_arena_max == 0) dev_arena_max = 16ull << 30;
  // Pinned segments are all of one size -- the largest object the writer produces by default (512 MiB chunks and
  // shard pieces) -- so any empty segment serves any request: an arena at its bound then waits for a block to
  // come back instead of returning odd-sized empty segments to the driver and pinning new ones (unregister +
  // unmap + first touch + cudaHostRegister, tens of milliseconds each, and cudaHostRegister stalls every other
  // CUDA call of the process meanwhile).  TSB200_PINNED_SEGMENT_BYTES overrides.
  uint64_t host_seg = 512ull << 20;
  if (const char* v = getenv("TSB200_PINNED_SEGMENT_BYTES")) {
    const long long x = atoll(v);
    if (x >= (1ll << 20)) host_seg = (uint64_t)x;
  }
  if (host_seg > host_arena_max) host_seg = host_arena_max;
  ctx->host_arena = new Arena(true, host_arena_max, host_seg, 4096);
  ctx->dev_arena = new Arena(false, dev_arena_max, 256ull << 20, 512);
  ctx->scratch_host = new Arena(true, 1ull << 30, 4ull << 20, 4096);
  ctx->scratch_dev = new Arena(false, 4ull << 30, 16ull << 20, 512);
  *out = ctx;
  return TSB_OK;
}

int tsb_ctx_destroy(TsbCtx* ctx) {
  CHECK_CTX(ctx);
  for (int i = 0; i < TSB_NUM_STREAMS; ++i) cudaStreamSynchronize(ctx->streams[i]);
  reap_scratch(ctx, true);
  for (auto e : ctx->events) cudaEventDestroy(e);
  for (auto e : ctx->spare_scratch_events) cudaEventDestroy(e);
  for (int i = 0; i < TSB_NUM_STREAMS; ++i) cudaStreamDestroy(ctx->streams[i]);
  for (auto& kv : ctx->standalone) cudaFree((void*)(uintptr_t)kv.first);
  ctx->standalone.clear();
  ctx->host_arena->destroy();
  ctx->dev_arena->destroy();
  ctx->scratch_host->destroy();
  ctx->scratch_dev->destroy();
  delete ctx->host_arena;
  delete ctx->dev_arena;
  delete ctx->scratch_host;
  delete ctx->scratch_dev;
  delete ctx;
  (void)cudaGetLastError();
  return TSB_OK;
}

int tsb_ctx_device(TsbCtx* ctx, int* device) {
  if (!ctx || !device) return set_err(TSB_EINVAL, "null argument");
  *device = ctx->device;
  return TSB_OK;
}

int tsb_ctx_sm_count(TsbCtx* ctx, int* sms) {
  if (!ctx || !sms) return set_err(TSB_EINVAL, "null argument");
  *sms = ctx->sm_count;
  return TSB_OK;
}

int tsb_ctx_counters(TsbCtx* ctx, uint64_t* kernel_launches, uint64_t* copy_calls, uint64_t* desc_upload_bytes,
                     uint64_t* kernel_payload_bytes) {
  if (!ctx) return set_err(TSB_EINVAL, "null context");
  std::lock_guard<std::mutex> g(ctx->mu);
  if (kernel_launches) *kernel_launches = ctx->n_kernel_launches;
  if (copy_calls) *copy_calls = ctx->n_copy_calls;
  if (desc_upload_bytes) *desc_upload_bytes = ctx->desc_upload_bytes;
  if (kernel_payload_bytes) *kernel_payload_bytes = ctx->kernel_payload_bytes;
  return TSB_OK;
}

int tsb_dev_alloc(TsbCtx* ctx, uint64_t nbytes, uint64_t* dptr) {
  CHECK_CTX(ctx);
  if (!dptr) return set_err(TSB_EINVAL, "null argument");
  int rc = ctx->dev_arena->alloc(nbytes, dptr);
  if (rc == TSB_EAGAIN) {
    reap_scratch(ctx, false);
    rc = ctx->dev_arena->alloc(nbytes, dptr);
  }
  return rc;
}

int tsb_dev_alloc_exportable(TsbCtx* ctx, uint64_t nbytes, uint64_t* dptr) {
  CHECK_CTX(ctx);
  if (!dptr) return set_err(TSB_EINVAL, "null argument");
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, nbytes ? nbytes : 1);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    // give cached staging segments back and try once more before reporting the shortage
    ctx->dev_arena->trim();
    e = cudaMalloc(&p, nbytes ? nbytes : 1);
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      return set_err(TSB_EAGAIN, "exportable device block of %llu bytes: %s", (unsigned long long)nbytes,
                     cudaGetErrorString(e));
    }
  }
  std::lock_guard<std::mutex> g(ctx->mu);
  ctx->standalone[(uint64_t)(uintptr_t)p] = nbytes;
  *dptr = (uint64_t)(uintptr_t)p;
  return TSB_OK;
}

int tsb_dev_free(TsbCtx* ctx, uint64_t dptr) {
  CHECK_CTX(ctx);
  bool standalone = false;
  {
    std::lock_guard<std::mutex> g(ctx->mu);
    auto it = ctx->standalone.find(dptr);
    if (it != ctx->standalone.end()) {
      ctx->standalone.erase(it);
      standalone = true;
    }
  }
  if (standalone) {  // cudaFree synchronises the device: never under ctx->mu
    CUDA_TRY(cudaFree((void*)(uintptr_t)dptr));
    return TSB_OK;
  }
  return ctx->dev_arena->free(dptr);
}

int tsb_host_alloc(TsbCtx* ctx, uint64_t nbytes, void** hptr) {
  CHECK_CTX(ctx);
  if (!hptr) return set_err(TSB_EINVAL, "null argument");
  uint64_t a = 0;
  int rc = ctx->host_arena->alloc(nbytes, &a);
  if (rc == TSB_EAGAIN) {
    reap_scratch(ctx, false);
    rc = ctx->host_arena->alloc(nbytes, &a);
  }
  *hptr = (void*)(uintptr_t)a;
  return rc;
}

int tsb_host_alloc_nogrow(TsbCtx* ctx, uint64_t nbytes, void** hptr) {
  CHECK_CTX(ctx);
  if (!hptr) return set_err(TSB_EINVAL, "null argument");
  uint64_t a = 0;
  int rc = ctx->host_arena->alloc(nbytes, &a, false);
  *hptr = (void*)(uintptr_t)a;
  return rc;
}

int tsb_host_free(TsbCtx* ctx, void* hptr) {
  CHECK_CTX(ctx);
  return ctx->host_arena->free((uint64_t)(uintptr_t)hptr);
}

int tsb_arena_stats(TsbCtx* ctx, int which, uint64_t* in_use, uint64_t* reserved, uint64_t* bound) {
  if (!ctx || !in_use || !reserved || !bound) return set_err(TSB_EINVAL, "null argument");
  (which == 0 ? ctx->host_arena : ctx->dev_arena)->stats(in_use, reserved, bound);
  return TSB_OK;
}

int tsb_arena_trim(TsbCtx* ctx) {
  CHECK_CTX(ctx);
  reap_scratch(ctx, false);
  ctx->host_arena->trim();
  ctx->dev_arena->trim();
  ctx->scratch_host->trim();
  ctx->scratch_dev->trim();
  return TSB_OK;
}

int tsb_event_record(TsbCtx* ctx, int stream, TsbEvent* ev) {
  CHECK_CTX(ctx);
  CHECK_STREAM(stream);
  if (!ev) return set_err(TSB_EINVAL, "null argument");
  uint32_t idx = 0;
  cudaEvent_t e = nullptr;
  {
    std::lock_guard<std::mutex> g(ctx->mu);
    if (!ctx->free_events.empty()) {
      idx = ctx->free_events.back();
      ctx->free_events.pop_back();
    } else {
      cudaEvent_t ne;
      cudaError_t ce = cudaEventCreate(&ne);
      if (ce != cudaSuccess) return set_err(TSB_ECUDA, "cudaEventCreate failed: %s", cudaGetErrorString(ce));
      ctx->events.push_back(ne);
      ctx->event_gen.push_back(1);
      idx = (uint32_t)ctx->events.size();
    }
    e = ctx->events[idx - 1];
    *ev = (uint64_t)idx | ((uint64_t)ctx->event_gen[idx - 1] << 32);
  }
  cudaError_t re = cudaEventRecord(e, ctx->streams[stream]);
  if (re != cudaSuccess) {
    tsb_event_release(ctx, *ev);  // the slot goes back to the free list; the handle handed out is stale now
    *ev = 0;
    return set_err(TSB_ECUDA, "cudaEventRecord failed: %s", cudaGetErrorString(re));
  }
  return TSB_OK;
}

int tsb_stream_wait_event(TsbCtx* ctx, int stream, TsbEvent ev) {
  CHECK_CTX(ctx);
  CHECK_STREAM(stream);
  cudaEvent_t e = nullptr;
  int rc = event_lookup(ctx, ev, &e);
  if (rc != TSB_OK) return rc;
  CUDA_TRY(cudaStreamWaitEvent(ctx->streams[stream], e, 0));
  return TSB_OK;
}

int tsb_event_query(TsbCtx* ctx, TsbEvent ev) {
  CHECK_CTX(ctx);
  cudaEvent_t e = nullptr;
  int rc = event_lookup(ctx, ev, &e);
  if (rc != TSB_OK) return rc;
  cudaError_t q = cudaEventQuery(e);
  if (q == cudaSuccess) return TSB_OK;
  if (q == cudaErrorNotReady) {
    (void)cudaGetLastError();
    return TSB_EAGAIN;
  }
  return set_err(TSB_ECUDA, "cudaEventQuery failed: %s", cudaGetErrorString(q));
}

int tsb_event_sync(TsbCtx* ctx, TsbEvent ev) {
  CHECK_CTX(ctx);
  cudaEvent_t e = nullptr;
  int rc = event_lookup(ctx, ev, &e);
  if (rc != TSB_OK) return rc;
  CUDA_TRY(cudaEventSynchronize(e));
  return TSB_OK;
}

int tsb_event_elapsed_ms(TsbCtx* ctx, TsbEvent start, TsbEvent stop, float* ms) {
  CHECK_CTX(ctx);
  if (!ms) return set_err(TSB_EINVAL, "null argument");
  cudaEvent_t a = nullptr, b = nullptr;
  int rc = event_lookup(ctx, start, &a);
  if (rc != TSB_OK) return rc;
  rc = event_lookup(ctx, stop, &b);
  if (rc != TSB_OK) return rc;
  CUDA_TRY(cudaEventElapsedTime(ms, a, b));
  return TSB_OK;
}

int tsb_event_release(TsbCtx* ctx, TsbEvent ev) {
  if (!ctx) return set_err(TSB_EINVAL, "null context");
  const uint32_t idx = (uint32_t)(ev & 0xffffffffu), gen = (uint32_t)(ev >> 32);
  std::lock_guard<std::mutex> g(ctx->mu);
  if (idx == 0 || idx > ctx->events.size() || ctx->event_gen[idx - 1] != gen)
    return set_err(TSB_EINVAL, "stale or invalid event handle 0x%llx", (unsigned long long)ev);
  ctx->event_gen[idx - 1] += 1;
  if (ctx->event_gen[idx - 1] == 0) ctx->event_gen[idx - 1] = 1;
  ctx->free_events.push_back(idx);
  return TSB_OK;
}

int tsb_stream_wait_external(TsbCtx* ctx, int stream, uint64_t cuda_stream) {
  CHECK_CTX(ctx);
  CHECK_STREAM(stream);
  cudaEvent_t e = nullptr;
  CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  cudaError_t r = cudaEventRecord(e, (cudaStream_t)(uintptr_t)cuda_stream);
  if (r == cudaSuccess) r = cudaStreamWaitEvent(ctx->streams[stream], e, 0);
  cudaEventDestroy(e);  // destruction is deferred by the runtime until the event completes
  if (r != cudaSuccess) return set_err(TSB_ECUDA, "stream_wait_external failed: %s", cudaGetErrorString(r));
  return TSB_OK;
}

int tsb_external_wait_stream(TsbCtx* ctx, uint64_t cuda_stream, int stream) {
  CHECK_CTX(ctx);
  CHECK_STREAM(stream);
  cudaEvent_t e = nullptr;
  CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  cudaError_t r = cudaEventRecord(e, ctx->streams[stream]);
  if (r == cudaSuccess) r = cudaStreamWaitEvent((cudaStream_t)(uintptr_t)cuda_stream, e, 0);
  cudaEventDestroy(e);
  if (r != cudaSuccess) return set_err(TSB_ECUDA, "external_wait_stream failed: %s", cudaGetErrorString(r));
  return TSB_OK;
}

int tsb_stream_sync(TsbCtx* ctx, int stream) {
  CHECK_CTX(ctx);
  CHECK_STREAM(stream);
  CUDA_TRY(cudaStreamSynchronize(ctx->streams[stream]));
  return TSB_OK;
}

static int copy_regions_impl(TsbCtx* ctx, int stream, const TsbRegion* regions, uint32_t n, const TsbEvent* kernel_start,
                             const TsbEvent* kernel_stop) {
  CHECK_CTX(ctx);
  CHECK_STREAM(stream);
  if (n == 0) return TSB_OK;
  if (!regions) return set_err(TSB_EINVAL, "null regions");
  reap_scratch(ctx, false);
  TilePlan plan;
  plan.flat_runs.reserve(n);
  int rc = build_tiles(regions, n, plan);
  if (rc != TSB_OK) return rc;
  if (plan.flat_runs.empty() && plan.strided.empty() && plan.transposed.empty()) return TSB_OK;
  if (plan.flat_tiles > 0x7fffffffull || plan.strided.size() > 0x7fffffffull || plan.transposed.size() > 0x7fffffffull)
    return set_err(TSB_EINVAL, "too many tiles in one launch");
  // uploaded scratch: [flat runs][strided tiles][transpose tiles][regions of both]; device-only: [flat tiles]
  const uint64_t run_bytes = (uint64_t)plan.flat_runs.size() * sizeof(TsbRun);
  const uint64_t str_bytes = (uint64_t)plan.strided.size() * sizeof(TsbTile);
  const uint64_t trn_bytes = (uint64_t)plan.transposed.size() * sizeof(TsbTile);
  const uint64_t reg_bytes = (uint64_t)plan.strided_regions.size() * sizeof(TsbRegion);
  const uint64_t upload = run_bytes + str_bytes + trn_bytes + reg_bytes;
  const uint64_t tile_bytes = plan.flat_tiles * sizeof(TsbTile);
  uint64_t hbuf = 0, dbuf = 0;
  rc = ctx->scratch_host->alloc(upload, &hbuf);
  if (rc == TSB_EAGAIN) {  // tables of earlier launches still pending: wait for them, then retry once
    reap_scratch(ctx, true);
    rc = ctx->scratch_host->alloc(upload, &hbuf);
  }
  if (rc != TSB_OK) return rc;
  rc = ctx->scratch_dev->alloc(upload + tile_bytes, &dbuf);
  if (rc == TSB_EAGAIN) {
    reap_scratch(ctx, true);
    rc = ctx->scratch_dev->alloc(upload + tile_bytes, &dbuf);
  }
  if (rc != TSB_OK) {
    ctx->scratch_host->free(hbuf);
    return rc;
  }
  char* h = (char*)(uintptr_t)hbuf;
  if (run_bytes) memcpy(h, plan.flat_runs.data(), run_bytes);
  if (str_bytes) memcpy(h + run_bytes, plan.strided.data(), str_bytes);
  if (trn_bytes) memcpy(h + run_bytes + str_bytes, plan.transposed.data(), trn_bytes);
  if (reg_bytes) memcpy(h + run_bytes + str_bytes + trn_bytes, plan.strided_regions.data(), reg_bytes);
  cudaStream_t s = ctx->streams[stream];
  cudaError_t e = cudaMemcpyAsync((void*)(uintptr_t)dbuf, h, upload, cudaMemcpyHostToDevice, s);
  cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
  if (e == cudaSuccess && plan.flat_tiles) {
    const uint32_t nt = (uint32_t)plan.flat_tiles;
    TsbTile* tiles = (TsbTile*)(uintptr_t)(dbuf + upload);
    tsb_expand_flat_kernel<<<(nt + 255) / 256, 256, 0, s>>>((const TsbRun*)(uintptr_t)dbuf,
                                                            (uint32_t)plan.flat_runs.size(), tiles, nt);
    if (kernel_start) {
      if (event_lookup(ctx, *kernel_start, &ev_start) == TSB_OK) cudaEventRecord(ev_start, s);
    }
    if (ctx->use_tma)
      tsb_copy_flat_tma_kernel<<<nt, TSB_FLAT_THREADS, TSB_TILE_BYTES, s>>>(tiles);
    else
      tsb_copy_flat_kernel<<<nt, TSB_FLAT_THREADS, 0, s>>>(tiles);
    e = cudaGetLastError();
  } else if (kernel_start && event_lookup(ctx, *kernel_start, &ev_start) == TSB_OK) {
    cudaEventRecord(ev_start, s);
  }
  const TsbRegion* d_regions = (const TsbRegion*)(uintptr_t)(dbuf + run_bytes + str_bytes + trn_bytes);
  if (e == cudaSuccess && !plan.strided.empty()) {
    tsb_copy_strided_kernel<<<(uint32_t)plan.strided.size(), TSB_STRIDED_THREADS, 0, s>>>(
        d_regions, (const TsbTile*)(uintptr_t)(dbuf + run_bytes), warp_run_min());
    e = cudaGetLastError();
  }
  if (e == cudaSuccess && !plan.transposed.empty()) {
    tsb_copy_transpose_kernel<<<(uint32_t)plan.transposed.size(), TSB_TRANSPOSE_THREADS, 0, s>>>(
        d_regions, (const TsbTile*)(uintptr_t)(dbuf + run_bytes + str_bytes));
    e = cudaGetLastError();
  }
  if (kernel_stop && event_lookup(ctx, *kernel_stop, &ev_stop) == TSB_OK) cudaEventRecord(ev_stop, s);
  {
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->n_copy_calls += 1;
    ctx->n_kernel_launches += (plan.flat_tiles ? 2 : 0) + (plan.strided.empty() ? 0 : 1) + (plan.transposed.empty() ? 0 : 1);
    ctx->desc_upload_bytes += upload;
    ctx->kernel_payload_bytes += plan.payload;
  }
  // the scratch is released once the launch has completed
  PendingScratch p;
  p.hbuf = hbuf;
  p.dbuf = dbuf;
  p.done = nullptr;
  {
    std::lock_guard<std::mutex> g(ctx->mu);
    if (!ctx->spare_scratch_events.empty()) {
      p.done = ctx->spare_scratch_events.back();
      ctx->spare_scratch_events.pop_back();
    }
  }
  if (p.done == nullptr && cudaEventCreateWithFlags(&p.done, cudaEventDisableTiming) != cudaSuccess) {
    cudaStreamSynchronize(s);
    ctx->scratch_host->free(hbuf);
    ctx->scratch_dev->free(dbuf);
    return set_err(TSB_ECUDA, "cudaEventCreate failed");
  }
  cudaEventRecord(p.done, s);
  {
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->pending.push_back(p);
  }
  if (e != cudaSuccess) return set_err(TSB_ECUDA, "copy_regions launch failed: %s", cudaGetErrorString(e));
  return TSB_OK;
}

int tsb_copy_regions(TsbCtx* ctx, int stream, const TsbRegion* regions, uint32_t n) {
  return copy_regions_impl(ctx, stream, regions, n, nullptr, nullptr);
}

int tsb_copy_regions_timed(TsbCtx* ctx, int stream, const TsbRegion* regions, uint32_t n, TsbEvent* kernel_start,
                           TsbEvent* kernel_stop) {
  if (!ctx || !kernel_start || !kernel_stop) return set_err(TSB_EINVAL, "null argument");
  // allocate the two event handles first (recorded on an idle point of the stream, then re-recorded
  // around the copy kernels inside the launch)
  int rc = tsb_event_record(ctx, stream, kernel_start);
  if (rc != TSB_OK) return rc;
  rc = tsb_event_record(ctx, stream, kernel_stop);
  if (rc != TSB_OK) {
    tsb_event_release(ctx, *kernel_start);
    *kernel_start = 0;
    return rc;
  }
  rc = copy_regions_impl(ctx, stream, regions, n, kernel_start, kernel_stop);
  if (rc != TSB_OK) {  // the caller gets no handles on failure, so they must not stay allocated
    char keep[sizeof(g_err)];
    memcpy(keep, g_err, sizeof(keep));
    tsb_event_release(ctx, *kernel_start);
    tsb_event_release(ctx, *kernel_stop);
    memcpy(g_err, keep, sizeof(keep));
    *kernel_start = *kernel_stop = 0;
  }
  return rc;
}

int tsb_memcpy_d2h(TsbCtx* ctx, int stream, void* hptr, uint64_t dptr, uint64_t nbytes) {
  CHECK_CTX(ctx);
  CHECK_STREAM(stream);
  if (nbytes == 0) return TSB_OK;
  CUDA_TRY(cudaMemcpyAsync(hptr, (const void*)(uintptr_t)dptr, nbytes, cudaMemcpyDeviceToHost, ctx->streams[stream]));
  return TSB_OK;
}

int tsb_memcpy_d2h_interleaved(TsbCtx* ctx, int stream, uint32_t n, const uint64_t* hptrs, const uint64_t* dptrs,
                               const uint64_t* nbytes, uint64_t chunk, TsbEvent* events) {
  CHECK_CTX(ctx);
  CHECK_STREAM(stream);
  if (n == 0) return TSB_OK;
  if (!hptrs || !dptrs || !nbytes || !events || chunk == 0) return set_err(TSB_EINVAL, "null argument");
  std::vector<uint64_t> pos(n, 0), slot(n, 0);
  uint64_t total = 0;
  for (uint32_t b = 0; b < n; ++b) {
    slot[b] = total;
    total += (nbytes[b] + chunk - 1) / chunk;
  }
  uint64_t recorded = 0;
  int rc = TSB_OK;
  for (bool left = true; left && rc == TSB_OK;) {
    left = false;
    for (uint32_t b = 0; b < n && rc == TSB_OK; ++b) {
      if (pos[b] >= nbytes[b]) continue;
      const uint64_t c = std::min(chunk, nbytes[b] - pos[b]);
      cudaError_t e = cudaMemcpyAsync((void*)(uintptr_t)(hptrs[b] + pos[b]), (const void*)(uintptr_t)(dptrs[b] + pos[b]), c,
                                      cudaMemcpyDeviceToHost, ctx->streams[stream]);
      if (e != cudaSuccess) {
        rc = set_err(TSB_ECUDA, "cudaMemcpyAsync failed: %s", cudaGetErrorString(e));
        break;
      }
      pos[b] += c;
      rc = tsb_event_record(ctx, stream, &events[slot[b]++]);
      if (rc == TSB_OK) ++recorded;
      left = true;
    }
  }
  if (rc != TSB_OK) {  // no handle survives a failed call
    char keep[sizeof(g_err)];
    memcpy(keep, g_err, sizeof(keep));
    uint64_t base = 0;
    for (uint32_t b = 0; b < n; ++b) {
      const uint64_t cnt = (nbytes[b] + chunk - 1) / chunk;
      for (uint64_t j = base; j < slot[b] && j < base + cnt; ++j)
        if (events[j]) tsb_event_release(ctx, events[j]);
      base += cnt;
    }
    memcpy(g_err, keep, sizeof(keep));
  }
  (void)recorded;
  return rc;
}

int tsb_memcpy_h2d(TsbCtx* ctx, int stream, uint64_t dptr, const void* hptr, uint64_t nbytes) {
  CHECK_CTX(ctx);
  CHECK_STREAM(stream);
  if (nbytes == 0) return TSB_OK;
  CUDA_TRY(cudaMemcpyAsync((void*)(uintptr_t)dptr, hptr, nbytes, cudaMemcpyHostToDevice, ctx->streams[stream]));
  return TSB_OK;
}

int tsb_memcpy_d2d(TsbCtx* ctx, int stream, uint64_t dst, uint64_t src, uint64_t nbytes) {
  CHECK_CTX(ctx);
  CHECK_STREAM(stream);
  if (nbytes == 0) return TSB_OK;
  CUDA_TRY(cudaMemcpyAsync((void*)(uintptr_t)dst, (const void*)(uintptr_t)src, nbytes, cudaMemcpyDefault,
                           ctx->streams[stream]));
  return TSB_OK;
}

// ---- copy without any allocation ----------------------------------------------------------------
//
// tsb_copy_regions needs scratch memory beyond its inputs and outputs: the descriptor tables it
// uploads and the tile table the expansion kernel writes.  It takes that scratch from the context's
// arenas, which may have to call cudaMalloc / cudaHostAlloc.  Some callers must not do that: a
// process that has already handed every byte of HBM to its own memory pool (a cudaMalloc there fails
// or steals from the pool), a hook that runs while another part of the program is in a phase where
// the implicit device-wide synchronisation of cudaMalloc / cudaFree would stall its streams, or an
// error / shutdown path that wants to push out a last copy after an out-of-memory condition.
// For them the same kernels run out of the fixed array below.  The array exists once per GPU that
// loads this library and stays for the life of the process, which is why it is small.
//
// Layout of the array:  [0, TSB_STATIC_UPLOAD_BYTES)   flat runs | strided tiles | strided regions
//                       [TSB_STATIC_UPLOAD_BYTES, end)  flat tile table written by the expansion kernel
// Nothing in it has to start at zero: every byte a kernel reads is written earlier in the same call
// (upload, then expansion), so the leftovers of the previous call are never observed.
static __device__ __align__(256) unsigned char g_static_scratch[TSB_STATIC_SCRATCH_BYTES];

#define TSB_STATIC_MAX_DEVICES 64
// `g_static_scratch` in host code is only a handle for the runtime; the address a kernel can use is
// looked up once per device and kept here.
static void* g_static_scratch_addr[TSB_STATIC_MAX_DEVICES];
static bool g_static_scratch_ready[TSB_STATIC_MAX_DEVICES];
static std::mutex g_static_mu;
// host staging for the upload: fixed storage as well, protected by g_static_mu
static unsigned char g_static_upload[TSB_STATIC_UPLOAD_BYTES] __attribute__((aligned(64)));

extern "C" int tsb_copy_regions_static(int device, uint64_t cuda_stream, const TsbRegion* regions, uint32_t n) {
  if (device < 0 || device >= TSB_STATIC_MAX_DEVICES) return set_err(TSB_EINVAL, "device %d out of range", device);
  if (n == 0) return TSB_OK;
  if (!regions) return set_err(TSB_EINVAL, "null regions");
  DeviceGuard guard(device);
  if (!guard.ok) return set_err(TSB_ECUDA, "cudaSetDevice(%d) failed", device);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_scratch_ready[device]) {
    // set-up, once per device, with that device current
    void* addr = nullptr;
    cudaError_t se = cudaGetSymbolAddress(&addr, g_static_scratch);
    if (se != cudaSuccess || addr == nullptr)
      return set_err(TSB_ECUDA, "cudaGetSymbolAddress(static scratch) failed on device %d: %s", device,
                     cudaGetErrorString(se));
    g_static_scratch_addr[device] = addr;
    g_static_scratch_ready[device] = true;
  }
  unsigned char* const dscratch = static_cast<unsigned char*>(g_static_scratch_addr[device]);
  {
    // The descriptor upload below reads a host buffer that the next call overwrites, so a captured copy of
    // this call would replay with whatever descriptors a later call left there: refuse instead.
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaError_t ce = cudaStreamIsCapturing((cudaStream_t)(uintptr_t)cuda_stream, &cap);
    if (ce != cudaSuccess) {
      (void)cudaGetLastError();
      return set_err(TSB_ECUDA, "cudaStreamIsCapturing failed: %s", cudaGetErrorString(ce));
    }
    if (cap != cudaStreamCaptureStatusNone)
      return set_err(TSB_EINVAL, "tsb_copy_regions_static: the stream is being captured; call it outside the capture");
  }

  // pass 1: count, so that an oversized request is refused before anything is written
  uint64_t n_runs = 0, n_flat_tiles = 0, n_str_tiles = 0, n_str_regions = 0;
  for (uint32_t i = 0; i < n; ++i) {
    const TsbRegion& r = regions[i];
    int rc = validate_region(r, i);
    if (rc != TSB_OK) return rc;
    uint64_t total;
    if (!region_payload(r, &total)) return set_err(TSB_EINVAL, "region %u: size overflow", i);
    if (total == 0) continue;
    if (r.ndim == 0) {
      n_runs += 1;
      n_flat_tiles += tsb_run_tiles(r.dst, r.run_bytes);
    } else if (r.run_bytes >= TSB_FLAT_RUN_MIN) {
      const uint64_t runs = total / r.run_bytes;
      n_runs += runs;
      // every run of the box: upper bound of its tiles (the exact count depends on each run's alignment)
      n_flat_tiles += runs * (2 + r.run_bytes / TSB_TILE_BYTES);
    } else {
      uint64_t per = TSB_TILE_BYTES / r.run_bytes;
      if (per == 0) per = 1;
      const uint64_t step = per * r.run_bytes;
      n_str_regions += 1;
      n_str_tiles += (total + step - 1) / step;
    }
    if (n_runs > TSB_STATIC_UPLOAD_BYTES || n_flat_tiles > (1ull << 40)) break;  // far beyond any limit already
  }
  const uint64_t upload_need = n_runs * sizeof(TsbRun) + n_str_tiles * sizeof(TsbTile) + n_str_regions * sizeof(TsbRegion);
  if (upload_need > TSB_STATIC_UPLOAD_BYTES)
    return set_err(TSB_EINVAL,
                   "tsb_copy_regions_static: descriptors need %llu bytes, the static scratch holds %llu "
                   "(use tsb_copy_regions, or split the call)",
                   (unsigned long long)upload_need, (unsigned long long)TSB_STATIC_UPLOAD_BYTES);
  if (n_flat_tiles > TSB_STATIC_MAX_TILES)
    return set_err(TSB_EINVAL,
                   "tsb_copy_regions_static: %llu tiles of %u bytes needed, the static scratch holds %llu "
                   "(about %llu MiB of payload per call; use tsb_copy_regions, or split the call)",
                   (unsigned long long)n_flat_tiles, (unsigned)TSB_TILE_BYTES, (unsigned long long)TSB_STATIC_MAX_TILES,
                   (unsigned long long)(TSB_STATIC_MAX_TILES * (uint64_t)TSB_TILE_BYTES >> 20));

  // pass 2: fill the fixed host buffer: [runs][strided tiles][strided regions]
  TsbRun* runs = reinterpret_cast<TsbRun*>(g_static_upload);
  TsbTile* stiles = reinterpret_cast<TsbTile*>(g_static_upload + n_runs * sizeof(TsbRun));
  TsbRegion* sregs = reinterpret_cast<TsbRegion*>(g_static_upload + n_runs * sizeof(TsbRun) + n_str_tiles * sizeof(TsbTile));
  uint64_t ri = 0, ti = 0, gi = 0, tiles = 0;
  for (uint32_t i = 0; i < n; ++i) {
    const TsbRegion& r = regions[i];
    uint64_t total = 0;
    region_payload(r, &total);
    if (total == 0) continue;
    if (r.ndim == 0) {
      runs[ri++] = TsbRun{r.src, r.dst, r.run_bytes, tiles};
      tiles += tsb_run_tiles(r.dst, r.run_bytes);
    } else if (r.run_bytes >= TSB_FLAT_RUN_MIN) {
      uint64_t idx[TSB_MAX_DIMS] = {0, 0, 0, 0};
      const uint64_t nr = total / r.run_bytes;
      for (uint64_t q = 0; q < nr; ++q) {
        int64_t so = 0, dofs = 0;
        for (uint32_t d = 0; d < r.ndim; ++d) {
          so += (int64_t)idx[d] * r.src_stride[d];
          dofs += (int64_t)idx[d] * r.dst_stride[d];
        }
        runs[ri++] = TsbRun{r.src + so, r.dst + dofs, r.run_bytes, tiles};
        tiles += tsb_run_tiles(r.dst + dofs, r.run_bytes);
        for (int d = (int)r.ndim - 1; d >= 0; --d) {
          if (++idx[d] < r.size[d]) break;
          idx[d] = 0;
        }
      }
    } else {
      uint64_t per = TSB_TILE_BYTES / r.run_bytes;
      if (per == 0) per = 1;
      const uint64_t step = per * r.run_bytes;
      for (uint64_t off = 0; off < total; off += step) {
        TsbTile t;
        t.src = 0;
        t.dst = 0;
        t.off = off;
        t.len = (uint32_t)std::min<uint64_t>(step, total - off);
        t.region = (uint32_t)gi;
        stiles[ti++] = t;
      }
      sregs[gi] = r;
      sregs[gi++].reserved = 0;
    }
  }
  if (ri == 0 && ti == 0) return TSB_OK;
  cudaStream_t s = (cudaStream_t)(uintptr_t)cuda_stream;
  // pageable source: the runtime first waits for the work already queued on the stream, then copies
  // g_static_upload out before it returns -- so this call blocks the host for that long, and the buffer can
  // be reused by the next call as soon as the mutex is released
  cudaError_t e = cudaMemcpyAsync(dscratch, g_static_upload, upload_need, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess && tiles) {
    TsbTile* dtiles = reinterpret_cast<TsbTile*>(dscratch + TSB_STATIC_UPLOAD_BYTES);
    tsb_expand_flat_kernel<<<((uint32_t)tiles + 255) / 256, 256, 0, s>>>(reinterpret_cast<const TsbRun*>(dscratch),
                                                                        (uint32_t)ri, dtiles, (uint32_t)tiles);
    tsb_copy_flat_kernel<<<(uint32_t)tiles, TSB_FLAT_THREADS, 0, s>>>(dtiles);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess && ti) {
    tsb_copy_strided_kernel<<<(uint32_t)ti, TSB_STRIDED_THREADS, 0, s>>>(
        reinterpret_cast<const TsbRegion*>(dscratch + ri * sizeof(TsbRun) + ti * sizeof(TsbTile)),
        reinterpret_cast<const TsbTile*>(dscratch + ri * sizeof(TsbRun)), warp_run_min());
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) return set_err(TSB_ECUDA, "tsb_copy_regions_static launch failed: %s", cudaGetErrorString(e));
  return TSB_OK;
}

// ---- host helpers ---------------------------------------------------------------------------
static int default_threads(uint64_t nbytes, int nthreads) {
  if (nthreads <= 0) {
    unsigned hc = std::thread::hardware_concurrency();
    nthreads = (int)std::min<unsigned>(hc ? hc : 4, 8);
  }
  const uint64_t min_chunk = 4ull << 20;
  uint64_t by_size = (nbytes + min_chunk - 1) / min_chunk;
  if (by_size < 1) by_size = 1;
  return (int)std::min<uint64_t>((uint64_t)nthreads, by_size);
}

int tsb_host_memcpy(void* dst, const void* src, uint64_t nbytes, int nthreads) {
  if (nbytes == 0) return TSB_OK;
  if (!dst || !src) return set_err(TSB_EINVAL, "null argument");
  const int nt = default_threads(nbytes, nthreads);
  if (nt <= 1) {
    memcpy(dst, src, nbytes);
    return TSB_OK;
  }
  std::vector<std::thread> th;
  const uint64_t chunk = ((nbytes + nt - 1) / nt + 4095) / 4096 * 4096;
  for (int i = 0; i < nt; ++i) {
    const uint64_t lo = (uint64_t)i * chunk;
    if (lo >= nbytes) break;
    const uint64_t len = std::min(chunk, nbytes - lo);
    th.emplace_back([=] { memcpy((char*)dst + lo, (const char*)src + lo, len); });
  }
  for (auto& t : th) t.join();
  return TSB_OK;
}

// Many independent host copies (the members of a CPU slab) spread over worker threads by BYTES: the flat
// byte range [0, total) of all copies is cut into equal shares, so one large member is split and many small
// ones are grouped.
int tsb_host_gather(const uint64_t* dst, const uint64_t* src, const uint64_t* nbytes, uint64_t count, int nthreads) {
  if (count == 0) return TSB_OK;
  if (!dst || !src || !nbytes) return set_err(TSB_EINVAL, "null argument");
  uint64_t total = 0;
  for (uint64_t i = 0; i < count; ++i) {
    if (nbytes[i] && (!dst[i] || !src[i])) return set_err(TSB_EINVAL, "copy %llu: null address", (unsigned long long)i);
    total += nbytes[i];
  }
  if (total == 0) return TSB_OK;
  const int nt = default_threads(total, nthreads);
  auto work = [=](uint64_t lo, uint64_t hi) {  // copies the bytes [lo, hi) of the flat range
    uint64_t pos = 0;
    for (uint64_t i = 0; i < count && pos < hi; ++i) {
      const uint64_t n = nbytes[i];
      const uint64_t a = std::max(lo, pos), b = std::min(hi, pos + n);
      if (a < b) memcpy((char*)(uintptr_t)dst[i] + (a - pos), (const char*)(uintptr_t)src[i] + (a - pos), b - a);
      pos += n;
    }
  };
  if (nt <= 1) {
    work(0, total);
    return TSB_OK;
  }
  const uint64_t share = ((total + nt - 1) / nt + 63) / 64 * 64;
  std::vector<std::thread> th;
  for (int t = 0; t < nt; ++t) {
    const uint64_t lo = (uint64_t)t * share;
    if (lo >= total) break;
    th.emplace_back(work, lo, std::min(total, lo + share));
  }
  for (auto& t : th) t.join();
  return TSB_OK;
}

static int rw_range(int fd, bool write, char* buf, uint64_t off, uint64_t len) {
  while (len) {
    ssize_t k = write ? pwrite(fd, buf, len, (off_t)off) : pread(fd, buf, len, (off_t)off);
    if (k < 0) {
      if (errno == EINTR) continue;
      return errno;
    }
    if (k == 0) return write ? EIO : ENODATA;
    buf += k;
    off += (uint64_t)k;
    len -= (uint64_t)k;
  }
  return 0;
}

// One file operation split over `nt` threads at 4 KiB-multiple boundaries.
static int rw_parallel(int fd, bool write, char* buf, uint64_t off, uint64_t nbytes, int nt) {
  if (nbytes == 0) return 0;
  if (nt <= 1) return rw_range(fd, write, buf, off, nbytes);
  std::vector<std::thread> th;
  std::vector<int> errs(nt, 0);
  const uint64_t chunk = ((nbytes + nt - 1) / nt + 4095) / 4096 * 4096;
  for (int i = 0; i < nt; ++i) {
    const uint64_t lo = (uint64_t)i * chunk;
    if (lo >= nbytes) break;
    const uint64_t len = std::min(chunk, nbytes - lo);
    int* e = &errs[i];
    th.emplace_back([=] { *e = rw_range(fd, write, buf + lo, off + lo, len); });
  }
  for (auto& t : th) t.join();
  for (int e : errs)
    if (e) return e;
  return 0;
}

static int file_rw(const char* path, bool write, uint64_t offset, void* hptr, uint64_t nbytes, int nthreads,
                   uint32_t flags) {
  if (!path) return set_err(TSB_EINVAL, "null path");
  if (flags & ~(uint32_t)TSB_FILE_DIRECT) return set_err(TSB_EINVAL, "unknown file flags 0x%x", flags);
  const int base_flags = write ? (O_WRONLY | O_CREAT | O_TRUNC) : O_RDONLY;
  int fd = open(path, base_flags, 0644);
  if (fd < 0) return set_err(TSB_EIO, "open(%s) failed: %s", path, strerror(errno));
  int err = 0;
  const int nt = default_threads(nbytes, nthreads);
  // TSB_FILE_DIRECT: the 4 KiB-aligned body bypasses the page cache through a second, O_DIRECT descriptor
  // (pinned arena blocks are 4 KiB aligned); an unaligned request, the sub-page tail, and file systems
  // that refuse O_DIRECT take the buffered descriptor.
  uint64_t direct_len = 0;
  if ((flags & TSB_FILE_DIRECT) && nbytes >= 4096 && offset % 4096 == 0 && (uintptr_t)hptr % 4096 == 0) {
    int dfd = open(path, (write ? O_WRONLY : O_RDONLY) | O_DIRECT);
    if (dfd >= 0) {
      direct_len = nbytes & ~4095ull;
      err = rw_parallel(dfd, write, (char*)hptr, offset, direct_len, nt);
      if (err == EINVAL) {  // the file system accepted the open but not the transfer: redo it buffered
        err = 0;
        direct_len = 0;
      }
      if (close(dfd) != 0 && !err) err = errno;
    }
  }
  if (!err) err = rw_parallel(fd, write, (char*)hptr + direct_len, offset + direct_len, nbytes - direct_len, nt);
  if (close(fd) != 0 && !err) err = errno;
  if (err) return set_err(TSB_EIO, "%s(%s) failed: %s", write ? "write" : "read", path, strerror(err));
  return TSB_OK;
}

int tsb_file_write_streamed(TsbCtx* ctx, const char* path, const void* hptr, uint64_t nbytes, const TsbEvent* events,
                            const uint64_t* ends, uint32_t n) {
  if (!ctx || !path || (nbytes && !hptr) || (n && (!events || !ends))) return set_err(TSB_EINVAL, "null argument");
  DeviceGuard guard(ctx->device);
  int fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
  if (fd < 0) return set_err(TSB_EIO, "open(%s) failed: %s", path, strerror(errno));
  uint64_t done = 0;
  int err = 0, rc = TSB_OK;
  for (uint32_t i = 0; i < n && !err && rc == TSB_OK; ++i) {
    if (ends[i] <= done) continue;
    cudaEvent_t e = nullptr;
    rc = event_lookup(ctx, events[i], &e);
    if (rc != TSB_OK) break;
    // Poll, asleep in between: up to 32 writer threads follow the copies at once, and spinning in
    // cudaEventSynchronize they took the cores away from the thread that waits for the pack kernel (async_take
    // blocked 65 ms instead of 13 on a 16-core host); blocking-sync events woke up late by the length of the
    // whole copy queue.  A chunk takes ~0.6 ms to arrive, so 100 us of extra latency costs nothing.
    cudaError_t ce;
    while ((ce = cudaEventQuery(e)) == cudaErrorNotReady) {
      (void)cudaGetLastError();
      usleep(100);
    }
    if (ce != cudaSuccess) {
      rc = set_err(TSB_ECUDA, "cudaEventQuery failed: %s", cudaGetErrorString(ce));
      break;
    }
    uint64_t upto = std::min(ends[i], nbytes);
    // everything whose copy has finished by now goes out in one write
    while (i + 1 < n && ends[i + 1] <= nbytes) {
      cudaEvent_t nx = nullptr;
      if (event_lookup(ctx, events[i + 1], &nx) != TSB_OK || cudaEventQuery(nx) != cudaSuccess) {
        (void)cudaGetLastError();
        break;
      }
      upto = ends[++i];
    }
    err = rw_range(fd, true, (char*)const_cast<void*>(hptr) + done, done, upto - done);
    done = upto;
  }
  if (rc == TSB_OK && !err && done < nbytes) err = EIO;  // the events did not cover the buffer
  if (close(fd) != 0 && !err) err = errno;
  if (rc != TSB_OK) return rc;
  if (err) return set_err(TSB_EIO, "write(%s) failed: %s", path, strerror(err));
  return TSB_OK;
}

int tsb_file_write(const char* path, const void* hptr, uint64_t nbytes, int nthreads, uint32_t flags) {
  return file_rw(path, true, 0, const_cast<void*>(hptr), nbytes, nthreads, flags);
}

int tsb_file_read(const char* path, uint64_t offset, void* hptr, uint64_t nbytes, int nthreads, uint32_t flags) {
  return file_rw(path, false, offset, hptr, nbytes, nthreads, flags);
}

// ---- peer access ------------------------------------------------------------------------------
int tsb_ipc_export(TsbCtx* ctx, uint64_t dptr, uint8_t* handle, uint64_t* offset) {
  CHECK_CTX(ctx);
  if (!handle || !offset) return set_err(TSB_EINVAL, "null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == TSB_IPC_HANDLE_BYTES, "ipc handle size");
  uint64_t base = 0, size = 0;
  bool found = false;
  {
    std::lock_guard<std::mutex> g(ctx->mu);
    auto it = ctx->standalone.upper_bound(dptr);
    if (it != ctx->standalone.begin()) {
      --it;
      if (dptr < it->first + std::max<uint64_t>(it->second, 1)) {
        base = it->first;
        size = it->second;
        found = true;
      }
    }
  }
  if (!found && !ctx->dev_arena->find_segment(dptr, &base, &size))
    return set_err(TSB_EINVAL, "ipc_export: 0x%llx is not device-arena memory", (unsigned long long)dptr);
  cudaIpcMemHandle_t h;
  CUDA_TRY(cudaIpcGetMemHandle(&h, (void*)(uintptr_t)base));
  memcpy(handle, &h, sizeof(h));
  *offset = dptr - base;
  return TSB_OK;
}

int tsb_ipc_open(TsbCtx* ctx, const uint8_t* handle, uint64_t* base) {
  CHECK_CTX(ctx);
  if (!handle || !base) return set_err(TSB_EINVAL, "null argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *base = (uint64_t)(uintptr_t)p;
  return TSB_OK;
}

int tsb_ipc_close(TsbCtx* ctx, uint64_t base) {
  CHECK_CTX(ctx);
  CUDA_TRY(cudaIpcCloseMemHandle((void*)(uintptr_t)base));
  return TSB_OK;
}

int tsb_peer_enable(TsbCtx* ctx, int peer_device) {
  CHECK_CTX(ctx);
  if (peer_device == ctx->device) return TSB_OK;
  int can = 0;
  CUDA_TRY(cudaDeviceCanAccessPeer(&can, ctx->device, peer_device));
  if (!can) return set_err(TSB_ECUDA, "device %d cannot access peer %d", ctx->device, peer_device);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    (void)cudaGetLastError();
    return TSB_OK;
  }
  if (e != cudaSuccess) return set_err(TSB_ECUDA, "cudaDeviceEnablePeerAccess failed: %s", cudaGetErrorString(e));
  return TSB_OK;
}

}  // extern "C"
