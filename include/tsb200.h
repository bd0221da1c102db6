/*
 * tsb200.h -- C-ABI of the B200-native checkpoint hot path (libtsb200.so).
 *
 * This is the boundary a torchsnapshot maintainer binds (ctypes from Python; see
 * INTEGRATION.md).  It replaces the byte-moving calls of the reference's hot path,
 * all of which are PyTorch library calls in the reference (file:line relative to the
 * reference repository):
 *
 *   K1  torchsnapshot/batcher.py:144        torch.cuda.ByteTensor(slab)   -> tsb_dev_alloc
 *   K2  torchsnapshot/batcher.py:152-155    per-tensor D2D copy_ into slab -> tsb_copy_regions
 *   K3  torchsnapshot/batcher.py:156        gpu_buf.cpu() (pageable D2H)  -> tsb_memcpy_d2h (pinned)
 *   K4  torchsnapshot/io_preparers/tensor.py:252-254, 351-353  tensor.to("cpu") -> tsb_copy_regions + tsb_memcpy_d2h
 *   K5  torchsnapshot/serialization.py:182-185  host .contiguous()        -> strided TsbRegion in tsb_copy_regions
 *   K6  torchsnapshot/io_preparers/tensor.py:329-338, 356-358  dst.copy_(src) (pageable H2D) -> tsb_memcpy_h2d + tsb_copy_regions
 *   K7  torchsnapshot/io_preparers/sharded_tensor.py:278-284, 296-309  narrow+copy_ per overlap -> tsb_copy_regions (also over peer memory)
 *
 * Conventions
 *   - plain C types only; device and host addresses travel as uint64_t / void*.
 *   - every function returns TSB_OK (0), TSB_EAGAIN (1: an arena is full, retry after a
 *     release) or a negative error; tsb_last_error() gives the thread-local message.
 *   - all work is asynchronous on streams owned by the context; completion is observed
 *     through events.  No function synchronises the device unless its name says so.
 *   - one context per (process, device).  All functions are thread-safe.
 *   - nothing here falls back to the CPU: if no CUDA device is present, tsb_ctx_create fails.
 */
#ifndef TSB200_H_
#define TSB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TSB_OK 0
#define TSB_EAGAIN 1
#define TSB_EINVAL (-1)
#define TSB_ECUDA (-2)
#define TSB_ENOMEM (-3)
#define TSB_EIO (-4)

#define TSB_MAX_DIMS 4

/* Streams owned by a context. */
#define TSB_STREAM_PACK 0   /* gather-pack / scatter-unpack / reshard kernels */
#define TSB_STREAM_D2H 1    /* device -> pinned host copies */
#define TSB_STREAM_H2D 2    /* pinned host -> device copies */
#define TSB_STREAM_AUX 3    /* peer copies, misc */
#define TSB_NUM_STREAMS 4

/*
 * One strided box to copy: `run_bytes` contiguous bytes, repeated over up to TSB_MAX_DIMS
 * outer dimensions with independent byte strides on the source and destination side.
 * A contiguous tensor is {ndim = 0, run_bytes = numel * element_size}.
 *
 * Payload byte order is row-major over size[0..ndim) then the run -- exactly the order in
 * which tensor_as_memoryview (torchsnapshot/serialization.py:162-188) lays out a tensor
 * after .contiguous(), so a pack region with a contiguous destination reproduces the
 * reference's bytes; an unpack region is the same box with src/dst exchanged
 * (torchsnapshot/io_preparers/tensor.py:329-338), and a reshard region has both sides
 * strided (torchsnapshot/io_preparers/sharded_tensor.py:271-284).
 *
 * src/dst are absolute device addresses (any byte alignment).
 */
typedef struct TsbRegion {
  uint64_t src;
  uint64_t dst;
  uint64_t run_bytes;
  uint32_t ndim;
  uint32_t reserved;
  uint64_t size[TSB_MAX_DIMS];
  int64_t src_stride[TSB_MAX_DIMS];
  int64_t dst_stride[TSB_MAX_DIMS];
} TsbRegion; /* 128 bytes */

typedef struct TsbCtx TsbCtx;
typedef uint64_t TsbEvent; /* opaque handle, 0 is never valid */

/* ---- library / context ------------------------------------------------------------- */

/* ABI version of this header (major * 1000 + minor). */
int tsb_version(void);

/* Message for the last error on the calling thread ("" if none). */
const char* tsb_last_error(void);

/* Number of CUDA devices visible to the process. */
int tsb_device_count(int* count);

/*
 * Create the context for `device`.  `host_arena_max` / `dev_arena_max` bound the pinned
 * host arena and the device staging arena (bytes; 0 = library default).  Arenas grow on
 * demand in segments; nothing is pinned at creation.  Replaces the implicit use of the
 * current device/stream and the caching allocator (batcher.py:144).
 */
int tsb_ctx_create(int device, uint64_t host_arena_max, uint64_t dev_arena_max, TsbCtx** out);
int tsb_ctx_destroy(TsbCtx* ctx);
int tsb_ctx_device(TsbCtx* ctx, int* device);
/* Multiprocessor count of the context's device. */
int tsb_ctx_sm_count(TsbCtx* ctx, int* sms);
/*
 * Monotonic counters since context creation (any pointer may be NULL): kernels launched by
 * tsb_copy_regions (tile expansion + flat + strided), tsb_copy_regions calls that launched work,
 * descriptor bytes uploaded host->device, and payload bytes the kernels were asked to move.
 * No reference counterpart (bench.py reports them as gpu_launches / h2d bytes).
 */
int tsb_ctx_counters(TsbCtx* ctx, uint64_t* kernel_launches, uint64_t* copy_calls, uint64_t* desc_upload_bytes,
                     uint64_t* kernel_payload_bytes);

/* ---- arenas ------------------------------------------------------------------------ */

/*
 * Device staging memory (slabs).  Replaces torch.cuda.ByteTensor(slab_sz) (batcher.py:144).
 * TSB_EAGAIN when the arena is at its bound and fragmented/full: the caller retries after a
 * tsb_dev_free -- there is no CPU-slab fallback (the reference's is batcher.py:145-149).
 */
int tsb_dev_alloc(TsbCtx* ctx, uint64_t nbytes, uint64_t* dptr);
int tsb_dev_free(TsbCtx* ctx, uint64_t dptr);
/*
 * A device block that is a driver allocation of its own (not carved out of an arena segment), for
 * blocks a peer process maps with tsb_ipc_open: mapping a peer allocation costs 10-20 ms per GB of
 * the WHOLE allocation, so what a peer maps should hold only what that peer reads (the per-destination
 * buffers of a reshard, reshard.py).  Freed with tsb_dev_free (an implicit device synchronisation).
 * No counterpart in the reference, which never moves payload between ranks.
 */
int tsb_dev_alloc_exportable(TsbCtx* ctx, uint64_t nbytes, uint64_t* dptr);

/*
 * Pinned host memory the staged buffers live in (what BufferStager.stage_buffer returns a
 * memoryview of, io_types.py:19-26; the reference uses pageable memory, tensor.py:248-251).
 */
int tsb_host_alloc(TsbCtx* ctx, uint64_t nbytes, void** hptr);
/*
 * The same, but only out of segments the arena already holds: TSB_EAGAIN instead of pinning new memory
 * (first touch + cudaHostRegister cost ~40 ms per GiB).  For callers on a latency-critical thread -- the event
 * loop of an async_take -- which then grow the arena from a worker thread with tsb_host_alloc.  No reference
 * counterpart (the reference allocates pageable host memory per copy, tensor.py:248-254).
 */
int tsb_host_alloc_nogrow(TsbCtx* ctx, uint64_t nbytes, void** hptr);
int tsb_host_free(TsbCtx* ctx, void* hptr);

/* Bytes currently handed out / reserved from the driver, per arena (0 = host, 1 = device). */
int tsb_arena_stats(TsbCtx* ctx, int which, uint64_t* in_use, uint64_t* reserved, uint64_t* bound);
/* Return all unused arena segments to the driver. */
int tsb_arena_trim(TsbCtx* ctx);

/* ---- streams and events -------------------------------------------------------------- */

/* Record an event on one of the context's streams. */
int tsb_event_record(TsbCtx* ctx, int stream, TsbEvent* ev);
/* Make a context stream wait for an event recorded on another context stream. */
int tsb_stream_wait_event(TsbCtx* ctx, int stream, TsbEvent ev);
/* TSB_OK if complete, TSB_EAGAIN if still pending. */
int tsb_event_query(TsbCtx* ctx, TsbEvent ev);
/* Block the calling host thread until the event completes. */
int tsb_event_sync(TsbCtx* ctx, TsbEvent ev);
int tsb_event_elapsed_ms(TsbCtx* ctx, TsbEvent start, TsbEvent stop, float* ms);
int tsb_event_release(TsbCtx* ctx, TsbEvent ev);

/*
 * Order a context stream after all work currently enqueued on a foreign stream
 * (`cuda_stream` is a cudaStream_t, e.g. torch.cuda.current_stream().cuda_stream; 0 = the
 * legacy default stream), and the reverse.  The reference gets this ordering implicitly from
 * running on the current stream and from the synchronous .cpu() (batcher.py:156).
 */
int tsb_stream_wait_external(TsbCtx* ctx, int stream, uint64_t cuda_stream);
int tsb_external_wait_stream(TsbCtx* ctx, uint64_t cuda_stream, int stream);
int tsb_stream_sync(TsbCtx* ctx, int stream);

/* ---- the hot path --------------------------------------------------------------------- */

/*
 * Multi-region strided copy, ONE kernel launch for all `n` regions on context stream
 * `stream`.  `regions` is a host array; it is copied before the call returns (the caller may
 * reuse it immediately).  The descriptor and tile tables live in device scratch owned by the
 * context until the launch completes.
 *
 *   gather-pack   (batcher.py:152-155; tensor.py:252-254 + serialization.py:182-185):
 *       src = tensor storage (strided), dst = slab + byte offset (contiguous)
 *   scatter-unpack (tensor.py:329-338, 356-358):
 *       src = slab + byte offset, dst = destination tensor (strided)
 *   reshard       (sharded_tensor.py:271-284, 296-309):
 *       src = box in the persisted piece, dst = box in the local shard; src may be memory of a
 *       peer GPU mapped with tsb_ipc_open.
 *
 * Byte-exact; no alignment requirement on any address or length.
 */
int tsb_copy_regions(TsbCtx* ctx, int stream, const TsbRegion* regions, uint32_t n);

/*
 * Same launch, plus two events that bracket ONLY the copy kernels on the device (descriptor upload
 * and tile-table expansion excluded): what bench.py divides the algorithmic bytes by for the
 * roofline.  Release both with tsb_event_release.  No reference counterpart (the reference has no
 * device timing).
 */
int tsb_copy_regions_timed(TsbCtx* ctx, int stream, const TsbRegion* regions, uint32_t n, TsbEvent* kernel_start,
                           TsbEvent* kernel_stop);

/*
 * The same multi-region copy for callers that must not allocate.  No counterpart in the reference.
 *
 * tsb_copy_regions takes its descriptor and tile tables from the context's arenas, which can end in
 * cudaMalloc / cudaHostAlloc.  That is not acceptable everywhere: a trainer whose caching allocator
 * already owns all of HBM (an extra cudaMalloc fails or eats into the pool), code that cannot afford
 * the device-wide synchronisation cudaMalloc / cudaFree imply while other streams are busy, or an
 * emergency path that wants to save state right after an out-of-memory error.  This entry point needs
 * no context and allocates nothing, on the device or on the host: its scratch is one fixed array of
 * TSB_STATIC_SCRATCH_BYTES inside the library, present on every GPU that loads the library for as long
 * as the process lives.
 *
 *   device       CUDA device the addresses in `regions` belong to
 *   cuda_stream  a cudaStream_t of that device (0 = legacy default stream); the copy is ordered on it
 *   regions, n   as for tsb_copy_regions (host array, consumed before the call returns)
 *
 * Limits, checked before anything is launched (TSB_EINVAL and a message naming the limit): the
 * descriptors of one call must fit TSB_STATIC_UPLOAD_BYTES (32 B per contiguous run, 32 B per 32 KiB of
 * short-run strided payload plus 128 B per such region) and the contiguous payload must fit
 * TSB_STATIC_MAX_TILES tiles of 32 KiB (1.5 GiB).  Split larger jobs into several calls on one stream.
 *
 * The first call on each device looks up the device address of the array (cudaGetSymbolAddress), which also
 * makes the runtime load the library's device code and with it the array; make that first call early --
 * before any stream capture begins, and before memory runs short if the call is meant for an
 * out-of-memory path (a one-region copy of a few bytes will do).  A call on a stream that is being
 * captured is refused with TSB_EINVAL.  Unlike the other entry points this one can block the host: the
 * descriptors are uploaded from ordinary (pageable) memory, so the call returns only after the work queued
 * on `cuda_stream` before it has finished and the upload has been handed to the copy engine; the copy
 * kernels themselves run asynchronously.  All calls share the array and it keeps whatever
 * the previous call left in it; nothing needs zeroing because each call rewrites every entry it reads,
 * but two calls whose kernels can overlap on the GPU (different streams, no ordering between them)
 * race on it -- keep such calls on one stream or order them with events.
 */
#define TSB_STATIC_SCRATCH_BYTES (2u * 1024u * 1024u)
#define TSB_STATIC_UPLOAD_BYTES (512u * 1024u)
#define TSB_STATIC_MAX_TILES ((TSB_STATIC_SCRATCH_BYTES - TSB_STATIC_UPLOAD_BYTES) / 32u)
int tsb_copy_regions_static(int device, uint64_t cuda_stream, const TsbRegion* regions, uint32_t n);

/* Device <-> pinned host copies on a context stream (K3 / K6). `hptr` should come from tsb_host_alloc. */
int tsb_memcpy_d2h(TsbCtx* ctx, int stream, void* hptr, uint64_t dptr, uint64_t nbytes);
int tsb_memcpy_h2d(TsbCtx* ctx, int stream, uint64_t dptr, const void* hptr, uint64_t nbytes);
/*
 * `n` device -> pinned host copies issued INTERLEAVED on one stream, `chunk` bytes at a time round-robin over
 * the copies, with an event recorded after every chunk: events[] receives, copy after copy,
 * ceil(nbytes[b] / chunk) handles (release each with tsb_event_release).  One call per batch of files, so that
 * every file of the batch starts arriving at once and a writer can follow each of them
 * (tsb_file_write_streamed).  Replaces one gpu_buf.cpu() per slab (batcher.py:156) / tensor.to("cpu") per
 * large tensor (io_preparers/tensor.py:252-254), which finish one file before the next one starts.
 */
int tsb_memcpy_d2h_interleaved(TsbCtx* ctx, int stream, uint32_t n, const uint64_t* hptrs, const uint64_t* dptrs,
                               const uint64_t* nbytes, uint64_t chunk, TsbEvent* events);
/* Device -> device (same or peer device) on a context stream. */
int tsb_memcpy_d2d(TsbCtx* ctx, int stream, uint64_t dst, uint64_t src, uint64_t nbytes);

/*
 * Host-side parallel memcpy (dst/src are host addresses) with `nthreads` worker threads
 * (0 = library default).  Used to move bytes a third-party StoragePlugin returned in a
 * BytesIO (io_types.py:68-72) into pinned memory; replaces the `buf[a:b]` slicing copies of
 * BatchedBufferConsumer (batcher.py:364-375).
 */
int tsb_host_memcpy(void* dst, const void* src, uint64_t nbytes, int nthreads);
/*
 * `count` independent host copies dst[i] <- src[i] (nbytes[i] bytes each; host addresses), balanced over
 * `nthreads` workers by bytes.  The CPU-tensor side of the path: fills a host slab from its member tensors
 * (BatchedBufferStager.stage_buffer, batcher.py:63-90: a zero-filled bytearray plus one interpreted slice
 * assignment per member on the event-loop thread) and scatters a slab that was read back into its member
 * tensors (BatchedBufferConsumer.consume_buffer, batcher.py:364-375 + tensor.py:329-338).
 */
int tsb_host_gather(const uint64_t* dst, const uint64_t* src, const uint64_t* nbytes, uint64_t count, int nthreads);

/* ---- peer access (resharding over NVLink) ---------------------------------------------- */

#define TSB_IPC_HANDLE_BYTES 64
/* Export a device allocation (base of a tsb_dev_alloc block or any cudaMalloc'd pointer) so a
 * peer process on the same box can map it.  `handle` receives TSB_IPC_HANDLE_BYTES bytes and
 * `offset` the offset of dptr within the exported allocation. */
int tsb_ipc_export(TsbCtx* ctx, uint64_t dptr, uint8_t* handle, uint64_t* offset);
/* Map a peer allocation; returns its base address in this process. */
int tsb_ipc_open(TsbCtx* ctx, const uint8_t* handle, uint64_t* base);
int tsb_ipc_close(TsbCtx* ctx, uint64_t base);
/* Enable direct access from the context's device to `peer_device` (same process). */
int tsb_peer_enable(TsbCtx* ctx, int peer_device);

/* ---- storage sink fast path (behind the unchanged StoragePlugin ABI) ------ */

/* File flags.  TSB_FILE_DIRECT: the 4 KiB-aligned body of the transfer bypasses the page cache
 * (O_DIRECT) when `hptr` and the file offset are 4 KiB aligned -- pinned arena blocks are; the
 * sub-page tail, unaligned requests and file systems without O_DIRECT use the buffered path, so the
 * bytes on disk never depend on the flag. */
#define TSB_FILE_DIRECT 1u

/* Write `nbytes` at `hptr` to a new file `path` (parents must exist) with `nthreads` parallel
 * pwrite workers (0 = library default).  Replaces the aiofiles write of FSStoragePlugin.write
 * (storage_plugins/fs.py:26-36). */
int tsb_file_write(const char* path, const void* hptr, uint64_t nbytes, int nthreads, uint32_t flags);
/*
 * Write a buffer that is still being filled: bytes [0, ends[i]) are valid once events[i] (in order) has
 * completed.  The calling thread waits for each event in turn and writes what has arrived, so the file is
 * complete shortly after the last copy.  Same file as tsb_file_write produces (FSStoragePlugin.write,
 * fs.py:26-36, which needs the whole buffer before it starts).
 */
int tsb_file_write_streamed(TsbCtx* ctx, const char* path, const void* hptr, uint64_t nbytes, const TsbEvent* events,
                            const uint64_t* ends, uint32_t n);
/* Read `nbytes` at `offset` of `path` into `hptr` (FSStoragePlugin.read, fs.py:38-49). */
int tsb_file_read(const char* path, uint64_t offset, void* hptr, uint64_t nbytes, int nthreads, uint32_t flags);

#ifdef __cplusplus
}
#endif
#endif /* TSB200_H_ */
