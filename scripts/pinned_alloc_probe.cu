// How fast can a pinned host segment be obtained?  cudaHostAlloc vs. anonymous mmap (+ MADV_HUGEPAGE),
// parallel first touch and cudaHostRegister; and is D2H into the registered mapping as fast?
//   nvcc -O2 -o scripts/pinned_alloc_probe scripts/pinned_alloc_probe.cu -lpthread
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include <sys/mman.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static void touch(char* p, size_t n, int threads) {
    std::vector<std::thread> th; size_t per = (n / threads + 4095) & ~size_t(4095);
    for (int t = 0; t < threads; ++t) th.emplace_back([=] { size_t lo = t * per, hi = lo + per < n ? lo + per : n; for (size_t o = lo; o < hi; o += 4096) p[o] = 0; });
    for (auto& t : th) t.join();
}
static double d2h_gbps(void* h, void* d, size_t n, cudaStream_t s) {
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    CK(cudaMemcpyAsync(h, d, n, cudaMemcpyDeviceToHost, s)); CK(cudaStreamSynchronize(s));
    CK(cudaEventRecord(a, s)); for (int r = 0; r < 4; ++r) CK(cudaMemcpyAsync(h, d, n, cudaMemcpyDeviceToHost, s));
    CK(cudaEventRecord(b, s)); CK(cudaEventSynchronize(b)); float ms; CK(cudaEventElapsedTime(&ms, a, b)); return 4.0 * n / ms / 1e6;
}
int main() {
    CK(cudaSetDevice(0)); CK(cudaFree(0));
    const size_t N = 1ull << 30; void* d; CK(cudaMalloc(&d, N)); CK(cudaMemset(d, 7, N)); cudaStream_t s; CK(cudaStreamCreate(&s));
    for (int rep = 0; rep < 2; ++rep) {
        double t0 = now(); void* h; CK(cudaHostAlloc(&h, N, cudaHostAllocPortable)); double t1 = now();
        double bw = d2h_gbps(h, d, N, s); double t2 = now(); CK(cudaFreeHost(h)); double t3 = now();
        printf("{\"how\": \"cudaHostAlloc\", \"alloc_ms_per_GiB\": %.1f, \"free_ms\": %.1f, \"d2h_gbps\": %.1f}\n", (t1 - t0) * 1e3, (t3 - t2) * 1e3, bw);
    }
    for (int huge = 0; huge < 2; ++huge) for (int threads : {1, 4, 16}) {
        double t0 = now();
        char* p = (char*)mmap(nullptr, N + (2u << 20), PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        if (p == MAP_FAILED) { printf("mmap failed\n"); return 1; }
        char* q = (char*)(((uintptr_t)p + (2u << 20) - 1) & ~uintptr_t((2u << 20) - 1));
        if (huge) madvise(q, N, MADV_HUGEPAGE);
        double t1 = now(); touch(q, N, threads); double t2 = now();
        CK(cudaHostRegister(q, N, cudaHostRegisterPortable)); double t3 = now();
        double bw = d2h_gbps(q, d, N, s); double t4 = now();
        CK(cudaHostUnregister(q)); double t5 = now(); munmap(p, N + (2u << 20)); double t6 = now();
        printf("{\"how\": \"mmap+touch+register\", \"hugepage_advice\": %d, \"touch_threads\": %d, \"touch_ms\": %.1f, \"register_ms\": %.1f, \"total_ms_per_GiB\": %.1f, \"unregister_ms\": %.1f, \"munmap_ms\": %.1f, \"d2h_gbps\": %.1f}\n",
               huge, threads, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t3 - t0) * 1e3, (t5 - t4) * 1e3, (t6 - t5) * 1e3, bw);
        fflush(stdout);
    }
    // register without touching first (the driver faults the pages in itself)
    for (int huge = 0; huge < 2; ++huge) {
        double t0 = now();
        char* p = (char*)mmap(nullptr, N + (2u << 20), PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        char* q = (char*)(((uintptr_t)p + (2u << 20) - 1) & ~uintptr_t((2u << 20) - 1));
        if (huge) madvise(q, N, MADV_HUGEPAGE);
        CK(cudaHostRegister(q, N, cudaHostRegisterPortable)); double t1 = now();
        double bw = d2h_gbps(q, d, N, s);
        CK(cudaHostUnregister(q)); munmap(p, N + (2u << 20));
        printf("{\"how\": \"mmap+register (no touch)\", \"hugepage_advice\": %d, \"total_ms_per_GiB\": %.1f, \"d2h_gbps\": %.1f}\n", huge, (t1 - t0) * 1e3, bw);
    }
    // two cudaHostAlloc calls of 512 MiB from two threads at once
    { double t0 = now(); void* h[4]; std::vector<std::thread> th;
      for (int i = 0; i < 4; ++i) th.emplace_back([&, i] { CK(cudaSetDevice(0)); CK(cudaHostAlloc(&h[i], N / 4, cudaHostAllocPortable)); });
      for (auto& t : th) t.join(); double t1 = now();
      printf("{\"how\": \"4 threads x cudaHostAlloc(256 MiB)\", \"total_ms_per_GiB\": %.1f}\n", (t1 - t0) * 1e3);
      for (int i = 0; i < 4; ++i) CK(cudaFreeHost(h[i])); }
    return 0;
}
