// Does the NUMA placement of the pinned destination explain why concurrent D2H does not scale on this box?
// One process, one thread per GPU; each copies 1 GiB device -> pinned host 6x, all GPUs at the same time.
// The pinned buffer is allocated under set_mempolicy(MPOL_BIND, node) for several placements.
//   nvcc -O2 -o scripts/numa_d2h_probe scripts/numa_d2h_probe.cu -lpthread
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include <atomic>
#include <unistd.h>
#include <sys/syscall.h>
#include <dirent.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)
static const int MPOL_DEFAULT_ = 0, MPOL_BIND_ = 2, MPOL_INTERLEAVE_ = 3;
static long set_policy(int mode, unsigned long mask) {
    return syscall(SYS_set_mempolicy, mode, mask ? &mask : nullptr, mask ? 64UL : 0UL);
}
static int read_int(const std::string& path, int dflt) {
    FILE* f = fopen(path.c_str(), "r"); if (!f) return dflt; int v = dflt; if (fscanf(f, "%d", &v) != 1) v = dflt; fclose(f); return v;
}
static std::string read_line(const std::string& path) {
    FILE* f = fopen(path.c_str(), "r"); if (!f) return "?"; char b[256] = {0}; if (!fgets(b, 255, f)) b[0] = 0; fclose(f);
    size_t n = strlen(b); if (n && b[n - 1] == '\n') b[n - 1] = 0; return b;
}
int main(int argc, char** argv) {
    int ndev = 0; CK(cudaGetDeviceCount(&ndev));
    int nnodes = 0;
    for (int i = 0; i < 16; ++i) { if (access(("/sys/devices/system/node/node" + std::to_string(i)).c_str(), F_OK) == 0) nnodes = i + 1; }
    printf("gpus %d numa_nodes %d online_cpus %ld\n", ndev, nnodes, sysconf(_SC_NPROCESSORS_ONLN));
    for (int i = 0; i < nnodes; ++i)
        printf("  node%d cpus %s\n", i, read_line("/sys/devices/system/node/node" + std::to_string(i) + "/cpulist").c_str());
    std::vector<int> gpu_node(ndev, -1);
    for (int d = 0; d < ndev; ++d) {
        char bdf[32]; CK(cudaDeviceGetPCIBusId(bdf, sizeof bdf, d));
        for (char* p = bdf; *p; ++p) *p = tolower(*p);
        gpu_node[d] = read_int(std::string("/sys/bus/pci/devices/") + bdf + "/numa_node", -1);
        printf("  gpu%d %s numa_node %d\n", d, bdf, gpu_node[d]);
    }
    const size_t N = 1ull << 30; const int reps = 6;
    struct Placement { const char* name; int mode; };  // mode: -1 default, -2 gpu-local, -3 interleave, >=0 that node
    std::vector<Placement> pl = {{"default", -1}, {"interleave", -3}};
    if (nnodes > 1) { pl.push_back({"gpu_local", -2}); pl.push_back({"node0", 0}); pl.push_back({"node1", 1}); }
    std::vector<void*> dbuf(ndev); std::vector<cudaStream_t> st(ndev);
    for (int d = 0; d < ndev; ++d) { CK(cudaSetDevice(d)); CK(cudaMalloc(&dbuf[d], N)); CK(cudaMemset(dbuf[d], d + 1, N)); CK(cudaStreamCreateWithFlags(&st[d], cudaStreamNonBlocking)); }
    for (auto& p : pl) {
        std::vector<void*> h(ndev);
        for (int d = 0; d < ndev; ++d) {
            CK(cudaSetDevice(d));
            int node = p.mode >= 0 ? p.mode : (p.mode == -2 ? gpu_node[d] : -1);
            if (p.mode == -3) set_policy(MPOL_INTERLEAVE_, (1ul << nnodes) - 1);
            else if (node >= 0) set_policy(MPOL_BIND_, 1ul << node);
            CK(cudaHostAlloc(&h[d], N, cudaHostAllocPortable));
            memset(h[d], 0, N);
            set_policy(MPOL_DEFAULT_, 0);
        }
        for (int active = 1; active <= ndev; active *= 2) {
            std::vector<float> gbps(active, 0.f); std::atomic<int> ready{0}; std::vector<std::thread> th;
            for (int d = 0; d < active; ++d) th.emplace_back([&, d] {
                CK(cudaSetDevice(d)); cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
                CK(cudaMemcpyAsync(h[d], dbuf[d], N, cudaMemcpyDeviceToHost, st[d])); CK(cudaStreamSynchronize(st[d]));
                ready++; while (ready.load() < active) {}
                CK(cudaEventRecord(a, st[d]));
                for (int r = 0; r < reps; ++r) CK(cudaMemcpyAsync(h[d], dbuf[d], N, cudaMemcpyDeviceToHost, st[d]));
                CK(cudaEventRecord(b, st[d])); CK(cudaEventSynchronize(b));
                float ms; CK(cudaEventElapsedTime(&ms, a, b)); gbps[d] = reps * (double)N / ms / 1e6;
            });
            for (auto& t : th) t.join();
            float sum = 0, mn = 1e9; for (float g : gbps) { sum += g; mn = g < mn ? g : mn; }
            printf("{\"placement\": \"%s\", \"active_gpus\": %d, \"d2h_sum_gbps\": %.1f, \"d2h_min_gbps\": %.1f, \"per_gpu\": [", p.name, active, sum, mn);
            for (int d = 0; d < active; ++d) printf("%s%.1f", d ? ", " : "", gbps[d]);
            printf("]}\n"); fflush(stdout);
        }
        for (int d = 0; d < ndev; ++d) CK(cudaFreeHost(h[d]));
    }
    return 0;
}
