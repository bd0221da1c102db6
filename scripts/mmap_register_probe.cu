// Probe: can a tmpfs file mapping be cudaHostRegister'ed and used as a DMA target/source? (scratch, not product)
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <fcntl.h>
#include <sys/mman.h>
#include <unistd.h>
#include <thread>
#include <vector>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("ERR %s: %s\n",#x,cudaGetErrorString(e)); return 1;} }while(0)
static double now(){ return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main(int argc,char**argv){
  const char* path = argc>1?argv[1]:"/dev/shm/tsb_probe.bin";
  const size_t N = (size_t)(argc>2?atof(argv[2]):1.0)*(1ull<<30);
  int nthreads = argc>3?atoi(argv[3]):8;
  void* d; CK(cudaMalloc(&d,N)); CK(cudaMemset(d,0x5a,N));
  cudaStream_t s; CK(cudaStreamCreate(&s));
  // ---- write direction: new file, D2H straight into the page cache ----
  unlink(path);
  double t0=now();
  int fd=open(path,O_RDWR|O_CREAT|O_TRUNC,0644); if(fd<0){perror("open");return 1;}
  if(ftruncate(fd,N)){perror("ftruncate");return 1;}
  void* p=mmap(nullptr,N,PROT_READ|PROT_WRITE,MAP_SHARED,fd,0); if(p==MAP_FAILED){perror("mmap");return 1;}
  double t1=now();
  // fault the pages in, in parallel (page allocation + zeroing)
  { std::vector<std::thread> th; size_t chunk=(N/nthreads+4095)/4096*4096;
    for(int i=0;i<nthreads;++i) th.emplace_back([=]{ size_t lo=i*chunk, hi=lo+chunk>N?N:lo+chunk; if(lo<hi) madvise((char*)p+lo,hi-lo,MADV_POPULATE_WRITE); });
    for(auto&t:th) t.join(); }
  double t2=now();
  cudaError_t e=cudaHostRegister(p,N,cudaHostRegisterDefault);
  double t3=now();
  printf("write: open+truncate+mmap %.1f ms, populate(%d thr) %.1f ms (%.1f GB/s), cudaHostRegister %.1f ms -> %s\n",(t1-t0)*1e3,nthreads,(t2-t1)*1e3,N/(t2-t1)/1e9,(t3-t2)*1e3,cudaGetErrorString(e));
  if(e==cudaSuccess){
    cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a,s); CK(cudaMemcpyAsync(p,d,N,cudaMemcpyDeviceToHost,s)); cudaEventRecord(b,s); CK(cudaEventSynchronize(b));
    float ms; cudaEventElapsedTime(&ms,a,b); printf("  D2H into the mapping: %.1f ms = %.1f GB/s\n",ms,N/ms/1e6);
    double u0=now(); CK(cudaHostUnregister(p)); double u1=now(); printf("  unregister %.1f ms; total write path %.1f ms = %.1f GB/s\n",(u1-u0)*1e3,(u1-t0)*1e3,N/(u1-t0)/1e9);
  } else { (void)cudaGetLastError(); }
  munmap(p,N); close(fd);
  // check content through the file API
  { int f2=open(path,O_RDONLY); unsigned char buf[4096]; ssize_t k=pread(f2,buf,4096,N-4096); bool ok=k==4096; for(int i=0;i<4096&&ok;++i) ok=buf[i]==0x5a; printf("  file content %s\n",ok?"ok":"BAD"); close(f2); }
  // ---- read direction: existing file -> H2D straight from the page cache ----
  double r0=now();
  fd=open(path,O_RDWR); p=mmap(nullptr,N,PROT_READ|PROT_WRITE,MAP_SHARED,fd,0);
  { std::vector<std::thread> th; size_t chunk=(N/nthreads+4095)/4096*4096;
    for(int i=0;i<nthreads;++i) th.emplace_back([=]{ size_t lo=i*chunk, hi=lo+chunk>N?N:lo+chunk; if(lo<hi) madvise((char*)p+lo,hi-lo,MADV_POPULATE_READ); });
    for(auto&t:th) t.join(); }
  double r1=now();
  e=cudaHostRegister(p,N,cudaHostRegisterDefault);
  double r2=now();
  printf("read: mmap+populate %.1f ms, cudaHostRegister %.1f ms -> %s\n",(r1-r0)*1e3,(r2-r1)*1e3,cudaGetErrorString(e));
  if(e==cudaSuccess){
    CK(cudaMemset(d,0,N));
    cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a,s); CK(cudaMemcpyAsync(d,p,N,cudaMemcpyHostToDevice,s)); cudaEventRecord(b,s); CK(cudaEventSynchronize(b));
    float ms; cudaEventElapsedTime(&ms,a,b); printf("  H2D from the mapping: %.1f ms = %.1f GB/s\n",ms,N/ms/1e6);
    double u0=now(); CK(cudaHostUnregister(p)); double u1=now(); printf("  unregister %.1f ms; total read path %.1f ms = %.1f GB/s\n",(u1-u0)*1e3,(u1-r0)*1e3,N/(u1-r0)/1e9);
    unsigned char h[16]; CK(cudaMemcpy(h,(char*)d+N-16,16,cudaMemcpyDeviceToHost)); printf("  device content %s\n",h[15]==0x5a?"ok":"BAD");
  } else { (void)cudaGetLastError(); }
  munmap(p,N); close(fd); unlink(path);
  return 0;
}
