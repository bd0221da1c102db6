// Scratch micro-benchmark: variants of the streaming copy inner loop (not product code).
// nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o copy_variants copy_variants.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <algorithm>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("ERR %s line %d: %s\n",#x,__LINE__,cudaGetErrorString(e)); exit(1);} }while(0)

// ---- V: persistent tiles, templated unroll/threads/cache ops -------------------------------
template<int MODE> __device__ __forceinline__ uint4 ld(const uint4* p){
  if (MODE==0) return __ldcs(p); if (MODE==1) return *p; if (MODE==2) return __ldg(p);
  uint4 r; asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];":"=r"(r.x),"=r"(r.y),"=r"(r.z),"=r"(r.w):"l"(p)); return r; }
template<int MODE> __device__ __forceinline__ void st(uint4* p, uint4 v){
  if (MODE==0) __stcs(p,v); else if (MODE==3) { asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"::"l"(p),"r"(v.x),"r"(v.y),"r"(v.z),"r"(v.w)); } else *p=v; }

template<int THREADS,int U,int MODE,int MINB>
__global__ void __launch_bounds__(THREADS,MINB) k_persist(uint4* __restrict__ d, const uint4* __restrict__ s, uint64_t nvec, uint32_t tile_vec){
  uint64_t ntiles=(nvec+tile_vec-1)/tile_vec;
  for(uint64_t t=blockIdx.x;t<ntiles;t+=gridDim.x){
    uint64_t b=t*tile_vec, e=min(b+(uint64_t)tile_vec,nvec);
    uint64_t j=b+threadIdx.x;
    for(;j+(uint64_t)(U-1)*THREADS<e;j+=(uint64_t)U*THREADS){
      uint4 v[U];
      #pragma unroll
      for(int u=0;u<U;++u) v[u]=ld<MODE>(s+j+(uint64_t)u*THREADS);
      #pragma unroll
      for(int u=0;u<U;++u) st<MODE>(d+j+(uint64_t)u*THREADS,v[u]);
    }
    for(;j<e;j+=THREADS) st<MODE>(d+j,ld<MODE>(s+j));
  }
}
// ---- simple non persistent: each block U*THREADS vectors ------------------------------------
template<int THREADS,int U,int MODE>
__global__ void __launch_bounds__(THREADS) k_simple(uint4* __restrict__ d,const uint4* __restrict__ s,uint64_t nvec){
  uint64_t j=(uint64_t)blockIdx.x*U*THREADS+threadIdx.x;
  uint4 v[U];
  #pragma unroll
  for(int u=0;u<U;++u) if(j+(uint64_t)u*THREADS<nvec) v[u]=ld<MODE>(s+j+(uint64_t)u*THREADS);
  #pragma unroll
  for(int u=0;u<U;++u) if(j+(uint64_t)u*THREADS<nvec) st<MODE>(d+j+(uint64_t)u*THREADS,v[u]);
}
// ---- shifted (src misaligned by S bytes) two-load variant ----------------------------------
template<int SW> __device__ __forceinline__ uint4 realign(const uint4&a,const uint4&b,uint32_t sb){
  const uint32_t w[8]={a.x,a.y,a.z,a.w,b.x,b.y,b.z,b.w}; uint4 o;
  o.x=__funnelshift_r(w[SW],w[SW+1],sb); o.y=__funnelshift_r(w[SW+1],w[SW+2],sb);
  o.z=__funnelshift_r(w[SW+2],w[SW+3],sb); o.w=__funnelshift_r(w[SW+3],w[SW+4],sb); return o; }
template<int THREADS,int U,int MINB>
__global__ void __launch_bounds__(THREADS,MINB) k_shift2(uint4* __restrict__ d,const uint4* __restrict__ a4,uint64_t nvec,uint32_t tile_vec,uint32_t sb){
  uint64_t ntiles=(nvec+tile_vec-1)/tile_vec;
  for(uint64_t t=blockIdx.x;t<ntiles;t+=gridDim.x){
    uint64_t b0=t*tile_vec,e=min(b0+(uint64_t)tile_vec,nvec); uint64_t j=b0+threadIdx.x;
    for(;j+(uint64_t)(U-1)*THREADS<e;j+=(uint64_t)U*THREADS){
      uint4 a[U],b[U];
      #pragma unroll
      for(int u=0;u<U;++u){ a[u]=__ldg(a4+j+(uint64_t)u*THREADS); b[u]=__ldg(a4+j+(uint64_t)u*THREADS+1);}
      #pragma unroll
      for(int u=0;u<U;++u) __stcs(d+j+(uint64_t)u*THREADS,realign<1>(a[u],b[u],sb));
    }
    for(;j<e;j+=THREADS){ uint4 a=__ldg(a4+j),b=__ldg(a4+j+1); __stcs(d+j,realign<1>(a,b,sb)); }
  }
}
// shuffle variant: one load per vector (+1 for lane 0), neighbours through shfl
template<int THREADS,int U,int MINB>
__global__ void __launch_bounds__(THREADS,MINB) k_shiftshfl(uint4* __restrict__ d,const uint4* __restrict__ a4,uint64_t nvec,uint32_t tile_vec,uint32_t sb){
  uint64_t ntiles=(nvec+tile_vec-1)/tile_vec; const uint32_t lane=threadIdx.x&31;
  for(uint64_t t=blockIdx.x;t<ntiles;t+=gridDim.x){
    uint64_t b0=t*tile_vec,e=min(b0+(uint64_t)tile_vec,nvec); uint64_t j=b0+threadIdx.x;
    for(;j+(uint64_t)(U-1)*THREADS<e;j+=(uint64_t)U*THREADS){
      uint4 a[U]; uint4 x[U];
      #pragma unroll
      for(int u=0;u<U;++u){ a[u]=__ldcs(a4+j+(uint64_t)u*THREADS); }
      #pragma unroll
      for(int u=0;u<U;++u){ if(lane==31) x[u]=__ldcs(a4+j+(uint64_t)u*THREADS+1); }
      #pragma unroll
      for(int u=0;u<U;++u){
        uint4 b; b.x=__shfl_down_sync(0xffffffffu,a[u].x,1); b.y=__shfl_down_sync(0xffffffffu,a[u].y,1);
        b.z=__shfl_down_sync(0xffffffffu,a[u].z,1); b.w=__shfl_down_sync(0xffffffffu,a[u].w,1);
        if(lane==31) b=x[u];
        __stcs(d+j+(uint64_t)u*THREADS,realign<1>(a[u],b,sb));
      }
    }
    for(;j<e;j+=THREADS){ uint4 a=__ldg(a4+j),b=__ldg(a4+j+1); __stcs(d+j,realign<1>(a,b,sb)); }
  }
}
// ---- TMA bulk pipeline: one warp per CTA issues cp.async.bulk g2s then s2g ------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p){ return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b,int c){ asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;"::"r"(smem_u32(b)),"r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b,uint32_t bytes){ asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"::"r"(smem_u32(b)),"r"(bytes):"memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b,uint32_t parity){
  asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}"::"r"(smem_u32(b)),"r"(parity):"memory"); }
__device__ __forceinline__ void bulk_g2s(void* sdst,const void* gsrc,uint32_t bytes,uint64_t* bar){
  asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"::"r"(smem_u32(sdst)),"l"(gsrc),"r"(bytes),"r"(smem_u32(bar)):"memory"); }
__device__ __forceinline__ void bulk_s2g(void* gdst,const void* ssrc,uint32_t bytes){
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"::"l"(gdst),"r"(smem_u32(ssrc)),"r"(bytes):"memory"); }
__device__ __forceinline__ void bulk_commit(){ asm volatile("cp.async.bulk.commit_group;":::"memory"); }
template<int N> __device__ __forceinline__ void bulk_wait_read(){ asm volatile("cp.async.bulk.wait_group.read %0;"::"n"(N):"memory"); }

template<int STAGES,int CHUNK>
__global__ void __launch_bounds__(32) k_tma(uint8_t* __restrict__ d,const uint8_t* __restrict__ s,uint64_t nbytes){
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t full[STAGES];
  if(threadIdx.x==0){ for(int i=0;i<STAGES;++i) mbar_init(&full[i],1); asm volatile("fence.mbarrier_init.release.cluster;":::"memory"); }
  __syncwarp();
  if(threadIdx.x!=0) return;
  const uint64_t nchunks=(nbytes+CHUNK-1)/CHUNK;
  // chunks owned by this CTA: c = blockIdx.x + k*gridDim.x
  uint64_t k_issue=0,k_done=0; 
  const uint64_t mine=(nchunks>blockIdx.x)?(nchunks-blockIdx.x+gridDim.x-1)/gridDim.x:0;
  // prologue
  for(;k_issue<mine&&k_issue<STAGES;++k_issue){
    uint64_t c=blockIdx.x+k_issue*gridDim.x; uint32_t len=(uint32_t)min((uint64_t)CHUNK,nbytes-c*CHUNK);
    int st=k_issue%STAGES; mbar_expect(&full[st],len); bulk_g2s(smem+st*CHUNK,s+c*CHUNK,len,&full[st]);
  }
  for(;k_done<mine;++k_done){
    int st=k_done%STAGES; uint32_t par=(k_done/STAGES)&1;
    mbar_wait(&full[st],par);
    uint64_t c=blockIdx.x+k_done*gridDim.x; uint32_t len=(uint32_t)min((uint64_t)CHUNK,nbytes-c*CHUNK);
    asm volatile("fence.proxy.async.shared::cta;":::"memory");
    bulk_s2g(d+c*CHUNK,smem+st*CHUNK,len); bulk_commit();
    // refill the stage that was stored one iteration ago (its store has been read out of smem)
    if(k_issue<mine && k_done>=1){
      bulk_wait_read<1>();
      uint64_t c2=blockIdx.x+k_issue*gridDim.x; uint32_t len2=(uint32_t)min((uint64_t)CHUNK,nbytes-c2*CHUNK);
      int st2=k_issue%STAGES; mbar_expect(&full[st2],len2); bulk_g2s(smem+st2*CHUNK,s+c2*CHUNK,len2,&full[st2]); ++k_issue;
    }
  }
  bulk_wait_read<0>();
  asm volatile("cp.async.bulk.wait_group 0;":::"memory");
}


// ---- simple with an occupancy cap (dynamic smem) to see how much residency matters ----------
// ---- simple + one dependent descriptor load per CTA (tile table) ----------------------------
struct __align__(32) Tile32 { uint64_t src,dst; uint32_t len,region; uint64_t off; };
template<int THREADS,int U>
__global__ void __launch_bounds__(THREADS) k_tiled(const Tile32* __restrict__ tiles){
  const uint4* tp=reinterpret_cast<const uint4*>(tiles+blockIdx.x);
  uint4 t0=__ldg(tp), t1=__ldg(tp+1);
  const uint4* s=(const uint4*)(((uint64_t)t0.y<<32)|t0.x); uint4* d=(uint4*)(((uint64_t)t0.w<<32)|t0.z);
  uint32_t nvec=t1.x>>4; uint32_t j=threadIdx.x;
  uint4 v[U];
  #pragma unroll
  for(int u=0;u<U;++u) if(j+u*THREADS<nvec) v[u]=__ldcs(s+j+u*THREADS);
  #pragma unroll
  for(int u=0;u<U;++u) if(j+u*THREADS<nvec) __stcs(d+j+u*THREADS,v[u]);
}
// ---- persistent with a dynamic (atomic) tile counter -----------------------------------------
template<int THREADS,int U,int MINB>
__global__ void __launch_bounds__(THREADS,MINB) k_dyn(uint4* __restrict__ d,const uint4* __restrict__ s,uint64_t nvec,unsigned long long* ctr){
  __shared__ unsigned long long cur;
  const uint64_t tile=(uint64_t)U*THREADS; const uint64_t ntiles=(nvec+tile-1)/tile;
  for(;;){
    if(threadIdx.x==0) cur=atomicAdd(ctr,1ull);
    __syncthreads(); uint64_t t=cur; __syncthreads();
    if(t>=ntiles) break;
    uint64_t j=t*tile+threadIdx.x; uint4 v[U];
    #pragma unroll
    for(int u=0;u<U;++u) if(j+(uint64_t)u*THREADS<nvec) v[u]=__ldcs(s+j+(uint64_t)u*THREADS);
    #pragma unroll
    for(int u=0;u<U;++u) if(j+(uint64_t)u*THREADS<nvec) __stcs(d+j+(uint64_t)u*THREADS,v[u]);
  }
}
// ---- non-persistent TMA: one warp per CTA, one bulk load + one bulk store --------------------
template<int CHUNK>
__global__ void __launch_bounds__(32) k_tma_simple(uint8_t* __restrict__ d,const uint8_t* __restrict__ s,uint64_t nbytes){
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t full;
  if(threadIdx.x!=0) return;
  mbar_init(&full,1); asm volatile("fence.mbarrier_init.release.cluster;":::"memory");
  uint64_t c=blockIdx.x; uint32_t len=(uint32_t)min((uint64_t)CHUNK,nbytes-c*CHUNK);
  mbar_expect(&full,len); bulk_g2s(smem,s+c*CHUNK,len,&full);
  mbar_wait(&full,0);
  asm volatile("fence.proxy.async.shared::cta;":::"memory");
  bulk_s2g(d+c*CHUNK,smem,len); bulk_commit(); bulk_wait_read<0>();
}
template<int THREADS,int U>
__global__ void __launch_bounds__(THREADS) k_simple_cap(uint4* __restrict__ d,const uint4* __restrict__ s,uint64_t nvec){
  extern __shared__ uint8_t dummy[];
  uint64_t j=(uint64_t)blockIdx.x*U*THREADS+threadIdx.x;
  uint4 v[U];
  #pragma unroll
  for(int u=0;u<U;++u) if(j+(uint64_t)u*THREADS<nvec) v[u]=__ldcs(s+j+(uint64_t)u*THREADS);
  #pragma unroll
  for(int u=0;u<U;++u) if(j+(uint64_t)u*THREADS<nvec) __stcs(d+j+(uint64_t)u*THREADS,v[u]);
}
// shifted simple: 2 loads per vector
template<int THREADS,int U>
__global__ void __launch_bounds__(THREADS) k_tiled_shift(const Tile32* __restrict__ tiles,uint32_t sb){
  const uint4* tp=reinterpret_cast<const uint4*>(tiles+blockIdx.x);
  uint4 t0=__ldg(tp), t1=__ldg(tp+1);
  const uint4* a4=(const uint4*)(((uint64_t)t0.y<<32)|t0.x); uint4* d=(uint4*)(((uint64_t)t0.w<<32)|t0.z);
  uint32_t nvec=(t1.x>>4)-1; uint32_t j=threadIdx.x;
  uint4 a[U],b[U];
  #pragma unroll
  for(int u=0;u<U;++u) if(j+u*THREADS<nvec){ a[u]=__ldg(a4+j+u*THREADS); b[u]=__ldg(a4+j+u*THREADS+1);}
  #pragma unroll
  for(int u=0;u<U;++u) if(j+u*THREADS<nvec) __stcs(d+j+u*THREADS,realign<1>(a[u],b[u],sb));
}

struct Timer{ cudaEvent_t a,b; Timer(){cudaEventCreate(&a);cudaEventCreate(&b);} };
template<class F> void bench(const char* name,uint64_t bytes,F f){
  Timer t; for(int i=0;i<3;++i) f(); CK(cudaDeviceSynchronize()); float best=1e9,tot=0; int it=10;
  for(int i=0;i<it;++i){ cudaEventRecord(t.a); f(); cudaEventRecord(t.b); cudaEventSynchronize(t.b); float ms; cudaEventElapsedTime(&ms,t.a,t.b); best=std::min(best,ms); tot+=ms; }
  CK(cudaGetLastError());
  printf("%-44s best %.4f ms  %.1f GB/s (r+w)   avg %.1f GB/s\n",name,best,2.0*bytes/best/1e6,2.0*bytes/(tot/it)/1e6); fflush(stdout);
}
int main(){
  const uint64_t N=4ull<<30; uint8_t *s,*d; CK(cudaMalloc(&s,N+256)); CK(cudaMalloc(&d,N+256));
  CK(cudaMemset(s,1,N+256)); CK(cudaMemset(d,0,N+256));
  int sms; cudaDeviceGetAttribute(&sms,cudaDevAttrMultiProcessorCount,0);
  const uint64_t nvec=N/16; const uint32_t tv=4096; // 64 KB tile
  bench("cudaMemcpy D2D",N,[&]{cudaMemcpyAsync(d,s,N,cudaMemcpyDeviceToDevice);});
#define P(T,U,M,B,G) bench("persist T" #T " U" #U " mode" #M " minb" #B " grid x" #G,N,[&]{k_persist<T,U,M,B><<<sms*G,T>>>((uint4*)d,(const uint4*)s,nvec,tv);});
  P(256,4,0,4,4) P(256,4,1,4,4) P(256,4,2,4,4) P(256,4,3,4,4)
  P(256,8,0,4,4) P(256,8,1,4,4) P(256,4,0,8,8) P(256,4,1,8,8) P(256,2,1,8,8) P(512,4,1,4,4) P(512,4,0,4,4) P(1024,4,1,2,2) P(128,4,1,16,16) P(256,8,3,4,4)
  bench("persist T256 U4 m1 tile 256KB",N,[&]{k_persist<256,4,1,8><<<sms*8,256>>>((uint4*)d,(const uint4*)s,nvec,16384);});
  bench("persist T256 U8 m1 tile 256KB",N,[&]{k_persist<256,8,1,4><<<sms*4,256>>>((uint4*)d,(const uint4*)s,nvec,16384);});
  bench("persist T256 U4 m1 tile 16KB",N,[&]{k_persist<256,4,1,8><<<sms*8,256>>>((uint4*)d,(const uint4*)s,nvec,1024);});
#define S(T,U,M) bench("simple T" #T " U" #U " mode" #M,N,[&]{k_simple<T,U,M><<<(unsigned)((nvec+(uint64_t)T*U-1)/((uint64_t)T*U)),T>>>((uint4*)d,(const uint4*)s,nvec);});
  S(128,4,1) S(128,4,0) S(256,4,1) S(256,8,1) S(128,8,1) S(256,4,3) S(512,4,1)
  bench("shift2 T256 U4 minb4",N,[&]{k_shift2<256,4,4><<<sms*4,256>>>((uint4*)d,(const uint4*)s,nvec-1,tv,0);});
  bench("shift2 T256 U2 minb8",N,[&]{k_shift2<256,2,8><<<sms*8,256>>>((uint4*)d,(const uint4*)s,nvec-1,tv,0);});
  bench("shiftshfl T256 U4 minb4",N,[&]{k_shiftshfl<256,4,4><<<sms*4,256>>>((uint4*)d,(const uint4*)s,nvec-1,tv,0);});
  bench("shiftshfl T256 U4 minb8",N,[&]{k_shiftshfl<256,4,8><<<sms*8,256>>>((uint4*)d,(const uint4*)s,nvec-1,tv,0);});
  bench("shiftshfl T256 U8 minb4",N,[&]{k_shiftshfl<256,8,4><<<sms*4,256>>>((uint4*)d,(const uint4*)s,nvec-1,tv,0);});
#define TM(ST,CH,G) { CK(cudaFuncSetAttribute(k_tma<ST,CH>,cudaFuncAttributeMaxDynamicSharedMemorySize,ST*CH)); \
  bench("tma stages" #ST " chunk" #CH " ctas/sm" #G,N,[&]{k_tma<ST,CH><<<sms*G,32,ST*CH>>>(d,s,N);}); }
  TM(4,16384,2) TM(4,16384,3) TM(6,16384,2) TM(3,32768,2) TM(4,8192,4) TM(8,8192,3) TM(4,32768,1) TM(6,32768,1) TM(2,65536,1) TM(3,65536,1) TM(8,16384,1) TM(12,16384,1) TM(4,49152,1)

  {
    unsigned long long* ctr; CK(cudaMalloc(&ctr,8));
    bench("dyn T256 U4 minb8",N,[&]{cudaMemsetAsync(ctr,0,8); k_dyn<256,4,8><<<sms*8,256>>>((uint4*)d,(const uint4*)s,nvec,ctr);});
    bench("dyn T512 U4 minb4",N,[&]{cudaMemsetAsync(ctr,0,8); k_dyn<512,4,4><<<sms*4,512>>>((uint4*)d,(const uint4*)s,nvec,ctr);});
  }
#define CAP(T,U,SM) { CK(cudaFuncSetAttribute(k_simple_cap<T,U>,cudaFuncAttributeMaxDynamicSharedMemorySize,SM)); \
  bench("simple_cap T" #T " U" #U " smem" #SM,N,[&]{k_simple_cap<T,U><<<(unsigned)((nvec+(uint64_t)T*U-1)/((uint64_t)T*U)),T,SM>>>((uint4*)d,(const uint4*)s,nvec);}); }
  CAP(256,4,0) CAP(256,4,28000) CAP(256,4,56000) CAP(256,4,110000) CAP(512,4,56000) CAP(512,4,110000)
#define TL(T,U) { uint64_t tb=(uint64_t)T*U*16; uint64_t nt=N/tb; std::vector<Tile32> h(nt); for(uint64_t i=0;i<nt;++i){h[i].src=(uint64_t)s+i*tb;h[i].dst=(uint64_t)d+i*tb;h[i].len=(uint32_t)tb;h[i].region=0xffffffffu;h[i].off=0;} \
  Tile32* dt; CK(cudaMalloc(&dt,nt*32)); CK(cudaMemcpy(dt,h.data(),nt*32,cudaMemcpyHostToDevice)); \
  bench("tiled T" #T " U" #U,N,[&]{k_tiled<T,U><<<(unsigned)nt,T>>>(dt);}); \
  bench("tiled_shift T" #T " U" #U,N,[&]{k_tiled_shift<T,U><<<(unsigned)nt,T>>>(dt,0);}); cudaFree(dt); }
  TL(256,4) TL(512,4) TL(256,8) TL(512,8) TL(1024,4) TL(128,4)
#define TS(CH) { CK(cudaFuncSetAttribute(k_tma_simple<CH>,cudaFuncAttributeMaxDynamicSharedMemorySize,CH)); \
  bench("tma_simple chunk" #CH,N,[&]{k_tma_simple<CH><<<(unsigned)((N+CH-1)/CH),32,CH>>>(d,s,N);}); }
  TS(8192) TS(16384) TS(32768) TS(65536)
  // correctness of the TMA variant
  CK(cudaMemset(d,0,N)); k_tma<4,16384><<<sms*2,32,4*16384>>>(d,s,N); CK(cudaDeviceSynchronize());
  std::vector<uint8_t> h(1<<20); CK(cudaMemcpy(h.data(),d+N-(1<<20),1<<20,cudaMemcpyDeviceToHost)); bool ok=true; for(auto x:h) ok&=(x==1);
  printf("tma tail check %s\n",ok?"ok":"BAD");
  return 0;
}
