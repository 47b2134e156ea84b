// Microbenchmark 2: random 4-byte gathers from (a) local shared memory, (b) distributed shared
// memory of a thread-block cluster, (c) a mix of DSMEM + global, on B200.
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
namespace cg = cooperative_groups;
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("ERR %s line %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)
__device__ __forceinline__ uint64_t mix(uint64_t x){ x+=0x9E3779B97F4A7C15ull; x=(x^(x>>30))*0xBF58476D1CE4E5B9ull; x=(x^(x>>27))*0x94D049BB133111EBull; return x^(x>>31);}
__global__ void make_idx(int32_t* idx, size_t n, uint32_t range){
  for(size_t i=blockIdx.x*(size_t)blockDim.x+threadIdx.x;i<n;i+=(size_t)gridDim.x*blockDim.x) idx[i]=(int32_t)(mix(i)%range);
}
__device__ __forceinline__ float ld_dsmem(uint32_t addr){ float v; asm volatile("ld.shared::cluster.f32 %0,[%1];":"=f"(v):"r"(addr)); return v; }
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank){ uint32_t r; asm volatile("mapa.shared::cluster.u32 %0,%1,%2;":"=r"(r):"r"(addr),"r"(rank)); return r; }

// PER = entries held per CTA (power of two); cache covers ids [0, PER*cluster_size); ids beyond go to global
template<int THREADS,int LOGPER> __global__ void __launch_bounds__(THREADS) dsmem_k(const int32_t* __restrict__ idx,const float* __restrict__ x,size_t n,int cached,float* out){
  extern __shared__ float sx[];
  cg::cluster_group cl=cg::this_cluster();
  const uint32_t crank=cl.block_rank(), csize=cl.num_blocks();
  constexpr int PER=1<<LOGPER;
  for(int i=threadIdx.x;i<PER;i+=THREADS) sx[i]=x[crank*PER+i];
  cl.sync();
  const uint32_t base_addr=(uint32_t)__cvta_generic_to_shared(sx);
  float acc=0.f;
  const size_t gthreads=(size_t)gridDim.x*THREADS;
  for(size_t b=(blockIdx.x*(size_t)THREADS+threadIdx.x)*4; b+gthreads*4+3<n; b+=gthreads*8){
    int4 i0=*reinterpret_cast<const int4*>(idx+b); int4 i1=*reinterpret_cast<const int4*>(idx+b+gthreads*4);
    int id[8]={i0.x,i0.y,i0.z,i0.w,i1.x,i1.y,i1.z,i1.w};
    float v[8];
#pragma unroll
    for(int k=0;k<8;k++){
      int j=id[k];
      if(j<cached){ uint32_t r=(uint32_t)j>>LOGPER; uint32_t a=mapa(base_addr+((j&(PER-1))<<2), r); v[k]=ld_dsmem(a);} else v[k]=__ldg(x+j);
    }
#pragma unroll
    for(int k=0;k<8;k++) acc+=v[k];
  }
  if(acc==123.456f) out[0]=acc;
  cl.sync();
  (void)csize;
}
template<typename F> float timeit(F f,int reps=5){ cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b); f(); CK(cudaDeviceSynchronize()); float best=1e30f; for(int r=0;r<reps;r++){ cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms,a,b); if(ms<best) best=ms;} return best; }

template<int THREADS,int LOGPER> void run(int csize,const int32_t* idx,const float* x,size_t n,uint32_t range,float* out,const char* tag){
  auto k=dsmem_k<THREADS,LOGPER>;
  size_t smem=(size_t)(1<<LOGPER)*4;
  CK(cudaFuncSetAttribute(k,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)smem));
  if(csize>8) CK(cudaFuncSetAttribute(k,cudaFuncAttributeNonPortableClusterSizeAllowed,1));
  cudaLaunchConfig_t cfg{}; cfg.blockDim=dim3(THREADS); cfg.dynamicSmemBytes=smem;
  int nblk=(148/csize)*csize; cfg.gridDim=dim3(nblk);
  cudaLaunchAttribute at[1]; at[0].id=cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x=csize; at[0].val.clusterDim.y=1; at[0].val.clusterDim.z=1; cfg.attrs=at; cfg.numAttrs=1;
  int cached=(1<<LOGPER)*csize; if((uint32_t)cached>range) cached=range;
  cudaError_t e=cudaLaunchKernelEx(&cfg,k,idx,x,n,cached,out); if(e!=cudaSuccess){printf("  %s cluster=%d launch failed: %s\n",tag,csize,cudaGetErrorString(e)); cudaGetLastError(); return;}
  float ms=timeit([&]{cudaLaunchKernelEx(&cfg,k,idx,x,n,cached,out);});
  printf("  %-10s threads=%d cluster=%2d per-CTA %3d KB cached=%7d of range %9u (%.0f%% hits) blocks=%d: %.3f ms %.1f Ggather/s (%.2f/clk/SM)\n",tag,THREADS,csize,(int)(smem>>10),cached,range,100.0*cached/range,nblk,ms,n/ms/1e6,n/ms/1e6/nblk/1.9);
}
int main(){
  size_t n=(size_t)1<<27; int32_t* idx; float* x; float* out; CK(cudaMalloc(&idx,n*4)); CK(cudaMalloc(&out,64)); CK(cudaMalloc(&x,(size_t)1<<28)); CK(cudaMemset(x,0,(size_t)1<<28));
  for(uint32_t range: {1u<<15,1u<<18,1u<<19,1u<<20,1u<<24}){
    make_idx<<<148*8,256>>>(idx,n,range); CK(cudaDeviceSynchronize());
    printf("== uniform indices over %u entries\n",range);
    run<1024,15>(1,idx,x,n,range,out,"local/mix");
    run<1024,15>(2,idx,x,n,range,out,"dsmem");
    run<1024,15>(4,idx,x,n,range,out,"dsmem");
    run<1024,15>(8,idx,x,n,range,out,"dsmem");
    run<1024,15>(16,idx,x,n,range,out,"dsmem");
    run<512,15>(8,idx,x,n,range,out,"dsmem");
    run<1024,13>(8,idx,x,n,range,out,"dsmem");
  }
  return 0;
}
