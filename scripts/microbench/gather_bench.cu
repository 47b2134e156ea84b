// Microbenchmark: random-gather throughput on B200 by load flavour (decides the design of the
// gather phase of the pull kernel).  nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("ERR %s line %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)

__device__ __forceinline__ uint64_t mix(uint64_t x){ x+=0x9E3779B97F4A7C15ull; x=(x^(x>>30))*0xBF58476D1CE4E5B9ull; x=(x^(x>>27))*0x94D049BB133111EBull; return x^(x>>31);}

__global__ void make_idx(int32_t* idx, size_t n, uint32_t range, int mode){
  for(size_t i=blockIdx.x*(size_t)blockDim.x+threadIdx.x;i<n;i+=(size_t)gridDim.x*blockDim.x){
    uint64_t r=mix(i);
    uint32_t v;
    if(mode==0) v=(uint32_t)(r%range);               // uniform
    else { // power-law-ish: skewed to low ids (RMAT-like marginal): product of bits with p=0.24
      uint32_t x=0; int bits=0; uint32_t rr=range; while(rr>1){bits++; rr>>=1;}
      for(int b=0;b<bits;b++){ uint64_t q=mix(r+b*7919); if((q&0xffff) < (uint64_t)(0.24*65536)) x|=(1u<<b);} v=x%range; }
    idx[i]=(int32_t)v;
  }
}
template<int MODE> __device__ __forceinline__ float ld(const float* p, cudaTextureObject_t tex, int i){
  if(MODE==0) return __ldg(p+i);
  if(MODE==1) return __ldcg(p+i);
  if(MODE==2){ float v; asm volatile("ld.global.nc.L1::no_allocate.f32 %0,[%1];":"=f"(v):"l"(p+i)); return v;}
  if(MODE==3) return tex1Dfetch<float>(tex,i);
  if(MODE==4){ float v; asm volatile("ld.global.nc.L1::evict_last.f32 %0,[%1];":"=f"(v):"l"(p+i)); return v;}
  return 0.f;
}
// each thread: U independent gathers per step, indices streamed coalesced as int4
template<int MODE,int U> __global__ void __launch_bounds__(512) gather_k(const int32_t* __restrict__ idx,const float* __restrict__ x,cudaTextureObject_t tex,size_t n,float* out){
  float acc=0.f;
  size_t stride=(size_t)gridDim.x*blockDim.x*4*U;
  for(size_t base=(blockIdx.x*(size_t)blockDim.x+threadIdx.x)*4; base+ (size_t)(U-1)*gridDim.x*blockDim.x*4+3<n; base+=stride){
    int4 id[U];
#pragma unroll
    for(int u=0;u<U;u++) id[u]=*reinterpret_cast<const int4*>(idx+base+(size_t)u*gridDim.x*blockDim.x*4);
    float v[U*4];
#pragma unroll
    for(int u=0;u<U;u++){ v[4*u]=ld<MODE>(x,tex,id[u].x); v[4*u+1]=ld<MODE>(x,tex,id[u].y); v[4*u+2]=ld<MODE>(x,tex,id[u].z); v[4*u+3]=ld<MODE>(x,tex,id[u].w);}
#pragma unroll
    for(int u=0;u<4*U;u++) acc+=v[u];
  }
  if(acc==123.456f) out[0]=acc;
}
// shared-memory gather: indices < SM_N served from smem copy
template<int U> __global__ void __launch_bounds__(512) gather_smem_k(const int32_t* __restrict__ idx,const float* __restrict__ x,size_t n,int sm_n,float* out){
  extern __shared__ float sx[];
  for(int i=threadIdx.x;i<sm_n;i+=blockDim.x) sx[i]=x[i];
  __syncthreads();
  float acc=0.f;
  size_t stride=(size_t)gridDim.x*blockDim.x*4*U;
  for(size_t base=(blockIdx.x*(size_t)blockDim.x+threadIdx.x)*4; base+(size_t)(U-1)*gridDim.x*blockDim.x*4+3<n; base+=stride){
    int4 id[U];
#pragma unroll
    for(int u=0;u<U;u++) id[u]=*reinterpret_cast<const int4*>(idx+base+(size_t)u*gridDim.x*blockDim.x*4);
#pragma unroll
    for(int u=0;u<U;u++){
      int a=id[u].x,b=id[u].y,c=id[u].z,d=id[u].w;
      acc+= (a<sm_n? sx[a]:__ldg(x+a)) + (b<sm_n? sx[b]:__ldg(x+b)) + (c<sm_n? sx[c]:__ldg(x+c)) + (d<sm_n? sx[d]:__ldg(x+d));
    }
  }
  if(acc==123.456f) out[0]=acc;
}
template<typename F> float timeit(F f,int reps=5){ cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b); f(); cudaDeviceSynchronize(); float best=1e30f; for(int r=0;r<reps;r++){ cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms,a,b); if(ms<best) best=ms;} return best; }

int main(int argc,char**argv){
  size_t n= (size_t)1<<28; // 268M gathers
  int32_t* idx; float* x; float* out; CK(cudaMalloc(&idx,n*4)); CK(cudaMalloc(&out,64));
  const char* names[]={"ldg","ldcg","nc.no_allocate","tex1Dfetch","nc.L1evict_last"};
  for(int rlog: {20,24,26}){
    uint32_t range=1u<<rlog;
    CK(cudaMalloc(&x,(size_t)range*4)); CK(cudaMemset(x,0,(size_t)range*4));
    cudaResourceDesc rd{}; rd.resType=cudaResourceTypeLinear; rd.res.linear.devPtr=x; rd.res.linear.desc=cudaCreateChannelDesc<float>(); rd.res.linear.sizeInBytes=(size_t)range*4;
    cudaTextureDesc td{}; td.readMode=cudaReadModeElementType; cudaTextureObject_t tex; CK(cudaCreateTextureObject(&tex,&rd,&td,nullptr));
    for(int mode=0;mode<2;mode++){
      make_idx<<<148*8,256>>>(idx,n,range,mode); CK(cudaDeviceSynchronize());
      printf("== x range 2^%d (%zu MB), index distribution %s\n",rlog,(size_t)range*4>>20,mode?"skewed(p=.24/bit)":"uniform");
      int grid=148*2;
#define RUN(M,U) { float ms=timeit([&]{gather_k<M,U><<<grid,512>>>(idx,x,tex,n,out);}); printf("  %-16s U=%d : %.3f ms  %.1f Ggather/s  (%.2f /clk/SM @1.9GHz)\n",names[M],U*4,ms,n/ms/1e6,n/ms/1e6/148/1.9); }
      RUN(0,1) RUN(0,2) RUN(0,4) RUN(1,2) RUN(2,2) RUN(3,2) RUN(3,4) RUN(4,2)
      for(int smn: {8192,32768,49152}){
        CK(cudaFuncSetAttribute(gather_smem_k<2>,cudaFuncAttributeMaxDynamicSharedMemorySize,smn*4));
        float ms=timeit([&]{gather_smem_k<2><<<148,512*2>0?512:512,smn*4>>>(idx,x,n,smn,out);});
        printf("  smem-cache %6d entries (1 CTA/SM x512): %.3f ms %.1f Ggather/s\n",smn,ms,n/ms/1e6);
      }
      { // pure streaming baseline: read idx only
        float ms=timeit([&]{gather_smem_k<2><<<148*2,512,4*4>>>(idx,x,0,4,out);}); (void)ms; }
    }
    cudaDestroyTextureObject(tex); cudaFree(x);
  }
  return 0;
}
