// Microbenchmark 3 (round-2 preparation): do random gathers from
// shared memory (what bounds phase 1 of the banded SpMV) and random 4-byte gathers from global
// memory (what bounds phase 2: one L2 -> SM sector each) overlap inside one SM?
// One persistent CTA of 1024 threads per SM, 128 KB band in shared memory.  Warps [0, G) gather
// from global memory (16 M floats = 64 MB, L2-resident but not L1-resident), warps [G, 32) from
// the band.  Each role walks its own index stream with 128-bit index loads, 8 gathers per lane
// per step.  Reported: time of each role alone with that many warps, and of both together.  If
// the two paths are independent resources, together ~= max(alone); if they share one pipe,
// together ~= sum.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o overlap_bench overlap_bench.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("ERR %s line %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)
__device__ __forceinline__ uint64_t mix(uint64_t x){ x+=0x9E3779B97F4A7C15ull; x=(x^(x>>30))*0xBF58476D1CE4E5B9ull; x=(x^(x>>27))*0x94D049BB133111EBull; return x^(x>>31);}
__global__ void make_idx16(uint16_t* idx, size_t n){
  for(size_t i=blockIdx.x*(size_t)blockDim.x+threadIdx.x;i<n;i+=(size_t)gridDim.x*blockDim.x) idx[i]=(uint16_t)(mix(i)&0x7fff);
}
__global__ void make_idx32(uint32_t* idx, size_t n, uint32_t range){
  for(size_t i=blockIdx.x*(size_t)blockDim.x+threadIdx.x;i<n;i+=(size_t)gridDim.x*blockDim.x) idx[i]=(uint32_t)(mix(i^0xabcdef)%range);
}
constexpr int kThreads=1024, kBand=1<<15;
// mode bit 0: run the shared-memory role, bit 1: run the global role
__global__ void __launch_bounds__(kThreads,1) overlap_k(const uint16_t* __restrict__ sidx, size_t ns, const uint32_t* __restrict__ gidx, size_t ng,
                                                        const float* __restrict__ x, int gwarps, int mode, float* out){
  extern __shared__ float band[];
  for(int i=threadIdx.x;i<kBand;i+=kThreads) band[i]=x[i];
  __syncthreads();
  const int warp=threadIdx.x>>5, lane=threadIdx.x&31;
  float acc=0.f;
  if(warp<gwarps){
    if(mode&2){
      const size_t nw=(size_t)gridDim.x*gwarps, w=(size_t)blockIdx.x*gwarps+warp;
      // every global warp walks its own contiguous slice in steps of 256 indices
      const size_t per=(ng/nw)&~(size_t)255;
      const uint32_t* p=gidx+w*per;
      for(size_t s=0;s+256<=per;s+=256){
        const uint4 a=*reinterpret_cast<const uint4*>(p+s+lane*8), b=*reinterpret_cast<const uint4*>(p+s+lane*8+4);
        float v0=__ldg(x+a.x),v1=__ldg(x+a.y),v2=__ldg(x+a.z),v3=__ldg(x+a.w),v4=__ldg(x+b.x),v5=__ldg(x+b.y),v6=__ldg(x+b.z),v7=__ldg(x+b.w);
        acc+=((v0+v1)+(v2+v3))+((v4+v5)+(v6+v7));
      }
    }
  } else if(mode&1){
    const int sw=32-gwarps;
    const size_t nw=(size_t)gridDim.x*sw, w=(size_t)blockIdx.x*sw+(warp-gwarps);
    const size_t per=(ns/nw)&~(size_t)255;
    const uint16_t* p=sidx+w*per;
    for(size_t s=0;s+256<=per;s+=256){
      const uint4 c=*reinterpret_cast<const uint4*>(p+s+lane*8);
      float v0=band[c.x&0x7fff],v1=band[(c.x>>16)&0x7fff],v2=band[c.y&0x7fff],v3=band[(c.y>>16)&0x7fff];
      float v4=band[c.z&0x7fff],v5=band[(c.z>>16)&0x7fff],v6=band[c.w&0x7fff],v7=band[(c.w>>16)&0x7fff];
      acc+=((v0+v1)+(v2+v3))+((v4+v5)+(v6+v7));
    }
  }
  if(acc==123.456f) out[0]=acc;
}
// Variant 2: the two roles as two kernels on two streams - a 768-thread shared-memory kernel
// (band + nothing else, <= 64 registers) and a 256-thread global-gather kernel that has to fit
// into what is left of each SM (16 K registers, no shared memory).
__global__ void __launch_bounds__(768,1) smem_only_k(const uint16_t* __restrict__ sidx, size_t ns, const float* __restrict__ x, float* out){
  extern __shared__ float band[];
  for(int i=threadIdx.x;i<kBand;i+=768) band[i]=x[i];
  __syncthreads();
  const int warp=threadIdx.x>>5, lane=threadIdx.x&31;
  const size_t nw=(size_t)gridDim.x*24, w=(size_t)blockIdx.x*24+warp;
  const size_t per=(ns/nw)&~(size_t)255;
  const uint16_t* p=sidx+w*per;
  float acc=0.f;
  for(size_t s=0;s+256<=per;s+=256){
    const uint4 c=*reinterpret_cast<const uint4*>(p+s+lane*8);
    float v0=band[c.x&0x7fff],v1=band[(c.x>>16)&0x7fff],v2=band[c.y&0x7fff],v3=band[(c.y>>16)&0x7fff];
    float v4=band[c.z&0x7fff],v5=band[(c.z>>16)&0x7fff],v6=band[c.w&0x7fff],v7=band[(c.w>>16)&0x7fff];
    acc+=((v0+v1)+(v2+v3))+((v4+v5)+(v6+v7));
  }
  if(acc==123.456f) out[0]=acc;
}
__global__ void __launch_bounds__(256) global_only_k(const uint32_t* __restrict__ gidx, size_t ng, const float* __restrict__ x, float* out){
  const int warp=threadIdx.x>>5, lane=threadIdx.x&31;
  const size_t nw=(size_t)gridDim.x*8, w=(size_t)blockIdx.x*8+warp;
  const size_t per=(ng/nw)&~(size_t)255;
  const uint32_t* p=gidx+w*per;
  float acc=0.f;
  for(size_t s=0;s+256<=per;s+=256){
    const uint4 a=*reinterpret_cast<const uint4*>(p+s+lane*8), b=*reinterpret_cast<const uint4*>(p+s+lane*8+4);
    float v0=__ldg(x+a.x),v1=__ldg(x+a.y),v2=__ldg(x+a.z),v3=__ldg(x+a.w),v4=__ldg(x+b.x),v5=__ldg(x+b.y),v6=__ldg(x+b.z),v7=__ldg(x+b.w);
    acc+=((v0+v1)+(v2+v3))+((v4+v5)+(v6+v7));
  }
  if(acc==123.456f) out[1]=acc;
}
template<typename F> float timeit(F f,int reps=5){ cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b); f(); CK(cudaDeviceSynchronize()); float best=1e30f; for(int r=0;r<reps;r++){ cudaEventRecord(a); f(); cudaEventRecord(b); CK(cudaEventSynchronize(b)); float ms; cudaEventElapsedTime(&ms,a,b); if(ms<best) best=ms;} return best; }
int main(){
  // work sized like one SpMV of the scale-24 graph: 215 M shared gathers, 53 M global gathers
  const size_t ns=(size_t)215<<20, ng=(size_t)53<<20; const uint32_t range=1u<<24;
  uint16_t* sidx; uint32_t* gidx; float *x,*out;
  CK(cudaMalloc(&sidx,ns*2)); CK(cudaMalloc(&gidx,ng*4)); CK(cudaMalloc(&x,(size_t)range*4)); CK(cudaMalloc(&out,64));
  CK(cudaMemset(x,0,(size_t)range*4));
  make_idx16<<<1184,256>>>(sidx,ns); make_idx32<<<1184,256>>>(gidx,ng,range); CK(cudaDeviceSynchronize());
  int sms=0; CK(cudaDeviceGetAttribute(&sms,cudaDevAttrMultiProcessorCount,0));
  const size_t smem=(size_t)kBand*4;
  CK(cudaFuncSetAttribute(overlap_k,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)smem));
  printf("SMs %d; %zu M shared gathers, %zu M global gathers over %u floats\n",sms,ns>>20,ng>>20,range);
  printf("%7s %12s %12s %12s %10s\n","gwarps","shared ms","global ms","together ms","overlap");
  for(int g: {2,4,6,8,12,16}){
    float ts=timeit([&]{overlap_k<<<sms,kThreads,smem>>>(sidx,ns,gidx,ng,x,g,1,out);});
    float tg=timeit([&]{overlap_k<<<sms,kThreads,smem>>>(sidx,ns,gidx,ng,x,g,2,out);});
    float tb=timeit([&]{overlap_k<<<sms,kThreads,smem>>>(sidx,ns,gidx,ng,x,g,3,out);});
    CK(cudaGetLastError());
    // 1.0 = perfect overlap (together = max), 0.0 = none (together = sum)
    float hi=ts>tg?ts:tg, lo=ts>tg?tg:ts;
    printf("%7d %12.3f %12.3f %12.3f %9.2f\n",g,ts,tg,tb,lo>0?(ts+tg-tb)/lo:0.f);
  }
  {
    cudaStream_t sa,sb; CK(cudaStreamCreateWithFlags(&sa,cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&sb,cudaStreamNonBlocking));
    CK(cudaFuncSetAttribute(smem_only_k,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)smem));
    int occ_a=0,occ_b=0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_a,smem_only_k,768,smem)); CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_b,global_only_k,256,0));
    cudaEvent_t e0,e1,e2; cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
    auto both=[&](bool a,bool b)->float{
      float best=1e30f;
      for(int r=0;r<6;r++){
        CK(cudaDeviceSynchronize());
        cudaEventRecord(e0,sa); cudaStreamWaitEvent(sb,e0,0);
        if(a) smem_only_k<<<sms,768,smem,sa>>>(sidx,ns,x,out);
        if(b) global_only_k<<<sms,256,0,sb>>>(gidx,ng,x,out);
        cudaEventRecord(e1,sb); cudaStreamWaitEvent(sa,e1,0); cudaEventRecord(e2,sa);
        CK(cudaEventSynchronize(e2)); float ms; cudaEventElapsedTime(&ms,e0,e2); if(r>0&&ms<best) best=ms;
      }
      return best; };
    float ta=both(true,false), tb2=both(false,true), tab=both(true,true);
    float lo=ta>tb2?tb2:ta;
    printf("two kernels on two streams (768-thread shared-memory kernel, occupancy %d/SM; 256-thread global kernel, occupancy %d/SM alone):\n",occ_a,occ_b);
    printf("  shared alone %.3f ms, global alone (1 CTA/SM) %.3f ms, together %.3f ms, overlap %.2f\n",ta,tb2,tab,lo>0?(ta+tb2-tab)/lo:0.f);
    CK(cudaGetLastError());
  }
  printf("(shared gathers/clk/SM alone at 32 warps and global sectors/clk/SM follow from ns, ng, the times and 1.9 GHz)\n");
  return 0;
}
