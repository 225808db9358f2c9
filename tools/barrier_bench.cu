// Micro-benchmark: grid-barrier variants for the persistent tracking kernel (148 SMs x k CTAs).
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;
__device__ __forceinline__ unsigned ld_acq(const unsigned* p){unsigned v;asm volatile("ld.acquire.gpu.global.u32 %0,[%1];":"=r"(v):"l"(p):"memory");return v;}
__device__ __forceinline__ unsigned ld_rlx(const unsigned* p){unsigned v;asm volatile("ld.relaxed.gpu.global.u32 %0,[%1];":"=r"(v):"l"(p):"memory");return v;}
__device__ __forceinline__ void st_rel(unsigned* p,unsigned v){asm volatile("st.release.gpu.global.u32 [%0],%1;"::"l"(p),"r"(v):"memory");}
__device__ __forceinline__ unsigned atom_ar(unsigned* p,unsigned v){unsigned o;asm volatile("atom.add.acq_rel.gpu.global.u32 %0,[%1],%2;":"=r"(o):"l"(p),"r"(v):"memory");return o;}
__device__ __forceinline__ unsigned atom_rel(unsigned* p,unsigned v){unsigned o;asm volatile("atom.add.release.gpu.global.u32 %0,[%1],%2;":"=r"(o):"l"(p),"r"(v):"memory");return o;}

template<int MODE> __global__ void __launch_bounds__(256,2) k(unsigned* sync, int nbar, float* sink, int sleep_ns){
  __shared__ int s_last;
  unsigned gen=0; const unsigned n=gridDim.x;
  unsigned* arrive=sync; unsigned* g=sync+ (MODE==3?32:1);
  cg::grid_group grid = cg::this_grid();
  for(int b=0;b<nbar;++b){
    if(MODE==2){ grid.sync(); continue; }
    __syncthreads();
    if(threadIdx.x==0){ __threadfence(); unsigned t=atom_ar(arrive,1u); s_last=(t==n*(gen+1u)-1u); }
    __syncthreads();
    if(s_last){ __syncthreads(); if(threadIdx.x==0){ __threadfence(); st_rel(g,gen+1u);} }
    else {
      if(threadIdx.x==0){
        if(MODE==0){ while(ld_acq(g)<=gen) __nanosleep(20); }
        else { while(ld_rlx(g)<=gen) { if(sleep_ns) __nanosleep(sleep_ns);} __threadfence(); }
      }
      __syncthreads();
    }
    ++gen;
  }
  if(threadIdx.x==0 && blockIdx.x==0) sink[0]=gen;
}
template<int MODE> float run(int grid,int nbar,unsigned* sync,float* sink,int sleep_ns){
  cudaMemset(sync,0,1024);
  void* args[]={&sync,&nbar,&sink,&sleep_ns};
  cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaLaunchCooperativeKernel((void*)k<MODE>,dim3(grid),dim3(256),args,0,0); cudaDeviceSynchronize();
  cudaMemset(sync,0,1024);
  cudaEventRecord(a); cudaLaunchCooperativeKernel((void*)k<MODE>,dim3(grid),dim3(256),args,0,0); cudaEventRecord(b); cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms,a,b);
  cudaError_t e=cudaGetLastError(); if(e!=cudaSuccess) printf("err %s\n",cudaGetErrorString(e));
  return ms*1000.f/nbar;
}
int main(){
  unsigned* sync; float* sink; cudaMalloc(&sync,1024); cudaMalloc(&sink,4);
  int sms; cudaDeviceGetAttribute(&sms,cudaDevAttrMultiProcessorCount,0);
  for(int per=1;per<=2;++per){ int grid=sms*per; int nbar=200;
    printf("grid %d: acquire-poll+sleep20 %.2f us | relaxed-poll nosleep %.2f | relaxed sleep32 %.2f | relaxed sleep100 %.2f | relaxed gen-on-own-line %.2f | cg grid.sync %.2f us/barrier\n",grid,
      run<0>(grid,nbar,sync,sink,20), run<1>(grid,nbar,sync,sink,0), run<1>(grid,nbar,sync,sink,32), run<1>(grid,nbar,sync,sink,100), run<3>(grid,nbar,sync,sink,0), run<2>(grid,nbar,sync,sink,0));
  }
  return 0;
}
