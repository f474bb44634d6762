#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
__device__ __forceinline__ float ex2f(float x){float y; asm volatile("ex2.approx.ftz.f32 %0, %1;":"=f"(y):"f"(x)); return y;}
__device__ __forceinline__ unsigned ex2h2(unsigned x){unsigned y; asm volatile("ex2.approx.f16x2 %0, %1;":"=r"(y):"r"(x)); return y;}
__device__ __forceinline__ unsigned ex2b2(unsigned x){unsigned y; asm volatile("ex2.approx.ftz.bf16x2 %0, %1;":"=r"(y):"r"(x)); return y;}
template<int MODE> __global__ void k(float* out, int iters){
  float a[8]; unsigned u[8];
  for(int i=0;i<8;i++){a[i]=-0.001f*(threadIdx.x+i); u[i]=0xb800b800u+i+threadIdx.x;}
  for(int it=0;it<iters;it++){
#pragma unroll
    for(int i=0;i<8;i++){
      if(MODE==0) a[i]=ex2f(a[i]);
      if(MODE==1) u[i]=ex2h2(u[i]);
      if(MODE==2) u[i]=ex2b2(u[i]);
    }
  }
  float s=0; for(int i=0;i<8;i++) s+=a[i]+__uint_as_float(u[i]);
  out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
int main(){
  float* d; cudaMalloc(&d, 148*8*256*4);
  cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int iters=20000;
  for(int mode=0;mode<3;mode++){
    for(int rep=0;rep<2;rep++){
      cudaEventRecord(e0);
      if(mode==0) k<0><<<148*4,256>>>(d,iters);
      if(mode==1) k<1><<<148*4,256>>>(d,iters);
      if(mode==2) k<2><<<148*4,256>>>(d,iters);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms,e0,e1);
      double ops=(double)148*4*256*8*iters;
      if(rep) printf("mode %d: %.3f ms, %.1f G instr-lanes/s => %.2f lanes/clk/SM @1.9GHz\n",mode,ms,ops/ms/1e6,ops/ms/1e6/148/1.9);
    }
  }
  return 0;
}
