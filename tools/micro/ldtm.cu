// TMEM read bandwidth microbenchmark: how many bytes per clock can tcgen05.ld deliver per SM, as a function of
// the number of warps (4 per CTA, one per TMEM lane quarter) and CTAs per SM?  Settles whether the attention
// kernel's S read-out (32 KB per 128x64 block) is the bound.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ldtm ldtm.cu && ./ldtm
#include <cstdio>
#include <cstdint>
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols) {
  uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(slot));
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(a), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void ld32(uint32_t addr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,"
      "%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
        "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(addr));
}
__device__ __forceinline__ void ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// WARPS: 4 (one warpgroup) or 8 (two warpgroups reading the same lanes, different columns)
__global__ void k(uint32_t* out, int iters, uint32_t cols) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(&slot, cols);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = slot + (static_cast<uint32_t>((warp & 3) * 32) << 16) + (warp >> 2) * 64;
  uint32_t acc = 0;
  uint32_t a[32], b[32];
  for (int it = 0; it < iters; ++it) {
    ld32(base, a);       // two x32 loads in flight, like the softmax pass
    ld32(base + 32, b);
    ld_wait();
#pragma unroll
    for (int i = 0; i < 32; ++i) acc += a[i] ^ b[i];
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  __syncthreads();
  if (warp == 0) tmem_dealloc(slot, cols);
}
int main() {
  uint32_t* d;
  cudaMalloc(&d, 148 * 8 * 256 * 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int iters = 20000;
  int clk_khz = 0;
  cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  const int cfgs[][3] = {{1, 128, 512}, {2, 128, 256}, {4, 128, 128}, {1, 256, 512}, {2, 256, 256}, {4, 256, 128}};
  for (auto& c : cfgs) {
    const int per_sm = c[0], threads = c[1], cols = c[2];
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      k<<<148 * per_sm, threads>>>(d, iters, cols);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      const double bytes = static_cast<double>(148) * per_sm * threads * 64 * 4 * iters;
      if (rep)
        printf("%d CTA/SM x %d thr: %.3f ms  %.1f GB/s  => %.1f B/clk/SM at %.2f GHz (err %s)\n", per_sm, threads, ms, bytes / ms / 1e6,
               bytes / ms / 1e6 / 148 / (clk_khz / 1e6), clk_khz / 1e6, cudaGetErrorString(cudaGetLastError()));
    }
  }
  return 0;
}
