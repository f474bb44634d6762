// Peer windows and the in-graph cond/uncond exchange of the GPU-pair split (BASELINE config 5; SURVEY §8e:
// "direct peer store into a symmetric buffer + flag").  The reference has no counterpart (it never shards inference:
// runpod-worker/rp_handler.py:15,36); the split halves the UNet batch of pipelines/sd_unified_pipeline.py:467-482 across
// two GPUs, and this file is the one exchange per sampler step that rejoins the halves before the CFG combine (:484-486).
//
// One process per GPU: a "window" is a cudaMalloc'ed buffer exported with cudaIpcGetMemHandle; the peer maps it with
// cudaIpcOpenMemHandle and from then on stores into it over NVLink from inside its own kernels.  Per step, inside the
// captured step graph (no host involvement, no NCCL launch):
//     conv_out writes this rank's eps half into its own window slot
//     PAIR_EXCHANGE   waits until the peer has consumed the previous message (ack), copies the slot into the SAME slot of
//                     the peer's window with 16-byte stores, __threadfence_system, the last CTA releases the peer's flag
//                     (= step sequence number), then one thread spins on the LOCAL flag until the peer's half has landed
//     CFG_STEP        reads both halves from the local window
//     STEP_ADVANCE    also acknowledges to the peer (its slot may be overwritten) and bumps the sequence number
// Every spin has a time-out (a peer that died must not hang this GPU): on expiry the kernel records an error code in the
// state block and carries on; the host checks it at the end of the trajectory and raises.
#include <new>

#include "common.cuh"
#include "internal.h"

namespace b200sd {
namespace {

__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(int* p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// sync block (in the window, written by the PEER): [0] flag = sequence number of the newest message that has landed,
//                                                  [1] ack  = sequence number of the newest message the peer has consumed
// state block (local): [0] seq = steps completed, [1] err (0 ok, 1 ack time-out, 2 flag time-out), [2] CTA arrival counter
struct PXParams {
  const uint4* src; uint4* dst_remote; long long vecs; const int* sync_local; int* sync_remote; int* state; long long timeout_ns;
};

__device__ __forceinline__ bool spin_until(const int* p, int want, long long timeout_ns) {
  const uint64_t t0 = global_timer_ns();
  while (ld_acquire_sys(p) < want) {
    if (static_cast<long long>(global_timer_ns() - t0) > timeout_ns) return false;
    __nanosleep(64);
  }
  return true;
}

__global__ void __launch_bounds__(256) pair_exchange_kernel(const PXParams p) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ int s_last;
  const int seq = p.state[0] + 1;
  if (threadIdx.x == 0) {
    // the peer must have consumed message seq-1 before its slot is overwritten (it has, a whole UNet forward ago)
    if (!spin_until(p.sync_local + 1, seq - 1, p.timeout_ns)) atomicMax(p.state + 1, 1);
  }
  __syncthreads();
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < p.vecs;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    p.dst_remote[i] = p.src[i];
  __threadfence_system();  // this thread's peer stores are visible system-wide before the arrival below
  __syncthreads();
  if (threadIdx.x == 0) {
    const int prev = atomicAdd(p.state + 2, 1);
    s_last = (prev == static_cast<int>(gridDim.x) - 1);
    if (s_last) {
      p.state[2] = 0;
      __threadfence_system();
      st_release_sys(p.sync_remote, seq);                       // the peer's flag: my half of step `seq` is in its window
      if (!spin_until(p.sync_local, seq, p.timeout_ns)) atomicMax(p.state + 1, 2);  // ... and its half is in mine
    }
  }
}

struct PXPlan { PXParams p; int grid; };
int launch_pair_exchange(const void* plan, cudaStream_t s) {
  const PXPlan& pl = *static_cast<const PXPlan*>(plan);
  B200SD_CUDA(launch_kernel(pair_exchange_kernel, dim3(pl.grid), dim3(256), 0, s, 1, pl.p));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

int plan_pair_exchange(const b200sd_pair_exchange& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.src && g.dst_remote && g.sync_local && g.sync_remote && g.state, "pair_exchange: NULL pointer");
  B200SD_CHECK_ARG(g.bytes >= 16 && g.bytes % 16 == 0, "pair_exchange: bytes=%lld must be a positive multiple of 16", (long long)g.bytes);
  B200SD_CHECK_ARG(((reinterpret_cast<uintptr_t>(g.src) | reinterpret_cast<uintptr_t>(g.dst_remote)) & 15) == 0,
                   "pair_exchange: buffers must be 16-byte aligned");
  PXPlan* pl = new (std::nothrow) PXPlan();
  if (!pl) return fail(B200SD_ERR_CUDA, "out of host memory");
  pl->p = PXParams{static_cast<const uint4*>(g.src), static_cast<uint4*>(g.dst_remote), g.bytes / 16, g.sync_local, g.sync_remote,
                   g.state, static_cast<long long>(g.timeout_ms > 0 ? g.timeout_ms : 20000) * 1000000LL};
  // 32 KiB (B = 1, 512^2) .. 8 MiB (B = 64, 1024^2): a CTA moves 16 KiB per pass; a few CTAs saturate one NVLink direction
  pl->grid = static_cast<int>(std::max<long long>(1, std::min<long long>((g.bytes + 16383) / 16384, 32)));
  out->kind = B200SD_OP_PAIR_EXCHANGE;
  out->plan = pl;
  out->launch = launch_pair_exchange;
  out->destroy = destroy_plan<PXPlan>;
  out->launches = 1;
  return 0;
}

}  // namespace b200sd

using namespace b200sd;

extern "C" {

int b200sd_peer_alloc(int device, size_t bytes, void** ptr, unsigned char* handle) {
  if (!ptr || !handle || bytes == 0) return fail(B200SD_ERR_INVALID, "peer_alloc: bad arguments");
  int rc = b200sd_device_check(device);
  if (rc) return rc;
  B200SD_CUDA(cudaSetDevice(device));
  void* p = nullptr;
  B200SD_CUDA(cudaMalloc(&p, bytes));
  cudaError_t e = cudaMemset(p, 0, bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return fail(B200SD_ERR_CUDA, "peer_alloc: %s", cudaGetErrorString(e));
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == B200SD_IPC_HANDLE_BYTES, "IPC handle size");
  memcpy(handle, &h, sizeof(h));
  *ptr = p;
  return 0;
}

int b200sd_peer_open(int device, const unsigned char* handle, void** remote_ptr) {
  if (!handle || !remote_ptr) return fail(B200SD_ERR_INVALID, "peer_open: bad arguments");
  B200SD_CUDA(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return fail(B200SD_ERR_CUDA, "peer_open: cudaIpcOpenMemHandle failed: %s (no peer access between the two GPUs?)", cudaGetErrorString(e));
  }
  *remote_ptr = p;
  return 0;
}

int b200sd_peer_close(void* remote_ptr) {
  if (!remote_ptr) return 0;
  B200SD_CUDA(cudaIpcCloseMemHandle(remote_ptr));
  return 0;
}

int b200sd_peer_free(void* ptr) {
  if (!ptr) return 0;
  B200SD_CUDA(cudaFree(ptr));
  return 0;
}

}  // extern "C"
