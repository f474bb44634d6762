// HBM-bound kernels of the denoising loop (SURVEY §2.3 "bandwidth-bound pieces"): coalesced 16-byte vector
// accesses, fp32 math, warp-shuffle / shared-memory reductions.  Algorithmic bytes per unit are stated at
// each kernel; DESIGN.md carries the roofline table.
#include <algorithm>
#include <cstdlib>
#include <new>

#include "common.cuh"
#include "internal.h"

namespace b200sd {
namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ float load_as_float(const void* p, long long i, int dtype) {
  if (dtype == B200SD_F32) return reinterpret_cast<const float*>(p)[i];
  if (dtype == B200SD_BF16) return __bfloat162float(reinterpret_cast<const bf16*>(p)[i]);
  return __half2float(reinterpret_cast<const __half*>(p)[i]);
}
__device__ __forceinline__ void store_from_float(void* p, long long i, int dtype, float v) {
  if (dtype == B200SD_F32) reinterpret_cast<float*>(p)[i] = v;
  else if (dtype == B200SD_BF16) reinterpret_cast<bf16*>(p)[i] = __float2bfloat16_rn(v);
  else reinterpret_cast<__half*>(p)[i] = __float2half_rn(v);
}

// ======================================================================================================
// GroupNorm (+SiLU), NHWC bf16, one or two channel-concatenated sources.
// Reference: ResnetBlock2D norm1/norm2 + act, Transformer2DModel.norm, conv_norm_out (diffusers).
// Pass 1 (stats): 1 read.  Pass 2 (apply): 1 read + 1 write.  Algorithmic bytes = 1R + 1W of the tensor.
// Thread t owns channel-vector (t % cols) (8 channels = 16 B) and walks pixels with stride R, so a warp
// always touches consecutive 16-byte chunks of one pixel row.
// ======================================================================================================
struct GNParams {
  const bf16* x1;
  const bf16* x2;
  long long ld1, ld2;
  int C1, C2, C;
  int NB, HW, groups, cpg, silu;
  float eps;
  const float* gamma;
  const float* beta;
  bf16* out;
  float* partials;
  int chunks, ppc, cols, R;
  double inv_cnt;  // 1 / (HW * channels per group)
};

__device__ __forceinline__ const bf16* gn_src(const GNParams& p, int n, int pix, int c0) {
  long long row = static_cast<long long>(n) * p.HW + pix;
  return (c0 < p.C1) ? p.x1 + row * p.ld1 + c0 : p.x2 + row * p.ld2 + (c0 - p.C1);
}


// Affine (+SiLU) on two values at once: packed FFMA2 / FMUL2 / FADD2 (two results per issue slot) around the four MUFU operations;
// the same operation sequence as fmaf + silu_f (x * rcp(1 + ex2(-x * log2 e))), so the results are bit-identical to the scalar form.
__device__ __forceinline__ uint32_t gn_affine_pack(float x0, float x1, float a0, float a1, float b0, float b1, bool silu) {
  uint64_t Y = f32x2_fma(f32x2_pack(x0, x1), f32x2_pack(a0, a1), f32x2_pack(b0, b1));
  if (silu) {
    float e0, e1;
    f32x2_unpack(f32x2_mul(Y, f32x2_pack(-1.4426950408889634f, -1.4426950408889634f)), e0, e1);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(e0));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(e1));
    float d0, d1;
    f32x2_unpack(f32x2_add(f32x2_pack(e0, e1), f32x2_pack(1.0f, 1.0f)), d0, d1);
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(d0) : "f"(d0));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(d1) : "f"(d1));
    Y = f32x2_mul(Y, f32x2_pack(d0, d1));
  }
  float y0, y1;
  f32x2_unpack(Y, y0, y1);
  return pack_bf16x2(y0, y1);
}
// mean and 1/sqrt(var + eps) of one group from its fp32 totals: the cancellation-prone step (E[x^2] - mean^2) in fp64 (three
// operations), the reciprocal square root as MUFU.RSQ + one Newton step in fp32 (relative error ~1e-7) instead of an fp64
// divide + square root (~150 instructions on a 1/64-rate pipe, once per CTA of every launch).
__device__ __forceinline__ void gn_finalize(float S, float SS, double inv_cnt, float eps, float& mean, float& rstd) {
  const double m = static_cast<double>(S) * inv_cnt;
  const double var = static_cast<double>(SS) * inv_cnt - m * m;
  const float v = fmaxf(static_cast<float>(var), 0.f) + eps;
  float r = rsqrtf(v);
  r = r * fmaf(-0.5f * v * r, r, 1.5f);
  mean = static_cast<float>(m);
  rstd = r;
}

// Deterministic fold of the per-thread channel sums (s, ss: 8 channels of pixel-row slot r) into per-group sums.
// No atomics: every addition happens in an order fixed by the CTA geometry (cols, R, groups), which depends on the
// per-image shape only — so the statistics of an image do not depend on the batch it is computed in, nor on warp
// scheduling.  scratch: [2][R][C] channel sums + [parts][groups][2] partial group sums (GN_FOLD_FLOATS per thread).
constexpr int GN_FOLD_FLOATS = 18;
__device__ __forceinline__ void gn_fold_groups(const GNParams& p, float* scratch, const float (&s)[8], const float (&ss)[8],
                                               int r, int c0, int tid, int nthreads, float* gsum /*[groups][2]*/) {
  const int RC = p.R * p.C;
  float* ch_s = scratch;
  float* ch_ss = scratch + RC;
  float* red = scratch + 2 * RC;
  *reinterpret_cast<float4*>(ch_s + r * p.C + c0) = make_float4(s[0], s[1], s[2], s[3]);
  *reinterpret_cast<float4*>(ch_s + r * p.C + c0 + 4) = make_float4(s[4], s[5], s[6], s[7]);
  *reinterpret_cast<float4*>(ch_ss + r * p.C + c0) = make_float4(ss[0], ss[1], ss[2], ss[3]);
  *reinterpret_cast<float4*>(ch_ss + r * p.C + c0 + 4) = make_float4(ss[4], ss[5], ss[6], ss[7]);
  __syncthreads();
  const int parts = nthreads / p.groups;  // >= 1 (the plan keeps nthreads >= groups)
  const int g = tid % p.groups, part = tid / p.groups;
  if (part < parts) {
    float S = 0.f, SS = 0.f;
    const int n = p.cpg * p.R;
    for (int e = part; e < n; e += parts) {
      const int rr = e / p.cpg;
      const int c = g * p.cpg + (e - rr * p.cpg);
      S += ch_s[rr * p.C + c];
      SS += ch_ss[rr * p.C + c];
    }
    red[(part * p.groups + g) * 2] = S;
    red[(part * p.groups + g) * 2 + 1] = SS;
  }
  __syncthreads();
  if (tid < p.groups) {
    float S = 0.f, SS = 0.f;
    for (int q = 0; q < parts; ++q) { S += red[(q * p.groups + tid) * 2]; SS += red[(q * p.groups + tid) * 2 + 1]; }
    gsum[2 * tid] = S;
    gsum[2 * tid + 1] = SS;
  }
}

__global__ void gn_stats_kernel(const GNParams p) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ __align__(16) float gn_scratch[];
  __shared__ float gsum[64 * 2];
  const int chunk = blockIdx.x, n = blockIdx.y;
  const int tid = threadIdx.x;
  const int vcol = tid % p.cols, r = tid / p.cols;
  const int c0 = vcol * 8;
  float s[8], ss[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { s[j] = 0.f; ss[j] = 0.f; }
  const int pix_end = min((chunk + 1) * p.ppc, p.HW);
  int pix = chunk * p.ppc + r;
  const bf16* src = gn_src(p, n, pix, c0);                               // pointer walk: no 64-bit multiplies in the loop
  const long long step = static_cast<long long>(p.R) * (c0 < p.C1 ? p.ld1 : p.ld2);
  for (; pix + 3 * p.R < pix_end; pix += 4 * p.R, src += 4 * step) {  // 4 independent 16-byte loads in flight per thread
    uint4 v4[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v4[u] = ldg_stream(src + u * step);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      uint32_t w[4] = {v4[u].x, v4[u].y, v4[u].z, v4[u].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 f = unpack_bf16x2(w[j]);
        s[2 * j] += f.x; ss[2 * j] += f.x * f.x;
        s[2 * j + 1] += f.y; ss[2 * j + 1] += f.y * f.y;
      }
    }
  }
  for (; pix < pix_end; pix += p.R, src += step) {
    uint4 v = ldg_stream(src);
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float2 f = unpack_bf16x2(w[j]);
      s[2 * j] += f.x; ss[2 * j] += f.x * f.x;
      s[2 * j + 1] += f.y; ss[2 * j + 1] += f.y * f.y;
    }
  }
  gn_fold_groups(p, gn_scratch, s, ss, r, c0, tid, blockDim.x, gsum);
  if (tid < p.groups) {
    float* dst = p.partials + ((static_cast<long long>(n) * p.chunks + chunk) * p.groups + tid) * 2;
    dst[0] = gsum[2 * tid];
    dst[1] = gsum[2 * tid + 1];
  }
}

__global__ void gn_apply_kernel(const GNParams p) {
  pdl_launch_dependents();
  extern __shared__ float sm[];  // mean[groups], rstd[groups], red[parts][groups][2]
  float* smean = sm;
  float* srstd = smean + p.groups;
  const int chunk = blockIdx.x, n = blockIdx.y;
  const int tid = threadIdx.x;
  const int vcol = tid % p.cols, r = tid / p.cols;
  const int c0 = vcol * 8;
  // this thread's affine parameters: weights, not produced by the previous kernel -> fetched before the dependency wait
  float a[8], b[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { a[j] = __ldg(p.gamma + c0 + j); b[j] = __ldg(p.beta + c0 + j); }
  pdl_wait();
  {
    // all threads cooperate: thread -> (group, part); parts stride over the chunk partials (fp32 sums of fp32 partials, fixed
    // order), then a shared-memory fold and the finalize of gn_finalize
    float* red = srstd + p.groups;  // [parts][groups][2]
    const int parts = blockDim.x / p.groups;
    const int g = tid % p.groups, part = tid / p.groups;
    if (part < parts) {
      float S = 0.f, SS = 0.f;
      const float* src = p.partials + (static_cast<long long>(n) * p.chunks * p.groups + g) * 2;
      for (int c = part; c < p.chunks; c += parts) {
        const float2 v = *reinterpret_cast<const float2*>(src + static_cast<long long>(c) * p.groups * 2);
        S += v.x; SS += v.y;
      }
      red[(part * p.groups + g) * 2] = S;
      red[(part * p.groups + g) * 2 + 1] = SS;
    }
    __syncthreads();
    if (tid < p.groups) {
      float S = 0.f, SS = 0.f;
      for (int q = 0; q < parts; ++q) { S += red[(q * p.groups + tid) * 2]; SS += red[(q * p.groups + tid) * 2 + 1]; }
      gn_finalize(S, SS, p.inv_cnt, p.eps, smean[tid], srstd[tid]);
    }
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int g = (c0 + j) / p.cpg;
    a[j] *= srstd[g];
    b[j] -= smean[g] * a[j];
  }
  const int pix_end = min((chunk + 1) * p.ppc, p.HW);
  const bool silu = p.silu != 0;
  auto apply_store = [&](const uint4& v, bf16* dst) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(w[j]);
      o[j] = gn_affine_pack(f.x, f.y, a[2 * j], a[2 * j + 1], b[2 * j], b[2 * j + 1], silu);
    }
    *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
  };
  int pix = chunk * p.ppc + r;
  const bf16* src = gn_src(p, n, pix, c0);  // pointer walk: no 64-bit multiplies in the loop
  const long long sstep = static_cast<long long>(p.R) * (c0 < p.C1 ? p.ld1 : p.ld2);
  bf16* dst = p.out + (static_cast<long long>(n) * p.HW + pix) * p.C + c0;
  const long long dstep = static_cast<long long>(p.R) * p.C;
  for (; pix + 3 * p.R < pix_end; pix += 4 * p.R, src += 4 * sstep, dst += 4 * dstep) {
    uint4 v4[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v4[u] = ldg_stream(src + u * sstep);
#pragma unroll
    for (int u = 0; u < 4; ++u) apply_store(v4[u], dst + u * dstep);
  }
  for (; pix < pix_end; pix += p.R, src += sstep, dst += dstep) apply_store(ldg_stream(src), dst);
}

// Single-launch GroupNorm for small images (the 8x8 level of the UNet, where the two-kernel form is bound by its
// launches and dependent global round trips, not by bytes).  A thread-block cluster owns ONE image: its CTAs split the
// pixels, each keeps its slice in shared memory between the statistics and the apply pass (one global read instead of
// two), the per-group sums are exchanged through distributed shared memory.
// Measured in-graph on B200 (tools/bench_chain.py): 16x64x1280 9.2 us vs 10.4 for the two kernels, 2x64x1280 6.8 vs 8.4;
// but a cluster is at most 8 SMs per image, so larger slices lose to the 4-CTAs-per-SM grid of the two-kernel form
// (16x1024x640, 164 KB per CTA: 37.6 vs 22.2 us; 2x256x1280: 10.8 vs 9.0) — hence the 32 KB slice limit.
constexpr int GNF_MAX_CLUSTER = 8;
constexpr int GNF_HEADER_BYTES = 1024;           // gsum[64][2], mean[64], rstd[64]
constexpr int GNF_MAX_DATA_BYTES = 32 * 1024;    // slice of one CTA

__device__ __forceinline__ float ld_shared_cluster_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}

__global__ void __launch_bounds__(512) gn_fused_kernel(const GNParams p) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ __align__(16) unsigned char gnf_smem[];
  float* gsum = reinterpret_cast<float*>(gnf_smem);  // [groups][2]
  float* smean = gsum + 128;
  float* srstd = smean + 64;
  uint4* data = reinterpret_cast<uint4*>(gnf_smem + GNF_HEADER_BYTES);
  const int chunk = blockIdx.x, n = blockIdx.y;  // cluster = all chunks of image n (gridDim.x == cluster size)
  const int tid = threadIdx.x, nthreads = blockDim.x;
  const int vcol = tid % p.cols, r = tid / p.cols;
  const int c0 = vcol * 8;
  float s[8], ss[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { s[j] = 0.f; ss[j] = 0.f; }
  const int pix_end = min((chunk + 1) * p.ppc, p.HW);
  auto accumulate = [&](const uint4& v) {
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float2 f = unpack_bf16x2(w[j]);
      s[2 * j] += f.x; ss[2 * j] += f.x * f.x;
      s[2 * j + 1] += f.y; ss[2 * j + 1] += f.y * f.y;
    }
  };
  int pix = chunk * p.ppc + r;
  int it = 0;
  for (; pix + 3 * p.R < pix_end; pix += 4 * p.R, it += 4) {  // 4 independent 16-byte loads in flight per thread
    uint4 v4[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v4[u] = ldg_stream(gn_src(p, n, pix + u * p.R, c0));
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      data[(it + u) * nthreads + tid] = v4[u];
      accumulate(v4[u]);
    }
  }
  for (; pix < pix_end; pix += p.R, ++it) {
    const uint4 v = ldg_stream(gn_src(p, n, pix, c0));
    data[it * nthreads + tid] = v;
    accumulate(v);
  }
  const int my_iters = it;
  // group sums of this CTA's slice, deterministic order (scratch sits behind the slice data)
  gn_fold_groups(p, reinterpret_cast<float*>(gnf_smem + GNF_HEADER_BYTES + GNF_MAX_DATA_BYTES), s, ss, r, c0, tid, nthreads, gsum);
  cluster_sync_all();  // every CTA's group sums are complete and visible cluster-wide
  if (tid < p.groups) {
    float S = 0.f, SS = 0.f;
    const uint32_t local = smem_u32(gsum + 2 * tid);
    for (int q = 0; q < p.chunks; ++q) {  // same order in every CTA: identical statistics across the cluster
      const uint32_t remote = mapa_u32(local, static_cast<uint32_t>(q));
      S += ld_shared_cluster_f32(remote);
      SS += ld_shared_cluster_f32(remote + 4);
    }
    gn_finalize(S, SS, p.inv_cnt, p.eps, smean[tid], srstd[tid]);
  }
  cluster_sync_all();  // peers have finished reading this CTA's sums (it may exit); mean / rstd visible to the CTA
  float a[8], b[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = c0 + j;
    const int g = c / p.cpg;
    a[j] = srstd[g] * __ldg(p.gamma + c);
    b[j] = __ldg(p.beta + c) - smean[g] * a[j];
  }
  pix = chunk * p.ppc + r;
  for (it = 0; it < my_iters; ++it, pix += p.R) {
    const uint4 v = data[it * nthreads + tid];
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(w[j]);
      o[j] = gn_affine_pack(f.x, f.y, a[2 * j], a[2 * j + 1], b[2 * j], b[2 * j + 1], p.silu != 0);
    }
    bf16* dst = p.out + (static_cast<long long>(n) * p.HW + pix) * p.C + c0;
    *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

bool gn_fused_enabled() {
  static const bool on = [] {
    const char* e = getenv("B200SD_GN_FUSED");
    return !(e && e[0] == '0');
  }();
  return on;
}

struct GNPlan { GNParams p; int threads; size_t smem; size_t smem_stats; int fused; };
int launch_groupnorm(const void* plan, cudaStream_t s) {
  const GNPlan& pl = *static_cast<const GNPlan*>(plan);
  dim3 grid(pl.p.chunks, pl.p.NB);
  if (pl.fused) {
    B200SD_CUDA(launch_kernel(gn_fused_kernel, dim3(grid), dim3(pl.threads), pl.smem, s, pl.p.chunks, pl.p));
    B200SD_CUDA(cudaGetLastError());
    return 0;
  }
  B200SD_CUDA(launch_kernel(gn_stats_kernel, dim3(grid), dim3(pl.threads), pl.smem_stats, s, 1, pl.p));
  B200SD_CUDA(cudaGetLastError());
  B200SD_CUDA(launch_kernel(gn_apply_kernel, dim3(grid), dim3(pl.threads), pl.smem, s, 1, pl.p));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

// ======================================================================================================
// LayerNorm: warp per row, the row lives in registers (C <= 2048), two-pass mean/variance like torch.
// Reference: BasicTransformerBlock.norm1/2/3 (eps 1e-5, affine).  Algorithmic bytes = 1R + 1W.
// (A persistent variant with the next row group's loads in flight won on cold inputs — 65536x320 in-graph from HBM: 21.2 ->
// 19.2 us — but lost inside the sampler step, where the input was just written and sits in L2 and its 128 registers halve the
// occupancy of the small grids: 22.6 -> 23.5 us at 65536x320, 12.2 -> 14.0 at 4096x1280.  One-shot CTAs kept.)
// ======================================================================================================
struct LNParams {
  const bf16* x; bf16* out; long long ldx, ldo; int rows, C; float eps; const float* gamma; const float* beta;
};
constexpr int LN_MAXV = 8;  // up to 8 vectors (64 channels) per lane
// A row is owned by `lpr` lanes (8, 16 or 32): with C = 320 a warp normalises 4 rows at once with all lanes busy,
// and every lane has VPL independent 16-byte loads in flight.
template <int VPL>
__global__ void __launch_bounds__(256) layernorm_kernel(const LNParams p, int lpr) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int rows_per_warp = 32 / lpr;
  const int sub = lane / lpr, sl = lane % lpr;
  const int warp_global = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int row = warp_global * rows_per_warp + sub;
  const bool active = row < p.rows;
  const int nvec = p.C >> 3;
  const bf16* xr = p.x + static_cast<long long>(active ? row : 0) * p.ldx;
  uint4 raw[VPL];
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int vi = sl + lpr * i;
    raw[i] = (active && vi < nvec) ? ldg_stream(xr + vi * 8) : make_uint4(0, 0, 0, 0);
  }
  float v[VPL][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    uint32_t w[4] = {raw[i].x, raw[i].y, raw[i].z, raw[i].w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float2 f = unpack_bf16x2(w[j]);
      v[i][2 * j] = f.x; v[i][2 * j + 1] = f.y;
      sum += f.x + f.y;
    }
  }
  for (int o = lpr >> 1; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float mean = sum / static_cast<float>(p.C);
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    if (sl + lpr * i < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { float d = v[i][j] - mean; sq += d * d; }
    }
  }
  for (int o = lpr >> 1; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  const float rstd = rsqrtf(sq / static_cast<float>(p.C) + p.eps);
  if (!active) return;
  bf16* orow = p.out + static_cast<long long>(row) * p.ldo;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int vi = sl + lpr * i;
    if (vi < nvec) {
      const float4* g4 = reinterpret_cast<const float4*>(p.gamma + vi * 8);
      const float4* b4 = reinterpret_cast<const float4*>(p.beta + vi * 8);
      float4 g0 = __ldg(g4), g1 = __ldg(g4 + 1), b0 = __ldg(b4), b1 = __ldg(b4 + 1);
      float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        o[j] = pack_bf16x2((v[i][2 * j] - mean) * rstd * g[2 * j] + b[2 * j],
                           (v[i][2 * j + 1] - mean) * rstd * g[2 * j + 1] + b[2 * j + 1]);
      *reinterpret_cast<uint4*>(orow + vi * 8) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}
int launch_layernorm(const void* plan, cudaStream_t s) {
  const LNParams& p = *static_cast<const LNParams*>(plan);
  const int nvec = p.C >> 3;
  int lpr = 32;
  for (int cand : {8, 16}) {
    if (nvec % cand == 0 && nvec / cand <= LN_MAXV) { lpr = cand; break; }
  }
  const int vpl = (nvec + lpr - 1) / lpr;
  const int rows_per_cta = 8 * (32 / lpr);
  const int grid = (p.rows + rows_per_cta - 1) / rows_per_cta;
  switch (vpl) {
    case 1: B200SD_CUDA(launch_kernel(layernorm_kernel<1>, dim3(grid), dim3(256), 0, s, 1, p, lpr)); break;
    case 2: B200SD_CUDA(launch_kernel(layernorm_kernel<2>, dim3(grid), dim3(256), 0, s, 1, p, lpr)); break;
    case 3: B200SD_CUDA(launch_kernel(layernorm_kernel<3>, dim3(grid), dim3(256), 0, s, 1, p, lpr)); break;
    case 4: B200SD_CUDA(launch_kernel(layernorm_kernel<4>, dim3(grid), dim3(256), 0, s, 1, p, lpr)); break;
    case 5: B200SD_CUDA(launch_kernel(layernorm_kernel<5>, dim3(grid), dim3(256), 0, s, 1, p, lpr)); break;
    case 6: B200SD_CUDA(launch_kernel(layernorm_kernel<6>, dim3(grid), dim3(256), 0, s, 1, p, lpr)); break;
    case 7: B200SD_CUDA(launch_kernel(layernorm_kernel<7>, dim3(grid), dim3(256), 0, s, 1, p, lpr)); break;
    default: B200SD_CUDA(launch_kernel(layernorm_kernel<8>, dim3(grid), dim3(256), 0, s, 1, p, lpr)); break;
  }
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

// ======================================================================================================
// Sinusoidal timestep embedding (diffusers Timesteps / get_timestep_embedding)
// ======================================================================================================
struct TEParams {
  const float* t; const float* table; const int* counter; int t_stride, rows, dim, flip; float shift, max_period; bf16* out;
};
__global__ void timestep_embed_kernel(const TEParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const int half = p.dim / 2;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= p.rows * half) return;
  const int row = idx / half, k = idx - row * half;
  float t;
  if (p.t) t = p.t[row * p.t_stride];
  else t = p.table[static_cast<long long>(*p.counter) * B200SD_STEP_ROW + B200SD_ST_TIMESTEP];
  float exponent = -logf(p.max_period) * static_cast<float>(k);
  exponent = exponent / (static_cast<float>(half) - p.shift);
  const float arg = t * expf(exponent);
  const float sn = sinf(arg), cs = cosf(arg);
  bf16* o = p.out + static_cast<long long>(row) * p.dim;
  if (p.flip) { o[k] = __float2bfloat16_rn(cs); o[half + k] = __float2bfloat16_rn(sn); }
  else        { o[k] = __float2bfloat16_rn(sn); o[half + k] = __float2bfloat16_rn(cs); }
}
int launch_timestep_embed(const void* plan, cudaStream_t s) {
  const TEParams& p = *static_cast<const TEParams*>(plan);
  const int n = p.rows * (p.dim / 2);
  B200SD_CUDA(launch_kernel(timestep_embed_kernel, dim3((n + 127) / 128), dim3(128), 0, s, 1, p));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

// ======================================================================================================
// PREP_LATENTS: CFG duplicate (pipeline :467-469) + scale_model_input (:472) + NCHW -> NHWC(Cpad) bf16
// ======================================================================================================
struct PLParams {
  const void* lat; bf16* out; const float* table; const int* counter; float fixed_scale; int in_dtype, B, C, hw, Cpad, rep;
};
__global__ void prep_latents_kernel(const PLParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= p.B * p.hw) return;
  const int b = idx / p.hw, pix = idx - b * p.hw;
  float scale = p.fixed_scale;
  if (p.counter) scale = p.table[static_cast<long long>(*p.counter) * B200SD_STEP_ROW + B200SD_ST_IN_SCALE];
  for (int r = 0; r < p.rep; ++r) {
    bf16* o = p.out + (static_cast<long long>(r * p.B + b) * p.hw + pix) * p.Cpad;
    for (int c = 0; c < p.Cpad; ++c) {
      float v = 0.f;
      if (c < p.C) v = load_as_float(p.lat, (static_cast<long long>(b) * p.C + c) * p.hw + pix, p.in_dtype) * scale;
      o[c] = __float2bfloat16_rn(v);
    }
  }
}
int launch_prep_latents(const void* plan, cudaStream_t s) {
  const PLParams& p = *static_cast<const PLParams*>(plan);
  const int n = p.B * p.hw;
  B200SD_CUDA(launch_kernel(prep_latents_kernel, dim3((n + 127) / 128), dim3(128), 0, s, 1, p));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

// ======================================================================================================
// nearest-2x upsample and stride-2 phase split (NHWC, 16-byte vectors)
// ======================================================================================================
struct USParams { const bf16* x; bf16* out; int NB, H, W, C; };
__global__ void upsample2x_kernel(const USParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const int cv = p.C >> 3;
  const long long total = static_cast<long long>(p.NB) * p.H * p.W * cv;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int v = static_cast<int>(idx % cv);
    long long pix = idx / cv;
    const int w = static_cast<int>(pix % p.W);
    pix /= p.W;
    const int h = static_cast<int>(pix % p.H);
    const int n = static_cast<int>(pix / p.H);
    const uint4 val = ldg_stream(p.x + idx * 8);
    const long long ob = ((static_cast<long long>(n) * 2 * p.H + 2 * h) * 2 * p.W + 2 * w) * p.C + v * 8;
    const long long rowp = static_cast<long long>(2 * p.W) * p.C;
    *reinterpret_cast<uint4*>(p.out + ob) = val;
    *reinterpret_cast<uint4*>(p.out + ob + p.C) = val;
    *reinterpret_cast<uint4*>(p.out + ob + rowp) = val;
    *reinterpret_cast<uint4*>(p.out + ob + rowp + p.C) = val;
  }
}
__global__ void phase_split_kernel(const USParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const int cv = p.C >> 3;
  const int H2 = p.H >> 1, W2 = p.W >> 1;
  const long long total = static_cast<long long>(p.NB) * p.H * p.W * cv;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int v = static_cast<int>(idx % cv);
    long long pix = idx / cv;
    const int w = static_cast<int>(pix % p.W);
    pix /= p.W;
    const int h = static_cast<int>(pix % p.H);
    const int n = static_cast<int>(pix / p.H);
    const int ph = (h & 1) * 2 + (w & 1);
    const long long o = (((static_cast<long long>(ph) * p.NB + n) * H2 + (h >> 1)) * W2 + (w >> 1)) * p.C + v * 8;
    *reinterpret_cast<uint4*>(p.out + o) = ldg_stream(p.x + idx * 8);
  }
}
int launch_upsample2x(const void* plan, cudaStream_t s) {
  const USParams& p = *static_cast<const USParams*>(plan);
  const long long total = static_cast<long long>(p.NB) * p.H * p.W * (p.C >> 3);
  const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 16));
  B200SD_CUDA(launch_kernel(upsample2x_kernel, dim3(blocks), dim3(256), 0, s, 1, p));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}
int launch_phase_split(const void* plan, cudaStream_t s) {
  const USParams& p = *static_cast<const USParams*>(plan);
  const long long total = static_cast<long long>(p.NB) * p.H * p.W * (p.C >> 3);
  const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 16));
  B200SD_CUDA(launch_kernel(phase_split_kernel, dim3(blocks), dim3(256), 0, s, 1, p));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

// ======================================================================================================
// Row softmax (VAE mid-block attention, single head d=512: scores are materialised)
// ======================================================================================================
struct SMParams { const bf16* x; bf16* out; long long ldx, ldo; int rows, cols; float scale; };
__global__ void __launch_bounds__(256) softmax_rows_kernel(const SMParams p) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[8];
  __shared__ float bcast;
  const int row = blockIdx.x;
  const bf16* xr = p.x + static_cast<long long>(row) * p.ldx;
  bf16* orow = p.out + static_cast<long long>(row) * p.ldo;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int nvec = p.cols >> 3;
  float mx = -INFINITY;
  for (int v = tid; v < nvec; v += 256) {
    uint4 u = *reinterpret_cast<const uint4*>(xr + v * 8);
    uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) { float2 f = unpack_bf16x2(w[j]); mx = fmaxf(mx, fmaxf(f.x, f.y)); }
  }
  mx = warp_max(mx);
  if (lane == 0) red[wid] = mx;
  __syncthreads();
  if (tid == 0) { float m = red[0]; for (int i = 1; i < 8; ++i) m = fmaxf(m, red[i]); bcast = m; }
  __syncthreads();
  mx = bcast;
  __syncthreads();
  const float sc = p.scale * 1.4426950408889634f;
  float sum = 0.f;
  for (int v = tid; v < nvec; v += 256) {
    uint4 u = *reinterpret_cast<const uint4*>(xr + v * 8);
    uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) { float2 f = unpack_bf16x2(w[j]); sum += exp2f((f.x - mx) * sc) + exp2f((f.y - mx) * sc); }
  }
  sum = warp_sum(sum);
  if (lane == 0) red[wid] = sum;
  __syncthreads();
  if (tid == 0) { float t = 0.f; for (int i = 0; i < 8; ++i) t += red[i]; bcast = 1.0f / t; }
  __syncthreads();
  const float inv = bcast;
  for (int v = tid; v < nvec; v += 256) {
    uint4 u = *reinterpret_cast<const uint4*>(xr + v * 8);
    uint32_t w[4] = {u.x, u.y, u.z, u.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float2 f = unpack_bf16x2(w[j]);
      o[j] = pack_bf16x2(exp2f((f.x - mx) * sc) * inv, exp2f((f.y - mx) * sc) * inv);
    }
    *reinterpret_cast<uint4*>(orow + v * 8) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}
int launch_softmax_rows(const void* plan, cudaStream_t s) {
  const SMParams& p = *static_cast<const SMParams*>(plan);
  B200SD_CUDA(launch_kernel(softmax_rows_kernel, dim3(p.rows), dim3(256), 0, s, 1, p));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

// ======================================================================================================
// CFG combine + scheduler update, fused (pipeline :484-489; scheduler math SURVEY App. B):
//   eps = u + s*(c - u);  x0 = (x - K1*eps)*K2;  x' = CX*x + CE*eps + C0*x0 + C1*(x0 - x0_prev)
// and optionally the NEXT step's scaled, CFG-duplicated NHWC UNet input.  Thread per (image, pixel).
// Algorithmic bytes (cfg, C=4): 2 eps rows (8 B each) + x read/write in its dtype.
// ======================================================================================================
struct CSParams {
  const bf16* eps; long long ld_eps; const void* x; void* x_out; float* x0_prev; bf16* next_in;
  const float* table; const int* counter; float gs; int cfg, x_dtype, B, C, hw, Cpad, rep, nsteps, gs_from_table;
  const float* noise; float* eps_hist; float* x_saved;
  const float* blend_mask; const float* blend_image; const float* blend_noise;
};
__global__ void cfg_step_kernel(const CSParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= p.B * p.hw) return;
  const int b = idx / p.hw, pix = idx - b * p.hw;
  const int step = *p.counter;
  const float* row = p.table + static_cast<long long>(step) * B200SD_STEP_ROW;
  const float K1 = row[B200SD_ST_K1], K2 = row[B200SD_ST_K2], CX = row[B200SD_ST_CX], CE = row[B200SD_ST_CE],
              C0 = row[B200SD_ST_C0], C1 = row[B200SD_ST_C1];
  const float gs = p.gs_from_table ? row[B200SD_ST_GUIDANCE] : p.gs;
  const float CN = p.noise ? row[B200SD_ST_CN] : 0.f;
  const float H1 = row[B200SD_ST_H1], H2 = row[B200SD_ST_H2], H3 = row[B200SD_ST_H3];
  const int flags = static_cast<int>(row[B200SD_ST_FLAGS]);
  const long long plane = static_cast<long long>(p.B) * p.C * p.hw;
  float next_scale = 1.0f;
  if (p.next_in && step + 1 < p.nsteps) next_scale = p.table[static_cast<long long>(step + 1) * B200SD_STEP_ROW + B200SD_ST_IN_SCALE];
  const bf16* eu = p.eps + (static_cast<long long>(b) * p.hw + pix) * p.ld_eps;
  const bf16* ec = p.eps + (static_cast<long long>(p.B + b) * p.hw + pix) * p.ld_eps;
  for (int c = 0; c < p.C; ++c) {
    const float u = __bfloat162float(eu[c]);
    float e = u;
    if (p.cfg) e = gs * (__bfloat162float(ec[c]) - u) + u;
    const long long xi = (static_cast<long long>(b) * p.C + c) * p.hw + pix;
    const float x_now = load_as_float(p.x, xi, p.x_dtype);
    float xn;
    if (flags & B200SD_STF_UNIPC) {  // UniPC: correct the incoming (predicted) sample, then predict the next one
      if (!p.x0_prev || !p.eps_hist || !p.x_saved) __trap();
      const float x0 = (x_now - K1 * e) * K2;
      const float m1 = p.x0_prev[xi], m2 = p.eps_hist[xi];
      float xc = x_now;
      if (flags & B200SD_STF_UNIPC_CORRECT) xc = H1 * p.x_saved[xi] + H2 * m1 + H3 * (m2 - m1) + CE * (x0 - m1);
      xn = CX * xc + C0 * x0 + C1 * (x0 - m1);
      p.x_saved[xi] = xc;
      p.eps_hist[xi] = m1;
      p.x0_prev[xi] = x0;
    } else {
      const float x = (p.x_saved && (flags & B200SD_STF_USE_SAVED_X)) ? p.x_saved[xi] : x_now;
      const float x0 = (x - K1 * e) * K2;
      xn = CX * x + CE * e + C0 * x0;
      if (p.x0_prev) {
        if (C1 != 0.f) xn += C1 * (x0 - p.x0_prev[xi]);
        p.x0_prev[xi] = x0;
      }
      if (CN != 0.f) xn += CN * p.noise[static_cast<long long>(step) * plane + xi];
      if (p.eps_hist) {
        const float h1 = p.eps_hist[xi], h2 = p.eps_hist[plane + xi], h3 = p.eps_hist[2 * plane + xi];
        xn += H1 * h1 + H2 * h2 + H3 * h3;
        if (flags & B200SD_STF_PUSH_EPS) {
          p.eps_hist[2 * plane + xi] = h2;
          p.eps_hist[plane + xi] = h1;
          p.eps_hist[xi] = e;
        }
      }
      if (p.x_saved && (flags & B200SD_STF_SAVE_X)) p.x_saved[xi] = x_now;
    }
    if (p.blend_mask) {  // inpainting: outside the mask the latents are the original image re-noised to the next level
      const float m = p.blend_mask[static_cast<long long>(b) * p.hw + pix];
      const float proper = row[B200SD_ST_BA] * p.blend_image[xi] + row[B200SD_ST_BS] * p.blend_noise[xi];
      xn = (1.0f - m) * proper + m * xn;
    }
    store_from_float(p.x_out, xi, p.x_dtype, xn);
    if (p.next_in) {
      // the UNet consumes the value the pipeline would hand it: latents in their storage dtype, scaled
      const float xs = load_as_float(p.x_out, xi, p.x_dtype) * next_scale;
      for (int r = 0; r < p.rep; ++r)
        p.next_in[(static_cast<long long>(r * p.B + b) * p.hw + pix) * p.Cpad + c] = __float2bfloat16_rn(xs);
    }
  }
  if (p.next_in) {
    for (int r = 0; r < p.rep; ++r)
      for (int c = p.C; c < p.Cpad; ++c)
        p.next_in[(static_cast<long long>(r * p.B + b) * p.hw + pix) * p.Cpad + c] = __float2bfloat16_rn(0.f);
  }
}
int launch_cfg_step(const void* plan, cudaStream_t s) {
  const CSParams& p = *static_cast<const CSParams*>(plan);
  const int n = p.B * p.hw;
  B200SD_CUDA(launch_kernel(cfg_step_kernel, dim3((n + 127) / 128), dim3(128), 0, s, 1, p));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

// ======================================================================================================
// small utilities
// ======================================================================================================
struct AXParams { const void* x; const void* y; void* out; float a, b; int xd, yd, od; long long n; const float* coef; };
__global__ void axpby_kernel(const AXParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const float a = p.coef ? p.coef[0] : p.a, b = p.coef ? p.coef[1] : p.b;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < p.n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float v = a * load_as_float(p.x, i, p.xd);
    if (p.y) v += b * load_as_float(p.y, i, p.yd);
    store_from_float(p.out, i, p.od, v);
  }
}
int launch_axpby(const void* plan, cudaStream_t s) {
  const AXParams& p = *static_cast<const AXParams*>(plan);
  const int blocks = static_cast<int>(std::min<long long>((p.n + 255) / 256, 148 * 8));
  B200SD_CUDA(launch_kernel(axpby_kernel, dim3(blocks), dim3(256), 0, s, 1, p));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

// VAE posterior sample (diffusers DiagonalGaussianDistribution.sample; pipeline :98-106, :1027): elementwise, fp32.
struct GSParams { const void* mom; const void* noise; void* out; float scale; int md, nd, od, B, C, hw; };
__global__ void gauss_sample_kernel(const GSParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const long long per = static_cast<long long>(p.C) * p.hw;
  const long long total = static_cast<long long>(p.B) * per;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long b = i / per, rest = i - b * per;
    const float mean = load_as_float(p.mom, b * 2 * per + rest, p.md);
    float logvar = load_as_float(p.mom, b * 2 * per + per + rest, p.md);
    logvar = fminf(fmaxf(logvar, -30.0f), 20.0f);
    const float std_ = expf(0.5f * logvar);
    store_from_float(p.out, i, p.od, p.scale * (mean + std_ * load_as_float(p.noise, i, p.nd)));
  }
}
int launch_gauss_sample(const void* plan, cudaStream_t s) {
  const GSParams& p = *static_cast<const GSParams*>(plan);
  const long long total = static_cast<long long>(p.B) * p.C * p.hw;
  const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 8));
  B200SD_CUDA(launch_kernel(gauss_sample_kernel, dim3(blocks), dim3(256), 0, s, 1, p));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

struct IOParams { const bf16* x; void* out; long long ldx; int B, C, HW, od; };
__global__ void image_out_kernel(const IOParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const long long total = static_cast<long long>(p.B) * p.HW;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int b = static_cast<int>(idx / p.HW);
    const int pix = static_cast<int>(idx - static_cast<long long>(b) * p.HW);
    const bf16* src = p.x + idx * p.ldx;
    for (int c = 0; c < p.C; ++c)
      store_from_float(p.out, (static_cast<long long>(b) * p.C + c) * p.HW + pix, p.od, __bfloat162float(src[c]));
  }
}
int launch_image_out(const void* plan, cudaStream_t s) {
  const IOParams& p = *static_cast<const IOParams*>(plan);
  const long long total = static_cast<long long>(p.B) * p.HW;
  const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 8));
  B200SD_CUDA(launch_kernel(image_out_kernel, dim3(blocks), dim3(256), 0, s, 1, p));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

// ======================================================================================================
// Output stage: image tensor in [-1, 1] (NCHW) -> uint8 HWC, `convert_pt_to_numpy` of the serving handler
// (runpod-worker/handler_logic.py:21-29): (img / 2 + 0.5).clamp(0, 1), permute, * 255, .to(torch.uint8).
// Each intermediate is rounded to the tensor's element type, as torch's eager ops do, so the bytes are identical.
// Algorithmic bytes per pixel: C elements read + C bytes written.
// ======================================================================================================
struct U8Params { const void* x; uint8_t* out; int B, C, HW, xd; };
template <int XD>
__device__ __forceinline__ float round_like(float v) {
  if (XD == B200SD_BF16) return __bfloat162float(__float2bfloat16_rn(v));
  if (XD == B200SD_F16) return __half2float(__float2half_rn(v));
  return v;
}
template <int XD>
__global__ void image_u8_kernel(const U8Params p) {
  pdl_launch_dependents();
  pdl_wait();
  const long long total = static_cast<long long>(p.B) * p.HW;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long b = idx / p.HW, pix = idx - b * p.HW;
    for (int c = 0; c < p.C; ++c) {  // consecutive threads read consecutive pixels of one plane: coalesced
      float v = load_as_float(p.x, (b * p.C + c) * p.HW + pix, XD);
      v = round_like<XD>(v * 0.5f);        // img / 2
      v = round_like<XD>(v + 0.5f);        // + 0.5
      v = fminf(fmaxf(v, 0.f), 1.f);       // clamp(0, 1)   (NaN -> 0 on both sides is not needed: VAE output is finite)
      v = round_like<XD>(v * 255.f);       // * 255
      p.out[idx * p.C + c] = static_cast<uint8_t>(static_cast<int>(v));  // truncation
    }
  }
}
int launch_image_u8(const void* plan, cudaStream_t s) {
  const U8Params& p = *static_cast<const U8Params*>(plan);
  const long long total = static_cast<long long>(p.B) * p.HW;
  const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 8));
  switch (p.xd) {
    case B200SD_BF16: B200SD_CUDA(launch_kernel(image_u8_kernel<B200SD_BF16>, dim3(blocks), dim3(256), 0, s, 1, p)); break;
    case B200SD_F16: B200SD_CUDA(launch_kernel(image_u8_kernel<B200SD_F16>, dim3(blocks), dim3(256), 0, s, 1, p)); break;
    default: B200SD_CUDA(launch_kernel(image_u8_kernel<B200SD_F32>, dim3(blocks), dim3(256), 0, s, 1, p)); break;
  }
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

// ======================================================================================================
// CLIP text embeddings: token row + position row (transformers CLIPTextEmbeddings), 8 channels per thread.
// ======================================================================================================
struct EmbParams { const int* ids; const bf16* tok; const bf16* pos; bf16* out; int rows, L, D, vocab; };
__global__ void embed_kernel(const EmbParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const int vecs = p.D >> 3;
  const long long total = static_cast<long long>(p.rows) * vecs;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int row = static_cast<int>(idx / vecs), v = static_cast<int>(idx - static_cast<long long>(row) * vecs);
    const int id = __ldg(p.ids + row);
    if (id < 0 || id >= p.vocab) __trap();  // a bad token id must not read outside the table
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(p.tok + static_cast<long long>(id) * p.D) + v);
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(p.pos + static_cast<long long>(row % p.L) * p.D) + v);
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 fa = unpack_bf16x2(aw[j]), fb = unpack_bf16x2(bw[j]);
      o[j] = pack_bf16x2(fa.x + fb.x, fa.y + fb.y);
    }
    *(reinterpret_cast<uint4*>(p.out + static_cast<long long>(row) * p.D) + v) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}
int launch_embed(const void* plan, cudaStream_t s) {
  const EmbParams& p = *static_cast<const EmbParams*>(plan);
  const long long total = static_cast<long long>(p.rows) * (p.D >> 3);
  const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 8));
  B200SD_CUDA(launch_kernel(embed_kernel, dim3(blocks), dim3(256), 0, s, 1, p));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

struct SAParams { int* counter; int* ack_remote; int* state; };
__global__ void step_advance_kernel(const SAParams p) {
  pdl_launch_dependents();
  pdl_wait();
  *p.counter += 1;
  if (p.ack_remote) {  // pair split: the peer's eps half of this step has been consumed by CFG_STEP (it ran before us)
    const int seq = p.state[0] + 1;
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p.ack_remote), "r"(seq) : "memory");
    p.state[0] = seq;
  }
}
int launch_step_advance(const void* plan, cudaStream_t s) {
  B200SD_CUDA(launch_kernel(step_advance_kernel, dim3(1), dim3(1), 0, s, 1, *static_cast<const SAParams*>(plan)));
  B200SD_CUDA(cudaGetLastError());
  return 0;
}

template <class P>
int finish_plan(PlannedOp* out, int kind, const P& p, int (*launch)(const void*, cudaStream_t), int launches = 1) {
  P* h = new (std::nothrow) P(p);
  if (!h) return fail(B200SD_ERR_CUDA, "out of host memory");
  out->kind = kind;
  out->plan = h;
  out->launch = launch;
  out->destroy = destroy_plan<P>;
  out->launches = launches;
  return 0;
}
bool dtype_ok(int d) { return d == B200SD_BF16 || d == B200SD_F32 || d == B200SD_F16; }
bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace

int groupnorm_chunks(int NB, int HW, int C) {
  if (NB < 1 || HW < 1 || C < 8) return 1;
  // One full wave: both kernels keep 4 CTAs of <= 256 threads resident per SM, so NB * chunks is sized to 4 * SMs
  // (a grid of 1.2 or 1.5 waves runs as long as 2).  The chunking fixes the order in which an image's partial sums are
  // added, so the batch-invariant mode sizes it for the canonical batch instead of the actual one.
  if (batch_invariant()) NB = kCanonBatch;
  int target = std::max(1, (4 * sm_count()) / NB);
  const int R = std::max(1, 256 / (C / 8));           // pixel rows a CTA walks in parallel
  int max_chunks = std::max(1, HW / (4 * R));         // keep >= 4 loads per thread so one unrolled batch covers a chunk
  return std::max(1, std::min(target, max_chunks));
}

int plan_groupnorm(const b200sd_groupnorm& g, PlannedOp* out) {
  GNPlan pl;
  GNParams& p = pl.p;
  B200SD_CHECK_ARG(g.x1 && g.out && g.gamma && g.beta && g.partials, "groupnorm: NULL pointer");
  B200SD_CHECK_ARG(g.C1 > 0 && g.C1 % 8 == 0 && g.C2 >= 0 && g.C2 % 8 == 0, "groupnorm: C1=%d C2=%d must be multiples of 8", g.C1, g.C2);
  B200SD_CHECK_ARG(g.C2 == 0 || g.x2, "groupnorm: C2>0 needs x2");
  const int C = g.C1 + g.C2;
  B200SD_CHECK_ARG(g.groups >= 1 && g.groups <= 64 && C % g.groups == 0, "groupnorm: groups=%d does not divide C=%d", g.groups, C);
  B200SD_CHECK_ARG(g.NB >= 1 && g.HW >= 1, "groupnorm: NB=%d HW=%d", g.NB, g.HW);
  B200SD_CHECK_ARG(g.ld1 % 8 == 0 && g.ld1 >= g.C1 && (g.C2 == 0 || (g.ld2 % 8 == 0 && g.ld2 >= g.C2)), "groupnorm: bad pitches");
  B200SD_CHECK_ARG(aligned16(g.x1) && aligned16(g.out) && (g.C2 == 0 || aligned16(g.x2)), "groupnorm: pointers must be 16-byte aligned");
  B200SD_CHECK_ARG(C / 8 <= 1024, "groupnorm: C=%d too large", C);
  p.x1 = static_cast<const bf16*>(g.x1); p.x2 = static_cast<const bf16*>(g.x2);
  p.ld1 = g.ld1; p.ld2 = g.ld2; p.C1 = g.C1; p.C2 = g.C2; p.C = C;
  p.NB = g.NB; p.HW = g.HW; p.groups = g.groups; p.cpg = C / g.groups; p.silu = g.silu; p.eps = g.eps;
  p.inv_cnt = 1.0 / (static_cast<double>(g.HW) * (C / g.groups));
  p.gamma = g.gamma; p.beta = g.beta; p.out = static_cast<bf16*>(g.out); p.partials = g.partials;
  p.cols = C / 8;
  pl.fused = 0;
  if (gn_fused_enabled() && p.cols <= 512) {
    // fused form: the image must fit the shared memory of a cluster of <= 8 CTAs
    const int R = std::max(1, 512 / p.cols);
    int chunks = std::min(GNF_MAX_CLUSTER, std::max(1, g.HW / R));
    const int ppc = (g.HW + chunks - 1) / chunks;
    chunks = (g.HW + ppc - 1) / ppc;
    const int iters = (ppc + R - 1) / R;
    const size_t data_bytes = static_cast<size_t>(iters) * p.cols * R * 16;
    if (data_bytes <= static_cast<size_t>(GNF_MAX_DATA_BYTES)) {
      static bool attr_set = false;
      if (!attr_set) {
        if (cudaFuncSetAttribute(gn_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 GNF_HEADER_BYTES + GNF_MAX_DATA_BYTES + 512 * GN_FOLD_FLOATS * 4) == cudaSuccess) attr_set = true;
        else (void)cudaGetLastError();  // no device (planning arithmetic only): the launch will report it
      }
      p.chunks = chunks; p.ppc = ppc; p.R = R;
      pl.threads = p.cols * R;
      pl.smem = GNF_HEADER_BYTES + GNF_MAX_DATA_BYTES + static_cast<size_t>(pl.threads) * GN_FOLD_FLOATS * 4;
      pl.smem_stats = 0;
      pl.fused = 1;
      return finish_plan(out, B200SD_OP_GROUPNORM, pl, launch_groupnorm, 1);
    }
  }
  p.chunks = groupnorm_chunks(g.NB, g.HW, C);
  p.ppc = (g.HW + p.chunks - 1) / p.chunks;
  p.chunks = (g.HW + p.ppc - 1) / p.ppc;
  p.cols = C / 8;
  p.R = std::max(1, 256 / p.cols);
  pl.threads = p.cols * p.R;
  if (pl.threads < g.groups) { p.R = (g.groups + p.cols - 1) / p.cols; pl.threads = p.cols * p.R; }
  pl.smem = (2 * static_cast<size_t>(C) + 2 * g.groups) * sizeof(float) + 8 +
            static_cast<size_t>(pl.threads / g.groups + 1) * g.groups * 2 * sizeof(double);
  pl.smem_stats = static_cast<size_t>(pl.threads) * GN_FOLD_FLOATS * 4;
  if (pl.smem_stats > 48 * 1024) {
    static bool stats_attr_set = false;
    if (!stats_attr_set) {
      if (cudaFuncSetAttribute(gn_stats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 1024 * GN_FOLD_FLOATS * 4) == cudaSuccess)
        stats_attr_set = true;
      else (void)cudaGetLastError();
    }
  }
  const long long need = static_cast<long long>(g.NB) * groupnorm_chunks(g.NB, g.HW, C) * g.groups * 2 * 4;
  B200SD_CHECK_ARG(g.partials_bytes >= need, "groupnorm: partials workspace %lld < %lld bytes", (long long)g.partials_bytes, need);
  return finish_plan(out, B200SD_OP_GROUPNORM, pl, launch_groupnorm, 2);
}

int plan_layernorm(const b200sd_layernorm& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.x && g.out && g.gamma && g.beta, "layernorm: NULL pointer");
  B200SD_CHECK_ARG(g.C >= 8 && g.C % 8 == 0 && g.C <= LN_MAXV * 256, "layernorm: C=%d unsupported (multiple of 8, <= %d)", g.C, LN_MAXV * 256);
  B200SD_CHECK_ARG(g.rows >= 1 && g.ldx % 8 == 0 && g.ldo % 8 == 0 && g.ldx >= g.C && g.ldo >= g.C, "layernorm: bad shape");
  B200SD_CHECK_ARG(aligned16(g.x) && aligned16(g.out) && aligned16(g.gamma) && aligned16(g.beta), "layernorm: pointers must be 16-byte aligned");
  LNParams p{static_cast<const bf16*>(g.x), static_cast<bf16*>(g.out), g.ldx, g.ldo, g.rows, g.C, g.eps, g.gamma, g.beta};
  return finish_plan(out, B200SD_OP_LAYERNORM, p, launch_layernorm);
}

int plan_timestep_embed(const b200sd_timestep_embed& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.out && g.rows >= 1 && g.dim >= 2 && g.dim % 2 == 0, "timestep_embed: bad arguments");
  B200SD_CHECK_ARG(g.t || (g.step_table && g.step_counter), "timestep_embed: need t or a step table");
  TEParams p{g.t, g.step_table, g.step_counter, g.t_stride, g.rows, g.dim, g.flip_sin_to_cos, g.freq_shift,
             g.max_period > 0.f ? g.max_period : 10000.f, static_cast<bf16*>(g.out)};
  return finish_plan(out, B200SD_OP_TIMESTEP_EMBED, p, launch_timestep_embed);
}

int plan_prep_latents(const b200sd_prep_latents& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.latents && g.out && dtype_ok(g.in_dtype), "prep_latents: bad pointers/dtype");
  B200SD_CHECK_ARG(g.B >= 1 && g.C >= 1 && g.hw >= 1 && g.Cpad >= g.C && g.rep >= 1, "prep_latents: bad shape");
  B200SD_CHECK_ARG(!g.step_counter || g.step_table, "prep_latents: step_counter without table");
  PLParams p{g.latents, static_cast<bf16*>(g.out), g.step_table, g.step_counter, g.fixed_scale, g.in_dtype, g.B, g.C, g.hw, g.Cpad, g.rep};
  return finish_plan(out, B200SD_OP_PREP_LATENTS, p, launch_prep_latents);
}

int plan_upsample2x(const b200sd_upsample2x& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.x && g.out && aligned16(g.x) && aligned16(g.out), "upsample2x: bad pointers");
  B200SD_CHECK_ARG(g.NB >= 1 && g.H >= 1 && g.W >= 1 && g.C >= 8 && g.C % 8 == 0, "upsample2x: bad shape");
  USParams p{static_cast<const bf16*>(g.x), static_cast<bf16*>(g.out), g.NB, g.H, g.W, g.C};
  return finish_plan(out, B200SD_OP_UPSAMPLE2X, p, launch_upsample2x);
}

int plan_phase_split(const b200sd_phase_split& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.x && g.out && aligned16(g.x) && aligned16(g.out), "phase_split: bad pointers");
  B200SD_CHECK_ARG(g.NB >= 1 && g.H >= 2 && g.W >= 2 && g.H % 2 == 0 && g.W % 2 == 0 && g.C >= 8 && g.C % 8 == 0,
                   "phase_split: H=%d W=%d must be even, C=%d a multiple of 8", g.H, g.W, g.C);
  USParams p{static_cast<const bf16*>(g.x), static_cast<bf16*>(g.out), g.NB, g.H, g.W, g.C};
  return finish_plan(out, B200SD_OP_PHASE_SPLIT, p, launch_phase_split);
}

int plan_softmax_rows(const b200sd_softmax_rows& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.x && g.out && aligned16(g.x) && aligned16(g.out), "softmax_rows: bad pointers");
  B200SD_CHECK_ARG(g.rows >= 1 && g.cols >= 8 && g.cols % 8 == 0 && g.ldx % 8 == 0 && g.ldo % 8 == 0, "softmax_rows: bad shape");
  SMParams p{static_cast<const bf16*>(g.x), static_cast<bf16*>(g.out), g.ldx, g.ldo, g.rows, g.cols, g.scale};
  return finish_plan(out, B200SD_OP_SOFTMAX_ROWS, p, launch_softmax_rows);
}

int plan_cfg_step(const b200sd_cfg_step& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.eps && g.x && g.x_out && g.step_table && g.step_counter && dtype_ok(g.x_dtype), "cfg_step: bad pointers/dtype");
  B200SD_CHECK_ARG(g.B >= 1 && g.C >= 1 && g.hw >= 1 && g.ld_eps >= g.C && g.nsteps >= 1, "cfg_step: bad shape");
  B200SD_CHECK_ARG(!g.next_in || (g.Cpad >= g.C && g.rep >= 1), "cfg_step: next_in needs Cpad/rep");
  B200SD_CHECK_ARG(!g.blend_mask || (g.blend_image && g.blend_noise), "cfg_step: the inpaint blend needs mask, image latents and noise");
  // (a row with B200SD_STF_UNIPC needs x0_prev, eps_hist and x_saved: the kernel traps on a NULL one rather than corrupt memory)
  CSParams p{static_cast<const bf16*>(g.eps), g.ld_eps, g.x, g.x_out, g.x0_prev, static_cast<bf16*>(g.next_in),
             g.step_table, g.step_counter, g.guidance_scale, g.cfg, g.x_dtype, g.B, g.C, g.hw, g.Cpad, g.rep, g.nsteps, g.gs_from_table,
             g.noise, g.eps_hist, g.x_saved, g.blend_mask, g.blend_image, g.blend_noise};
  return finish_plan(out, B200SD_OP_CFG_STEP, p, launch_cfg_step);
}

int plan_axpby(const b200sd_axpby& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.x && g.out && g.n >= 1 && dtype_ok(g.x_dtype) && dtype_ok(g.out_dtype) && (!g.y || dtype_ok(g.y_dtype)), "axpby: bad arguments");
  AXParams p{g.x, g.y, g.out, g.a, g.b, g.x_dtype, g.y_dtype, g.out_dtype, g.n, g.coef};
  return finish_plan(out, B200SD_OP_AXPBY, p, launch_axpby);
}

int plan_image_out(const b200sd_image_out& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.x && g.out && g.B >= 1 && g.C >= 1 && g.HW >= 1 && g.ldx >= g.C && dtype_ok(g.out_dtype), "image_out: bad arguments");
  IOParams p{static_cast<const bf16*>(g.x), g.out, g.ldx, g.B, g.C, g.HW, g.out_dtype};
  return finish_plan(out, B200SD_OP_IMAGE_OUT, p, launch_image_out);
}

int plan_image_u8(const b200sd_image_u8& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.x && g.out, "image_u8: NULL pointer");
  B200SD_CHECK_ARG(g.B >= 1 && g.C >= 1 && g.HW >= 1 && dtype_ok(g.x_dtype), "image_u8: bad shape or dtype");
  U8Params p{g.x, g.out, g.B, g.C, g.HW, g.x_dtype};
  return finish_plan(out, B200SD_OP_IMAGE_U8, p, launch_image_u8);
}

int plan_embed(const b200sd_embed& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.ids && g.tok && g.pos && g.out, "embed: NULL pointer");
  B200SD_CHECK_ARG(g.rows >= 1 && g.L >= 1 && g.vocab >= 1 && g.D >= 8 && g.D % 8 == 0, "embed: bad shape (D must be a multiple of 8)");
  B200SD_CHECK_ARG(aligned16(g.tok) && aligned16(g.pos) && aligned16(g.out), "embed: tables and output must be 16-byte aligned");
  EmbParams p{g.ids, static_cast<const bf16*>(g.tok), static_cast<const bf16*>(g.pos), static_cast<bf16*>(g.out), g.rows, g.L, g.D, g.vocab};
  return finish_plan(out, B200SD_OP_EMBED, p, launch_embed);
}

int plan_gauss_sample(const b200sd_gauss_sample& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.moments && g.noise && g.out, "gauss_sample: NULL pointer");
  B200SD_CHECK_ARG(g.B >= 1 && g.C >= 1 && g.hw >= 1 && dtype_ok(g.m_dtype) && dtype_ok(g.n_dtype) && dtype_ok(g.out_dtype),
                   "gauss_sample: bad shape or dtype");
  GSParams p{g.moments, g.noise, g.out, g.scale == 0.0f ? 1.0f : g.scale, g.m_dtype, g.n_dtype, g.out_dtype, g.B, g.C, g.hw};
  return finish_plan(out, B200SD_OP_GAUSS_SAMPLE, p, launch_gauss_sample);
}

int plan_step_advance(const b200sd_step_advance& g, PlannedOp* out) {
  B200SD_CHECK_ARG(g.step_counter, "step_advance: NULL counter");
  B200SD_CHECK_ARG((g.ack_remote == nullptr) == (g.state == nullptr), "step_advance: ack_remote and state go together");
  SAParams p{g.step_counter, g.ack_remote, g.state};
  return finish_plan(out, B200SD_OP_STEP_ADVANCE, p, launch_step_advance);
}

}  // namespace b200sd

// ======================================================================================================
// b200sd_groupnorm_static_scratch — GroupNorm (+SiLU) for callers that cannot allocate device memory (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same two kernels as the program op (gn_stats_kernel -> gn_apply_kernel), but the
// per-(image, chunk, group) partial sums live in the file-scope device array below instead of a caller workspace, and the
// call plans on the stack: no cudaMalloc / cudaFree / allocating library routine, no host heap either.
//   * every GPU that loads the library holds B200SD_GN_STATIC_SCRATCH_BYTES (1 MiB) until the process exits;
//   * shapes whose partials exceed that are rejected with B200SD_ERR_INVALID (the largest UNet / VAE shape of the path,
//     64 images x 37 chunks x 32 groups, needs 0.6 MiB);
//   * the array is shared by every call and keeps the last call's contents.  Nothing has to start at zero: the stats kernel
//     overwrites every slot the apply kernel reads.  Calls that OVERLAP on the GPU (two streams, two threads) race on it —
//     issue them on one stream, or order them with events.
// In host code `g_gn_static_scratch` is only a stand-in for the symbol, not a device address: the address is fetched once
// per device, with that device current, through cudaGetSymbolAddress and handed to the kernels.
// ======================================================================================================
static __device__ __align__(256) float g_gn_static_scratch[B200SD_GN_STATIC_SCRATCH_BYTES / sizeof(float)];

namespace b200sd {
namespace {
constexpr int kMaxDevices = 64;
float* g_gn_scratch_ptr[kMaxDevices];  // device address of g_gn_static_scratch, per device (nullptr = not set up yet)

// once per device (the caller holds the device current): the symbol's address and the >48 KB shared-memory opt-in
int gn_static_setup(int device, float** ptr) {
  if (g_gn_scratch_ptr[device] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_gn_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      (void)cudaGetLastError();
      return fail(B200SD_ERR_CUDA, "groupnorm_static_scratch: cudaGetSymbolAddress failed on device %d: %s", device,
                  cudaGetErrorString(e));
    }
    B200SD_CUDA(cudaFuncSetAttribute(gn_stats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 1024 * GN_FOLD_FLOATS * 4));
    g_gn_scratch_ptr[device] = static_cast<float*>(p);
  }
  *ptr = g_gn_scratch_ptr[device];
  return 0;
}
}  // namespace
}  // namespace b200sd

extern "C" int b200sd_groupnorm_static_scratch(int device, const b200sd_groupnorm* gp, void* cuda_stream) {
  using namespace b200sd;
  if (!gp) return fail(B200SD_ERR_INVALID, "groupnorm_static_scratch: NULL op");
  const b200sd_groupnorm& g = *gp;
  // ---- validation (host arithmetic only; the same rules as the program op) ----
  B200SD_CHECK_ARG(g.x1 && g.out && g.gamma && g.beta, "groupnorm_static_scratch: NULL pointer");
  B200SD_CHECK_ARG(g.partials == nullptr, "groupnorm_static_scratch: `partials` must be NULL (the scratch is the library's static array)");
  B200SD_CHECK_ARG(g.C1 > 0 && g.C1 % 8 == 0 && g.C2 >= 0 && g.C2 % 8 == 0, "groupnorm: C1=%d C2=%d must be multiples of 8", g.C1, g.C2);
  B200SD_CHECK_ARG(g.C2 == 0 || g.x2, "groupnorm: C2>0 needs x2");
  const int C = g.C1 + g.C2;
  B200SD_CHECK_ARG(g.groups >= 1 && g.groups <= 64 && C % g.groups == 0, "groupnorm: groups=%d does not divide C=%d", g.groups, C);
  B200SD_CHECK_ARG(g.NB >= 1 && g.HW >= 1, "groupnorm: NB=%d HW=%d", g.NB, g.HW);
  B200SD_CHECK_ARG(g.ld1 % 8 == 0 && g.ld1 >= g.C1 && (g.C2 == 0 || (g.ld2 % 8 == 0 && g.ld2 >= g.C2)), "groupnorm: bad pitches");
  B200SD_CHECK_ARG(aligned16(g.x1) && aligned16(g.out) && (g.C2 == 0 || aligned16(g.x2)), "groupnorm: pointers must be 16-byte aligned");
  B200SD_CHECK_ARG(C / 8 <= 1024, "groupnorm: C=%d too large", C);
  B200SD_CHECK_ARG(device >= 0 && device < kMaxDevices, "groupnorm_static_scratch: device %d out of range", device);
  GNPlan pl;
  GNParams& p = pl.p;
  p.x1 = static_cast<const bf16*>(g.x1); p.x2 = static_cast<const bf16*>(g.x2);
  p.ld1 = g.ld1; p.ld2 = g.ld2; p.C1 = g.C1; p.C2 = g.C2; p.C = C;
  p.NB = g.NB; p.HW = g.HW; p.groups = g.groups; p.cpg = C / g.groups; p.silu = g.silu; p.eps = g.eps;
  p.inv_cnt = 1.0 / (static_cast<double>(g.HW) * (C / g.groups));
  p.gamma = g.gamma; p.beta = g.beta; p.out = static_cast<bf16*>(g.out);
  p.chunks = groupnorm_chunks(g.NB, g.HW, C);
  p.ppc = (g.HW + p.chunks - 1) / p.chunks;
  p.chunks = (g.HW + p.ppc - 1) / p.ppc;
  p.cols = C / 8;
  p.R = std::max(1, 256 / p.cols);
  pl.threads = p.cols * p.R;
  if (pl.threads < g.groups) { p.R = (g.groups + p.cols - 1) / p.cols; pl.threads = p.cols * p.R; }
  pl.smem = (2 * static_cast<size_t>(C) + 2 * g.groups) * sizeof(float) + 8 +
            static_cast<size_t>(pl.threads / g.groups + 1) * g.groups * 2 * sizeof(double);
  pl.smem_stats = static_cast<size_t>(pl.threads) * GN_FOLD_FLOATS * 4;
  pl.fused = 0;  // always the two-kernel form: it is the one that needs scratch
  const long long need = static_cast<long long>(g.NB) * p.chunks * g.groups * 2 * 4;
  if (need > static_cast<long long>(B200SD_GN_STATIC_SCRATCH_BYTES))
    return fail(B200SD_ERR_INVALID,
                "groupnorm_static_scratch: NB=%d HW=%d C=%d needs %lld bytes of partial sums, the static scratch holds %d "
                "(B200SD_GN_STATIC_SCRATCH_BYTES): use the program op with a caller workspace for this shape",
                g.NB, g.HW, C, need, (int)B200SD_GN_STATIC_SCRATCH_BYTES);
  // ---- device: current device, symbol address (once per device), two launches ----
  int rc = b200sd_device_check(device);
  if (rc) return rc;
  B200SD_CUDA(cudaSetDevice(device));
  float* scratch = nullptr;
  rc = gn_static_setup(device, &scratch);
  if (rc) return rc;
  p.partials = scratch;
  return launch_groupnorm(&pl, static_cast<cudaStream_t>(cuda_stream));
}
