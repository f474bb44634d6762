// tcgen05 implicit-GEMM for sm_100a:  C[M,N] = A[M,K] * W[N,K]^T  (+ fused epilogue)
//
// One kernel serves every dense contraction of the UNet / VAE (reference call sites: the conv2d / linear
// library calls behind pipelines/sd_unified_pipeline.py:475-482 and :523 — SURVEY §2.3):
//   * linear / 1x1 conv : A rows are tokens, 2D TMA boxes [128 rows x 64 K]
//   * 3x3 conv (stride 1, stride 2 after PHASE_SPLIT, any tap table): A rows are a [bn x bh x bw] patch of
//     output pixels; for every tap the producer issues ONE 5D TMA box shifted by (dh,dw) — the halo is
//     zero-filled by TMA's out-of-bounds handling, so no im2col buffer ever exists.
//   * K may be the concatenation of up to 3 segments (e.g. conv2's 9*C taps followed by the 1x1 shortcut's
//     C_in columns), each with its own activation and weight tensor map.
// Structure: persistent work units — one CTA, or a CTA PAIR (cluster of 2, tcgen05.mma.cta_group::2, 256 x BN tile:
// each CTA stages its own 128 rows of A and half of the weight tile).  12 warps: warp 0 TMA producer, warp 1 MMA
// issuer (elect.sync-gated single thread; fp32 accumulators double-buffered in TMEM so the epilogue of tile i
// overlaps the main loop of tile i+1), warps 2-9 epilogue in two groups that take alternate 32-column sub-tiles
// (tcgen05.ld -> bias / time-embedding row vector / SiLU / GEGLU -> residual from a TMA-prefetched smem slot ->
// bf16 -> swizzled smem ring -> TMA store), warps 10-11 issue the epilogue's TMA loads / stores.
// Split-K (for the M=128..512 low-resolution levels) writes fp32 partial tiles that a second small kernel
// reduces with the same epilogue; a weight-resident mode serves the short-K linears (see choose_gemm_tiling).
// The kernel is instantiated once per tile shape AND epilogue family (generic / GEGLU / bias / bias + residual / split-K partials /
// bias + row vector: epilogue_family): the epilogue is the critical path of the short-K GEMMs and is bound by its register allocation
// and dependent chains, so code of a family not taken costs every launch (profiles/r02_gemm_epilogue_families_ab.txt).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <new>

#include "common.cuh"
#include "internal.h"

namespace b200sd {

namespace {

constexpr int BM = 128;                     // rows per CTA tile (a CTA pair computes 256 x BN)
constexpr int BK = 64;
constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KiB
constexpr int GEMM_THREADS = 384;           // warp 0 TMA producer, warp 1 MMA issuer, warps 2-9 epilogue (2 per TMEM quadrant),
                                            // warps 10-11 epilogue TMA issue (residual loads + stores of warp group 0 / 1)
constexpr int EPI_THREADS = 256;

constexpr int RING_SLOTS = 8;              // epilogue staging ring: 8 x [128 rows x 32 bf16 columns] (SWIZZLE_64B)
constexpr int RING_SLOT_BYTES = BM * 64;   // 8 KiB
constexpr int RING_BYTES = RING_SLOTS * RING_SLOT_BYTES;
constexpr int RES_AHEAD = 4;               // residual sub-tiles in flight ahead of the epilogue (2 per warp group)
constexpr int GEMM_SMEM_LIMIT = 232448;
// per-CTA bytes of one pipeline stage: 128 rows of A, BN (one CTA) or BN/2 (CTA pair: each CTA holds half of B) rows of W
__host__ __device__ constexpr int gemm_stage_bytes(int bn, int ncta) { return A_STAGE_BYTES + (bn / ncta) * 128; }
__host__ __device__ constexpr int gemm_fixed_bytes(int bn) {
  return RING_BYTES + 2 * bn * 4 /*bias, one row per accumulator*/ + 1024 /*align slack*/ + 512 /*barriers*/;
}
__host__ __device__ constexpr int gemm_stages(int bn, int ncta) {
  int s = (GEMM_SMEM_LIMIT - gemm_fixed_bytes(bn)) / gemm_stage_bytes(bn, ncta);
  return s > 8 ? 8 : s;
}
__host__ __device__ constexpr int tmem_cols_for(int bn) {
  return 2 * bn <= 32 ? 32 : (2 * bn <= 64 ? 64 : (2 * bn <= 128 ? 128 : (2 * bn <= 256 ? 256 : 512)));
}
// Bytes of the operand-stage region.  The kernel always takes the whole SM (one CTA per SM anyway): in the
// weight-resident mode the region is re-carved as [all k-blocks of this CTA's weight rows][A-only stages].
__host__ __device__ constexpr int gemm_region_bytes(int bn) { return (GEMM_SMEM_LIMIT - gemm_fixed_bytes(bn)) & ~1023; }
__host__ __device__ constexpr int gemm_smem_bytes(int bn, int) { return gemm_region_bytes(bn) + gemm_fixed_bytes(bn); }
constexpr int MAX_STAGES = 8;
constexpr int WRES_MIN_STAGES = 4;  // A-only stages the weight-resident mode needs to keep the L2 -> SM stream busy
static_assert(gemm_stages(256, 1) >= 3 && gemm_stages(160, 1) >= 4 && gemm_stages(256, 2) >= 4, "pipeline depth");

struct alignas(64) GemmKParams {
  CUtensorMap tmA[B200SD_MAX_SEG];
  CUtensorMap tmB[B200SD_MAX_SEG];
  CUtensorMap tmC;  // output    (staged epilogue: 32-column boxes, SWIZZLE_64B)
  CUtensorMap tmR;  // residual  (same geometry)
  int staged;       // 1: epilogue goes registers -> swizzled smem ring -> TMA store (residual arrives by TMA load)
  int nseg;
  int seg_cb[B200SD_MAX_SEG];        // 64-wide channel blocks per tap
  int seg_taps[B200SD_MAX_SEG];
  int seg_kb_start[B200SD_MAX_SEG + 1];
  int8_t tap_phase[B200SD_MAX_SEG][12];
  int8_t tap_dh[B200SD_MAX_SEG][12];
  int8_t tap_dw[B200SD_MAX_SEG][12];
  int kb_total, kb_per_split, splits;
  int spatial;
  int M, N;
  int NB, H, W;
  int bw_log2, bh_log2, bn_log2;
  int tiles_w, tiles_h, tiles_nb;
  int num_m_tiles, num_n_tiles;
  int ncta;         // 1: one CTA per 128 x BN tile; 2: CTA pair (cluster of 2, tcgen05 cta_group::2) per 256 x BN tile
  int num_pm;       // ceil(num_m_tiles / ncta): M tiles of the work decomposition
  int n_fastest;    // work order: 1 = the N tiles of one M tile are adjacent (activation read from HBM once), 0 = M fastest
  // epilogue
  void* out;
  long long ldc;
  int OH, OW, sy, sx, oy, ox;
  const float* bias;
  const float* rowvec;
  long long ld_rowvec;
  int rows_per_group;
  const bf16* residual;
  long long ldr;
  float alpha;
  int flags;
  float* partial;
  // weight-resident mode (short-K linears): a work unit keeps ONE N tile for all its M tiles, its weight rows (all k-blocks)
  // are loaded once and stay in shared memory, the pipeline stages carry activations only
  int wres, wres_stages, units_per_nt;
};

struct TileCoord {
  int mt, nt, ks;
  int m0;          // linear
  int w0, h0, n0;  // spatial
};

__device__ __forceinline__ TileCoord decode_tile(const GemmKParams& p, int mt, int nt, int ks) {
  TileCoord t;
  t.mt = mt; t.nt = nt; t.ks = ks;
  t.m0 = t.mt * BM;
  if (p.spatial) {
    int tw = t.mt % p.tiles_w;
    int r2 = t.mt / p.tiles_w;
    int th = r2 % p.tiles_h;
    int tn = r2 / p.tiles_h;  // >= tiles_nb for the phantom tile of an odd pair: every row out of bounds
    t.w0 = tw << p.bw_log2;
    t.h0 = th << p.bh_log2;
    t.n0 = tn << p.bn_log2;
  } else {
    t.w0 = t.h0 = t.n0 = 0;
  }
  return t;
}
// work item -> tile of CTA `rank` of the pair.  Work items that run together should share what does NOT fit in L2:
// n_fastest keeps the N tiles of one M tile adjacent (large activations, small weights: the level-0 layers), the
// other order walks M for a fixed weight tile (small activations, large weights: the low-resolution levels).
__device__ __forceinline__ TileCoord decode_work(const GemmKParams& p, int work, int rank) {
  int pm, nt, ks;
  if (p.n_fastest) {
    nt = work % p.num_n_tiles;
    const int rest = work / p.num_n_tiles;
    pm = rest % p.num_pm;
    ks = rest / p.num_pm;
  } else {
    pm = work % p.num_pm;
    const int rest = work / p.num_pm;
    nt = rest % p.num_n_tiles;
    ks = rest / p.num_n_tiles;
  }
  return decode_tile(p, pm * p.ncta + rank, nt, ks);
}

// Work items of persistent unit `unit` (of `num_units`) and the it-th one.
__device__ __forceinline__ int unit_items(const GemmKParams& p, int unit, int num_units) {
  if (p.wres) {
    const int first = unit / p.num_n_tiles;
    return first < p.num_pm ? (p.num_pm - first + p.units_per_nt - 1) / p.units_per_nt : 0;
  }
  const int total_work = p.num_pm * p.num_n_tiles * p.splits;
  return unit < total_work ? (total_work - unit + num_units - 1) / num_units : 0;
}
__device__ __forceinline__ TileCoord unit_tile(const GemmKParams& p, int unit, int num_units, int it, int rank) {
  if (p.wres) return decode_tile(p, (unit / p.num_n_tiles + it * p.units_per_nt) * p.ncta + rank, unit % p.num_n_tiles, 0);
  return decode_work(p, unit + it * num_units, rank);
}

struct RowInfo {
  bool valid;
  long long out_off, res_off;
  int group;
  int m;
};

__device__ __forceinline__ RowInfo row_info(const GemmKParams& p, const TileCoord& t, int r) {
  RowInfo ri;
  if (p.spatial) {
    int iw = r & ((1 << p.bw_log2) - 1);
    int ih = (r >> p.bw_log2) & ((1 << p.bh_log2) - 1);
    int in = r >> (p.bw_log2 + p.bh_log2);
    int n = t.n0 + in, oh = t.h0 + ih, ow = t.w0 + iw;
    ri.valid = (n < p.NB) && (oh < p.H) && (ow < p.W);
    long long pix = (static_cast<long long>(n) * p.OH + (oh * p.sy + p.oy)) * p.OW + (ow * p.sx + p.ox);
    ri.out_off = pix * p.ldc;
    ri.res_off = pix * p.ldr;
    ri.group = n;
    ri.m = (n * p.H + oh) * p.W + ow;
  } else {
    int m = t.m0 + r;
    ri.valid = m < p.M;
    ri.out_off = static_cast<long long>(m) * p.ldc;
    ri.res_off = static_cast<long long>(m) * p.ldr;
    ri.group = p.rows_per_group > 0 ? m / p.rows_per_group : 0;
    ri.m = m;
  }
  return ri;
}

// Fused epilogue on 16 consecutive output columns [n, n+16) of one row.
__device__ __forceinline__ void epilogue_store16(const GemmKParams& p, const RowInfo& ri, int n, float (&v)[16]) {
  if (!ri.valid || n >= p.N) return;
  const int flags = p.flags;
  const bool full = (n + 16 <= p.N);
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] *= p.alpha;
  if (flags & B200SD_GEMM_BIAS) {
    if (full) {
      const float4* b4 = reinterpret_cast<const float4*>(p.bias + n);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float4 b = __ldg(b4 + j);
        v[4 * j + 0] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
      }
    } else {
      for (int j = 0; j < 16; ++j) if (n + j < p.N) v[j] += __ldg(p.bias + n + j);
    }
  }
  if (flags & B200SD_GEMM_BIAS_ROW) {
    float b = __ldg(p.bias + ri.m);
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] += b;
  }
  if (flags & B200SD_GEMM_ROWVEC) {
    const float* rv = p.rowvec + static_cast<long long>(ri.group) * p.ld_rowvec + n;
    if (full) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float4 b = __ldg(reinterpret_cast<const float4*>(rv) + j);
        v[4 * j + 0] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
      }
    } else {
      for (int j = 0; j < 16; ++j) if (n + j < p.N) v[j] += __ldg(rv + j);
    }
  }
  if (flags & B200SD_GEMM_SILU) {
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = silu_f(v[j]);
  }
  if (flags & B200SD_GEMM_QUICK_GELU) {
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = quick_gelu_f(v[j]);
  }
  if (flags & B200SD_GEMM_GELU) {
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = gelu_erf_f(v[j]);
  }
  const bool vec_ok = full && ((p.ldc & 7) == 0);
  if (flags & B200SD_GEMM_RESIDUAL) {
    const bf16* rp = p.residual + ri.res_off + n;
    if (full && ((p.ldr & 7) == 0)) {
      uint4 r0 = __ldg(reinterpret_cast<const uint4*>(rp));
      uint4 r1 = __ldg(reinterpret_cast<const uint4*>(rp) + 1);
      uint32_t w[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float2 f = unpack_bf16x2(w[j]);
        v[2 * j] += f.x; v[2 * j + 1] += f.y;
      }
    } else {
      for (int j = 0; j < 16; ++j) if (n + j < p.N) v[j] += __bfloat162float(rp[j]);
    }
  }
  if (flags & B200SD_GEMM_OUT_F32) {
    float* op = reinterpret_cast<float*>(p.out) + ri.out_off + n;
    if (full && ((p.ldc & 3) == 0)) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        reinterpret_cast<float4*>(op)[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    } else {
      for (int j = 0; j < 16; ++j) if (n + j < p.N) op[j] = v[j];
    }
  } else {
    bf16* op = reinterpret_cast<bf16*>(p.out) + ri.out_off + n;
    if (vec_ok) {
      uint4 o0, o1;
      o0.x = pack_bf16x2(v[0], v[1]);   o0.y = pack_bf16x2(v[2], v[3]);
      o0.z = pack_bf16x2(v[4], v[5]);   o0.w = pack_bf16x2(v[6], v[7]);
      o1.x = pack_bf16x2(v[8], v[9]);   o1.y = pack_bf16x2(v[10], v[11]);
      o1.z = pack_bf16x2(v[12], v[13]); o1.w = pack_bf16x2(v[14], v[15]);
      reinterpret_cast<uint4*>(op)[0] = o0;
      reinterpret_cast<uint4*>(op)[1] = o1;
    } else {
      for (int j = 0; j < 16; ++j) if (n + j < p.N) op[j] = __float2bfloat16_rn(v[j]);
    }
  }
}

// row bias, time-embedding row vector, SiLU on 32 consecutive columns [n, n+32) (alpha and bias already applied)
__device__ __forceinline__ void epilogue_extra32(const GemmKParams& p, const RowInfo& ri, int n, float (&v)[32]) {
  const int flags = p.flags;
  if (ri.valid && (flags & B200SD_GEMM_BIAS_ROW)) {
    const float b = __ldg(p.bias + ri.m);
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] += b;
  }
  if (ri.valid && (flags & B200SD_GEMM_ROWVEC) && n < p.N) {
    const float* rv = p.rowvec + static_cast<long long>(ri.group) * p.ld_rowvec + n;
    if (n + 32 <= p.N) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(rv) + j);
        v[4 * j + 0] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) if (n + j < p.N) v[j] += __ldg(rv + j);
    }
  }
  if (flags & B200SD_GEMM_SILU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = silu_f(v[j]);
  }
  if (flags & B200SD_GEMM_QUICK_GELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = quick_gelu_f(v[j]);
  }
  if (flags & B200SD_GEMM_GELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = gelu_erf_f(v[j]);
  }
}

// GEGLU epilogue: h and g are the matching 16-column chunks of the two halves of an N tile; `n` is the
// output column (N_out = N/2).  Reference: diffusers GEGLU = proj(x).chunk(2) -> h * gelu(g).
__device__ __forceinline__ void epilogue_geglu16(const GemmKParams& p, const RowInfo& ri, int n_h, int n_g, int n_out,
                                                 float (&h)[16], float (&g)[16]) {
  if (!ri.valid) return;
  float v[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    float hh = h[j] * p.alpha, gg = g[j] * p.alpha;
    if (p.flags & B200SD_GEMM_BIAS) {
      hh += __ldg(p.bias + n_h + j);
      gg += __ldg(p.bias + n_g + j);
    }
    v[j] = hh * gelu_erf_f(gg);
  }
  bf16* op = reinterpret_cast<bf16*>(p.out) + ri.out_off + n_out;
  uint4 o0, o1;
  o0.x = pack_bf16x2(v[0], v[1]);   o0.y = pack_bf16x2(v[2], v[3]);
  o0.z = pack_bf16x2(v[4], v[5]);   o0.w = pack_bf16x2(v[6], v[7]);
  o1.x = pack_bf16x2(v[8], v[9]);   o1.y = pack_bf16x2(v[10], v[11]);
  o1.z = pack_bf16x2(v[12], v[13]); o1.w = pack_bf16x2(v[14], v[15]);
  reinterpret_cast<uint4*>(op)[0] = o0;
  reinterpret_cast<uint4*>(op)[1] = o1;
}

// GEGLU: compile-time epilogue family.  The epilogue is the critical path of the short-K GEMMs and is bound by its register
// allocation and dependent chains, not by its instruction count: code of the family NOT taken costs every launch (measured:
// profiles/r02_ln_fold_experiment.txt), so the h * gelu(g) epilogue and the bias / residual / activation one are separate kernels.
// EPI: 0 = any flag combination but GEGLU (runtime branches), 1 = GEGLU, 2 = bias only (or nothing), 3 = bias + residual,
// 4 = split-K main kernel (fp32 partial tiles only; the reduce kernel owns the epilogue), 5 = bias + per-image row vector (the
// time-embedding add of every resnet's conv1).
template <int BN, int NCTA, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1) gemm_tc_kernel(const __grid_constant__ GemmKParams p) {
  constexpr bool GEGLU = EPI == 1;
  constexpr int B_ROWS = BN / NCTA;                   // weight rows this CTA stages (a CTA pair splits the N tile)
  constexpr int B_STAGE_BYTES = B_ROWS * 128;
  constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  constexpr int NSTAGE = gemm_stages(BN, NCTA);
  constexpr int TMEM_COLS = tmem_cols_for(BN);
  constexpr uint32_t IDESC = umma_idesc_bf16(BM * NCTA, BN, 0);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* ring = smem + gemm_region_bytes(BN);  // epilogue staging ring (1 KiB aligned)
  float* bias_s = reinterpret_cast<float*>(ring + RING_BYTES);  // [2 accumulators][BN]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bias_s + 2 * BN);
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* tfull_bar = empty_bar + MAX_STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* res_bar = tempty_bar + 2;            // [RING_SLOTS] residual sub-tile landed in the slot
  uint64_t* ready_bar = res_bar + RING_SLOTS;    // [RING_SLOTS] sub-tile written by its 128 epilogue threads
  uint64_t* free_bar = ready_bar + RING_SLOTS;   // [RING_SLOTS] the slot's TMA store has finished reading it
  uint64_t* wres_bar = free_bar + RING_SLOTS;    // resident weight rows landed (pair: the leader's collects both CTAs')
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wres_bar + 1);
  static_assert((2 * MAX_STAGES + 4 + 3 * RING_SLOTS + 1) * 8 + 4 <= 512, "barrier block");
  static_assert(NSTAGE * STAGE_BYTES <= gemm_region_bytes(BN) && NSTAGE <= MAX_STAGES, "stage region");
  // stage geometry: [A | W rows] per stage, or (weight-resident) [W rows of all k-blocks] followed by A-only stages
  const bool wres = p.wres != 0;
  const int nst = wres ? p.wres_stages : NSTAGE;
  const int stage_bytes = wres ? A_STAGE_BYTES : STAGE_BYTES;
  uint8_t* const stage_base = wres ? smem + p.kb_total * B_STAGE_BYTES : smem;

  pdl_launch_dependents();  // the next kernel may start its prologue as SMs free up
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = NCTA == 2 ? static_cast<int>(cluster_ctarank()) : 0;  // rank 0 = leader of the pair (issues the MMAs)
  const int unit = blockIdx.x / NCTA;                                     // persistent work unit = CTA or CTA pair
  const int num_units = gridDim.x / NCTA;

  if (threadIdx.x == 0) {
    for (int i = 0; i < MAX_STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(wres_bar, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], 8 * NCTA);  // one arrival per epilogue warp of every CTA that reads the accumulator
    }
    for (int i = 0; i < RING_SLOTS; ++i) { mbar_init(&res_bar[i], 1); mbar_init(&ready_bar[i], 128); mbar_init(&free_bar[i], 1); }
    fence_barrier_init();
  }
  if (warp == 1) {
    if (NCTA == 2) { tmem_alloc_2cta(tmem_slot, TMEM_COLS); tmem_relinquish_2cta(); }
    else { tmem_alloc(tmem_slot, TMEM_COLS); tmem_relinquish(); }
  }
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < p.nseg; ++s) {
      tma_prefetch_desc(&p.tmA[s]);
      tma_prefetch_desc(&p.tmB[s]);
    }
    if (p.staged) { tma_prefetch_desc(&p.tmC); tma_prefetch_desc(&p.tmR); }
  }
  tc_fence_before();
  if (NCTA == 2) cluster_sync_all(); else __syncthreads();  // barriers initialised cluster-wide before any remote arrive
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // everything above overlapped the previous kernel's tail; its results are visible from here on

  const int n_items = unit_items(p, unit, num_units);

  if (warp == 0) {
    // ===================== TMA producer (one lane; in a pair BOTH CTAs load: own A rows, own half of W) =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      if (wres && n_items > 0) {  // this unit's weight rows, every k-block, once (one segment, one tap: kb == channel block)
        const int n0 = (unit % p.num_n_tiles) * BN + rank * B_ROWS;
        if (NCTA == 1) {
          mbar_arrive_expect_tx(wres_bar, p.kb_total * B_STAGE_BYTES);
          for (int kb = 0; kb < p.kb_total; ++kb) tma_load_3d(smem + kb * B_STAGE_BYTES, &p.tmB[0], wres_bar, kb * BK, 0, n0);
        } else {
          if (rank == 0) mbar_arrive_expect_tx(wres_bar, 2 * p.kb_total * B_STAGE_BYTES);
          const uint32_t wb = mapa_u32(smem_u32(wres_bar), 0);
          for (int kb = 0; kb < p.kb_total; ++kb) tma_load_3d_2cta(smem + kb * B_STAGE_BYTES, &p.tmB[0], wb, kb * BK, 0, n0);
        }
      }
      for (int it = 0; it < n_items; ++it) {
        const TileCoord t = unit_tile(p, unit, num_units, it, rank);
        const int kb0 = t.ks * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.kb_total);
        int seg = 0;
        while (seg + 1 < p.nseg && kb0 >= p.seg_kb_start[seg + 1]) ++seg;
        int local = kb0 - p.seg_kb_start[seg];
        int tap = local / p.seg_cb[seg];
        int cb = local - tap * p.seg_cb[seg];
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = stage_base + stage * stage_bytes;
          uint8_t* sb = sa + A_STAGE_BYTES;
          if (NCTA == 1) {
            mbar_arrive_expect_tx(&full_bar[stage], stage_bytes);
            if (p.spatial) {
              tma_load_5d(sa, &p.tmA[seg], &full_bar[stage], cb * BK, t.w0 + p.tap_dw[seg][tap],
                          t.h0 + p.tap_dh[seg][tap], t.n0, p.tap_phase[seg][tap]);
            } else {
              tma_load_2d(sa, &p.tmA[seg], &full_bar[stage], cb * BK, t.m0);
            }
            if (!wres) tma_load_3d(sb, &p.tmB[seg], &full_bar[stage], cb * BK, tap, t.nt * BN);
          } else {
            // the leader's full barrier collects the bytes of both CTAs (tx-count may go negative until it has armed)
            if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * stage_bytes);
            const uint32_t fb = mapa_u32(smem_u32(&full_bar[stage]), 0);
            if (p.spatial) {
              tma_load_5d_2cta(sa, &p.tmA[seg], fb, cb * BK, t.w0 + p.tap_dw[seg][tap],
                               t.h0 + p.tap_dh[seg][tap], t.n0, p.tap_phase[seg][tap]);
            } else {
              tma_load_2d_2cta(sa, &p.tmA[seg], fb, cb * BK, t.m0);
            }
            if (!wres) tma_load_3d_2cta(sb, &p.tmB[seg], fb, cb * BK, tap, t.nt * BN + rank * B_ROWS);
          }
          if (++stage == nst) { stage = 0; phase ^= 1; }
          if (++cb == p.seg_cb[seg]) {
            cb = 0;
            if (++tap == p.seg_taps[seg]) { tap = 0; ++seg; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (rank == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      if (wres && n_items > 0) {
        mbar_wait(wres_bar, 0);
        tc_fence_after();
      }
      for (int it = 0; it < n_items; ++it) {
        const int ks = wres ? 0 : (unit + it * num_units) / (p.num_pm * p.num_n_tiles);
        const int kb0 = ks * p.kb_per_split;
        const int nkb = min(kb0 + p.kb_per_split, p.kb_total) - kb0;
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int i = 0; i < nkb; ++i) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t a_addr = smem_u32(stage_base + stage * stage_bytes);
            const uint32_t b_addr = wres ? smem_u32(smem + (kb0 + i) * B_STAGE_BYTES) : a_addr + A_STAGE_BYTES;
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t da = umma_smem_desc_sw128(a_addr + k * 32, 16, 1024);
              const uint64_t db = umma_smem_desc_sw128(b_addr + k * 32, 16, 1024);
              if (NCTA == 2) umma_bf16_ss_2cta(d_tmem, da, db, IDESC, (i | k) != 0 ? 1u : 0u);
              else umma_bf16_ss(d_tmem, da, db, IDESC, (i | k) != 0 ? 1u : 0u);
            }
            if (NCTA == 2) {
              umma_commit_2cta(&empty_bar[stage], 3);   // frees the stage in both CTAs
              if (i == nkb - 1) umma_commit_2cta(&tfull_bar[acc], 3);
            } else {
              umma_commit(&empty_bar[stage]);
              if (i == nkb - 1) umma_commit(&tfull_bar[acc]);
            }
          }
          __syncwarp();
          if (++stage == nst) { stage = 0; phase ^= 1; }
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (warp >= 10) {
    // ===================== epilogue TMA issue (one lane per warp group) =====================
    // Issuing a 128-row TMA box costs the issuing thread several hundred cycles, so it is kept off the epilogue warps:
    // this lane prefetches the residual sub-tiles of its group RES_AHEAD ahead into the ring slots and stores every
    // sub-tile its 128 epilogue threads have marked ready.
    if (EPI != 4 && (EPI >= 2 || p.staged) && n_items > 0 && elect_one()) {
      const int g = warp - 10;
      constexpr bool geglu = GEGLU;
#ifdef B200SD_GEMM_NORES  // diagnostic build only: residual sub-tiles are not loaded (nor added)
      const bool has_res = false;
#else
      const bool has_res = EPI == 3 || (EPI == 0 && (p.flags & B200SD_GEMM_RESIDUAL) != 0);
#endif
      const int bn_out = geglu ? BN / 2 : BN;
      const int nsub = bn_out / 32;
      const int total_q = n_items * nsub;
      auto locate = [&](int qq, int& sub) {
        const int k = qq / nsub;
        sub = qq - k * nsub;
        return unit_tile(p, unit, num_units, k, rank);
      };
      auto issue_residual = [&](int qq) {
        int sub;
        const TileCoord tt = locate(qq, sub);
        const int slot = qq % RING_SLOTS;
        mbar_arrive_expect_tx(&res_bar[slot], RING_SLOT_BYTES);
        const int col0 = tt.nt * bn_out + sub * 32;
        if (p.spatial) tma_load_4d(ring + slot * RING_SLOT_BYTES, &p.tmR, &res_bar[slot], col0, tt.w0, tt.h0, tt.n0);
        else tma_load_2d(ring + slot * RING_SLOT_BYTES, &p.tmR, &res_bar[slot], col0, tt.m0);
      };
      if (has_res) {
        for (int qq = g; qq < RES_AHEAD && qq < total_q; qq += 2) issue_residual(qq);
      }
      for (int qq = g; qq < total_q; qq += 2) {
        const int slot = qq % RING_SLOTS;
        int sub;
        const TileCoord tt = locate(qq, sub);
        mbar_wait(&ready_bar[slot], (qq / RING_SLOTS) & 1);
#ifdef B200SD_GEMM_NOSTORE  // diagnostic build only: how fast is everything but the output stores?
        if (false) {
#else
        if (tt.mt < p.num_m_tiles) {  // (the phantom tile of an odd pair has nothing to store)
#endif
          const int col0 = tt.nt * bn_out + sub * 32;
          if (p.spatial) tma_store_4d(&p.tmC, ring + slot * RING_SLOT_BYTES, col0, tt.w0, tt.h0, tt.n0);
          else tma_store_2d(&p.tmC, ring + slot * RING_SLOT_BYTES, col0, tt.m0);
        }
        bulk_commit_group();
        bulk_wait_group_read<1>();  // every store of this group but the latest has finished reading its slot
        if (qq >= 2) mbar_arrive(&free_bar[(qq - 2) % RING_SLOTS]);
        // slot of qq + RES_AHEAD = slot of qq - RES_AHEAD: its store was read-complete one iteration ago
        if (has_res && qq + RES_AHEAD < total_q) issue_residual(qq + RES_AHEAD);
      }
      bulk_wait_group<0>();
    }
  } else {
    // ===================== epilogue: 8 warps, two per TMEM lane quadrant (hardware rule: quadrant = warp_id % 4) ==========
    // The two warp groups (warps 2-5, warps 6-9) take alternate 32-column sub-tiles of every accumulator, so the
    // TMEM -> registers -> smem -> TMA-store chain of one sub-tile overlaps the next one's.
    const int q = warp & 3;
    const int g = (warp - 2) >> 2;
    const int r = q * 32 + lane;
    const int et = (warp - 2) * 32 + lane;  // 0..255
    int acc = 0;
    uint32_t acc_phase = 0;
    auto release_acc = [&](int a) {  // this warp has read everything it needs from accumulator a
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (NCTA == 2) mbar_arrive_cluster(mapa_u32(smem_u32(&tempty_bar[a]), 0));
        else mbar_arrive(&tempty_bar[a]);
      }
    };
    if (EPI != 4 && (EPI >= 2 || p.staged)) {  // (families 2 and 3 are only planned for staged outputs)
      // ---- staged epilogue: sub-tiles go through an 8-slot swizzled smem ring; the residual sub-tile is prefetched by
      // TMA RES_AHEAD sub-tiles ahead into the slot it is added in, results leave by TMA store (coalesced, OOB clipped).
      constexpr bool geglu = GEGLU;
#ifdef B200SD_GEMM_NORES
      const bool has_res = false;
#else
      const bool has_res = EPI == 3 || (EPI == 0 && (p.flags & B200SD_GEMM_RESIDUAL) != 0);
#endif
      const bool has_bias = (p.flags & B200SD_GEMM_BIAS) != 0;
      const bool extra = EPI == 0 && (p.flags & (B200SD_GEMM_BIAS_ROW | B200SD_GEMM_ROWVEC | B200SD_GEMM_SILU | B200SD_GEMM_QUICK_GELU | B200SD_GEMM_GELU)) != 0;
      constexpr bool rowvec_only = EPI == 5;
      const int bn_out = geglu ? BN / 2 : BN;
      const int nsub = bn_out / 32;
      const int sw = (r >> 1) & 3;  // SWIZZLE_64B: 16-byte chunk index ^= (row / 2) % 4
      int qbase = 0;
      for (int it = 0; it < n_items; ++it, qbase += nsub) {
        const TileCoord t = unit_tile(p, unit, num_units, it, rank);
        const RowInfo ri = row_info(p, t, r);
        float* bs = bias_s + acc * BN;
        if (has_bias) {  // bias row of this N tile -> smem while the MMAs of the tile are still running
          for (int i = et; i < BN; i += EPI_THREADS) {
            const int n = t.nt * BN + i;
            bs[i] = n < p.N ? __ldg(p.bias + n) : 0.f;
          }
        }
        named_bar_sync(3, EPI_THREADS);
        int my_last = -1;
        for (int sub = 0; sub < nsub; ++sub) if (((qbase + sub) & 1) == g) my_last = sub;
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
        if (my_last < 0) release_acc(acc);
#pragma unroll 1
        for (int sub = 0; sub < nsub; ++sub) {
          const int qq = qbase + sub;
          if ((qq & 1) != g) continue;
          const int slot = qq % RING_SLOTS;
          uint8_t* srow = ring + slot * RING_SLOT_BYTES + r * 64;
          float v[32];
#ifdef B200SD_GEMM_NOEPI  // diagnostic build only: the accumulator is neither read nor converted (main-loop rate)
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
          if (false) {
#else
          if constexpr (geglu) {
#endif
            uint32_t uh[32], ug[32];
            tmem_ld32(taddr + sub * 32, uh);
            tmem_ld32(taddr + BN / 2 + sub * 32, ug);
            tmem_ld_wait();
            const float* bh = bs + sub * 32;
            const float* bg = bh + BN / 2;
            if (p.alpha != 1.0f) {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                uh[j] = __float_as_uint(__uint_as_float(uh[j]) * p.alpha);
                ug[j] = __float_as_uint(__uint_as_float(ug[j]) * p.alpha);
              }
            }
#pragma unroll
            for (int j4 = 0; j4 < 8; ++j4) {
              float4 b_h = make_float4(0.f, 0.f, 0.f, 0.f), b_g = b_h;
              if (has_bias) {
                b_h = reinterpret_cast<const float4*>(bh)[j4];
                b_g = reinterpret_cast<const float4*>(bg)[j4];
              }
              const float bhv[4] = {b_h.x, b_h.y, b_h.z, b_h.w}, bgv[4] = {b_g.x, b_g.y, b_g.z, b_g.w};
#pragma unroll
              for (int jj = 0; jj < 4; jj += 2) {
                const int j = 4 * j4 + jj;
                geglu_f32x2(__uint_as_float(uh[j]) + bhv[jj], __uint_as_float(uh[j + 1]) + bhv[jj + 1],
                            __uint_as_float(ug[j]) + bgv[jj], __uint_as_float(ug[j + 1]) + bgv[jj + 1], v[j], v[j + 1]);
              }
            }
#ifdef B200SD_GEMM_NOEPI
          } else if (false) {
#else
          } else {
#endif
            uint32_t u[32];
            tmem_ld32(taddr + sub * 32, u);
            tmem_ld_wait();
            if (p.alpha != 1.0f) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(u[j]) * p.alpha;
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(u[j]);
            }
            if (has_bias) {
              const float4* b4 = reinterpret_cast<const float4*>(bs + sub * 32);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 b = b4[j];
                v[4 * j + 0] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
              }
            }
            if (extra) epilogue_extra32(p, ri, t.nt * BN + sub * 32, v);
            if constexpr (rowvec_only) {
              const int n = t.nt * BN + sub * 32;
              if (ri.valid && n < p.N) {
                const float* rv = p.rowvec + static_cast<long long>(ri.group) * p.ld_rowvec + n;
                if (n + 32 <= p.N) {
#pragma unroll
                  for (int j = 0; j < 8; ++j) {
                    const float4 b = __ldg(reinterpret_cast<const float4*>(rv) + j);
                    v[4 * j + 0] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
                  }
                } else {
#pragma unroll
                  for (int j = 0; j < 32; ++j) if (n + j < p.N) v[j] += __ldg(rv + j);
                }
              }
            }
          }
          if (sub == my_last) release_acc(acc);  // accumulator fully read by this warp: hand it back to the MMA warp
          if (has_res) {
            mbar_wait(&res_bar[slot], (qq / RING_SLOTS) & 1);  // (loaded only after the slot's previous store was read)
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const uint4 rr = *reinterpret_cast<const uint4*>(srow + ((c ^ sw) * 16));
              const uint32_t w[4] = {rr.x, rr.y, rr.z, rr.w};
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_bf16x2(w[j]);
                v[8 * c + 2 * j] += f.x; v[8 * c + 2 * j + 1] += f.y;
              }
            }
          } else if (qq >= RING_SLOTS) {
            mbar_wait(&free_bar[slot], ((qq / RING_SLOTS) - 1) & 1);  // the slot's previous store has been read
          }
#pragma unroll
          for (int c = 0; c < 4; ++c)
            *reinterpret_cast<uint4*>(srow + ((c ^ sw) * 16)) =
                make_uint4(pack_bf16x2(v[8 * c], v[8 * c + 1]), pack_bf16x2(v[8 * c + 2], v[8 * c + 3]),
                           pack_bf16x2(v[8 * c + 4], v[8 * c + 5]), pack_bf16x2(v[8 * c + 6], v[8 * c + 7]));
          fence_proxy_async_smem();       // generic-proxy writes -> visible to the TMA store
          mbar_arrive(&ready_bar[slot]);
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    } else {
      // ---- direct epilogue (split-K partials, fp32 / strided / narrow outputs): 16-column chunks alternate between groups
      for (int it = 0; it < n_items; ++it) {
        const TileCoord t = unit_tile(p, unit, num_units, it, rank);
        const RowInfo ri = row_info(p, t, r);
        const bool tile_exists = t.mt < p.num_m_tiles;
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
        if (!tile_exists) {
          // phantom tile of an odd pair: nothing to write
        } else if (EPI == 4 || (EPI == 0 && p.splits > 1)) {
          float* dst = p.partial +
                       (static_cast<long long>((t.ks * p.num_n_tiles + t.nt)) * p.num_m_tiles + t.mt) * (BM * BN) +
                       static_cast<long long>(r) * BN;
#pragma unroll 1
          for (int c = g; c < BN / 16; c += 2) {
            uint32_t u[16];
            tmem_ld16(taddr + c * 16, u);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 4; ++j)
              reinterpret_cast<uint4*>(dst + c * 16)[j] = make_uint4(u[4 * j], u[4 * j + 1], u[4 * j + 2], u[4 * j + 3]);
          }
        } else if (EPI == 4) {
          // (unreachable: family 4 only writes partial tiles)
        } else if (GEGLU) {
          if constexpr (GEGLU && BN % 32 == 0) {
#pragma unroll 1
            for (int c = g; c < BN / 32; c += 2) {
              uint32_t uh[16], ug[16];
              tmem_ld16(taddr + c * 16, uh);
              tmem_ld16(taddr + BN / 2 + c * 16, ug);
              tmem_ld_wait();
              float h[16], gg[16];
#pragma unroll
              for (int j = 0; j < 16; ++j) { h[j] = __uint_as_float(uh[j]); gg[j] = __uint_as_float(ug[j]); }
              epilogue_geglu16(p, ri, t.nt * BN + c * 16, t.nt * BN + BN / 2 + c * 16, t.nt * (BN / 2) + c * 16, h, gg);
            }
          }
        } else {
#pragma unroll 1
          for (int c = g; c < BN / 16; c += 2) {
            uint32_t u[16];
            tmem_ld16(taddr + c * 16, u);
            tmem_ld_wait();
            float v[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(u[j]);
            epilogue_store16(p, ri, t.nt * BN + c * 16, v);
          }
        }
        release_acc(acc);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  }

  tc_fence_before();
  if (NCTA == 2) cluster_sync_all(); else __syncthreads();  // the pair stays resident until both halves are done
  tc_fence_after();
  if (warp == 1) {
    if (NCTA == 2) tmem_dealloc_2cta(tmem_base, TMEM_COLS);
    else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// Split-K reduction + epilogue.  grid = (tiles, row groups), block = 128.  Consecutive threads take consecutive 16-column chunks of
// a tile row (a warp reads 1-2 whole rows of every partial tile: 64-byte pieces back to back, and writes its outputs the same way);
// round 1 mapped a thread to a row, which made every warp access 32 pieces a kilobyte apart.
__host__ __device__ constexpr int splitk_rows_per_block(int bn, bool geglu) { return 128 / (geglu ? bn / 32 : bn / 16); }
template <int BN>
__global__ void __launch_bounds__(128) gemm_splitk_reduce_kernel(const __grid_constant__ GemmKParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const int tile = blockIdx.x;  // nt * num_m_tiles + mt
  const bool is_geglu = (p.flags & B200SD_GEMM_GEGLU) != 0;
  const int nchunks = is_geglu ? BN / 32 : BN / 16;
  const int rpb = 128 / nchunks;
  const int c = threadIdx.x % nchunks;
  const int r = blockIdx.y * rpb + threadIdx.x / nchunks;
  if (static_cast<int>(threadIdx.x) >= rpb * nchunks || r >= BM) return;
  const TileCoord t = decode_tile(p, tile % p.num_m_tiles, tile / p.num_m_tiles, 0);
  const RowInfo ri = row_info(p, t, r);
  const long long tile_elems = static_cast<long long>(BM) * BN;
  const long long split_stride = static_cast<long long>(p.num_m_tiles) * p.num_n_tiles * tile_elems;
  const float* src = p.partial + static_cast<long long>(tile) * tile_elems + static_cast<long long>(r) * BN;
  if (p.flags & B200SD_GEMM_GEGLU) {
    float h[16], g[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) { h[j] = 0.f; g[j] = 0.f; }
    for (int s = 0; s < p.splits; ++s) {
      const float4* ph = reinterpret_cast<const float4*>(src + s * split_stride + c * 16);
      const float4* pg = reinterpret_cast<const float4*>(src + s * split_stride + BN / 2 + c * 16);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float4 a = ph[j], b = pg[j];
        h[4 * j] += a.x; h[4 * j + 1] += a.y; h[4 * j + 2] += a.z; h[4 * j + 3] += a.w;
        g[4 * j] += b.x; g[4 * j + 1] += b.y; g[4 * j + 2] += b.z; g[4 * j + 3] += b.w;
      }
    }
    epilogue_geglu16(p, ri, t.nt * BN + c * 16, t.nt * BN + BN / 2 + c * 16, t.nt * (BN / 2) + c * 16, h, g);
  } else {
    float v[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = 0.f;
    for (int s = 0; s < p.splits; ++s) {
      const float4* ps = reinterpret_cast<const float4*>(src + s * split_stride + c * 16);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float4 a = ps[j];
        v[4 * j] += a.x; v[4 * j + 1] += a.y; v[4 * j + 2] += a.z; v[4 * j + 3] += a.w;
      }
    }
    epilogue_store16(p, ri, t.nt * BN + c * 16, v);
  }
}

struct GemmPlan {
  GemmKParams kp;
  int block_n;
  int ncta;
  int grid;
  int reduce_grid_x, reduce_grid_y;
  int epi;  // epilogue family the kernel is instantiated for (epilogue_family)
};

template <int BN, int NCTA, int EPI>
int prepare_gemm_bn() {  // called at plan time (outside any stream capture)
  static bool attr_set = false;
  if (!attr_set) {
    B200SD_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, NCTA, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     gemm_smem_bytes(BN, NCTA)));
    attr_set = true;
  }
  return 0;
}
// Epilogue family of a plan (the specialised families exist for the tile widths the UNet / VAE layers use).
int epilogue_family(int bn, int flags, int staged, int splits) {
  if (splits > 1) return 4;
  if (flags & B200SD_GEMM_GEGLU) return 1;
  if (!staged || bn < 128) return 0;
  const int other = flags & ~(B200SD_GEMM_BIAS | B200SD_GEMM_RESIDUAL);
  if (other == B200SD_GEMM_ROWVEC && !(flags & B200SD_GEMM_RESIDUAL)) return 5;
  if (other) return 0;
  return (flags & B200SD_GEMM_RESIDUAL) ? 3 : 2;
}
template <int BN, int NCTA, int EPI>
int launch_gemm_epi(const GemmPlan& pl, cudaStream_t s) {
  B200SD_CUDA(launch_kernel(gemm_tc_kernel<BN, NCTA, EPI>, dim3(pl.grid), dim3(GEMM_THREADS), gemm_smem_bytes(BN, NCTA), s, NCTA, pl.kp));
  if (pl.kp.splits > 1) {
    B200SD_CUDA(launch_kernel(gemm_splitk_reduce_kernel<BN>, dim3(pl.reduce_grid_x, pl.reduce_grid_y), dim3(128), 0, s, 1, pl.kp));
  }
  return 0;
}
// One switch for both uses: LAUNCH = false sets the kernel attribute, true launches.
template <bool LAUNCH, int BN, int NCTA>
int dispatch_gemm_epi(int epi, const GemmPlan* pl, cudaStream_t s) {
#define B200SD_GEMM_EPI(E) \
  if (epi == E) { if constexpr (LAUNCH) return launch_gemm_epi<BN, NCTA, E>(*pl, s); else return prepare_gemm_bn<BN, NCTA, E>(); }
  B200SD_GEMM_EPI(0)
  if constexpr (BN % 64 == 0) { B200SD_GEMM_EPI(1) }   // (a GEGLU tile holds whole 32-column h | g halves)
  if constexpr (BN >= 128) { B200SD_GEMM_EPI(2) B200SD_GEMM_EPI(3) B200SD_GEMM_EPI(5) }
  B200SD_GEMM_EPI(4)
#undef B200SD_GEMM_EPI
  return fail(B200SD_ERR_INVALID, "gemm: epilogue family %d not built for block_n %d", epi, BN);
}
template <bool LAUNCH, int NCTA>
int dispatch_gemm_n(int bn, int epi, const GemmPlan* pl, cudaStream_t s) {
  switch (bn) {
    case 32: return dispatch_gemm_epi<LAUNCH, 32, NCTA>(epi, pl, s);
    case 64: return dispatch_gemm_epi<LAUNCH, 64, NCTA>(epi, pl, s);
    case 128: return dispatch_gemm_epi<LAUNCH, 128, NCTA>(epi, pl, s);
    case 160: return dispatch_gemm_epi<LAUNCH, 160, NCTA>(epi, pl, s);
    case 256: return dispatch_gemm_epi<LAUNCH, 256, NCTA>(epi, pl, s);
  }
  return fail(B200SD_ERR_INVALID, "gemm: unsupported block_n %d", bn);
}
int prepare_gemm(int bn, int ncta, int epi) {
  return ncta == 2 ? dispatch_gemm_n<false, 2>(bn, epi, nullptr, nullptr) : dispatch_gemm_n<false, 1>(bn, epi, nullptr, nullptr);
}
int launch_gemm(const void* plan, cudaStream_t s) {
  const GemmPlan& pl = *static_cast<const GemmPlan*>(plan);
  return pl.ncta == 2 ? dispatch_gemm_n<true, 2>(pl.block_n, pl.epi, &pl, s) : dispatch_gemm_n<true, 1>(pl.block_n, pl.epi, &pl, s);
}

int ilog2(int v) {
  int l = 0;
  while ((1 << (l + 1)) <= v) ++l;
  return l;
}

}  // namespace

// ------------------------------------------------------------------------------------------------------
// Tiling / split-K choice (pure host arithmetic, also used by the workspace-size entry point)
// ------------------------------------------------------------------------------------------------------
int choose_gemm_tiling(const b200sd_gemm& g, GemmTiling* t) {
  B200SD_CHECK_ARG(g.nseg >= 1 && g.nseg <= B200SD_MAX_SEG, "gemm: nseg=%d out of range", g.nseg);
  B200SD_CHECK_ARG(g.N >= 1, "gemm: N=%d", g.N);
  int kb_total = 0;
  for (int s = 0; s < g.nseg; ++s) {
    const b200sd_gemm_seg& sg = g.seg[s];
    B200SD_CHECK_ARG(sg.channels >= 8 && sg.channels % 8 == 0, "gemm: segment %d channels=%d must be a multiple of 8",
                     s, sg.channels);
    B200SD_CHECK_ARG(sg.taps >= 1 && sg.taps <= B200SD_MAX_TAPS, "gemm: segment %d taps=%d", s, sg.taps);
    kb_total += sg.taps * ((sg.channels + BK - 1) / BK);
  }
  const bool geglu = (g.flags & B200SD_GEMM_GEGLU) != 0;
  t->kb_total = kb_total;
  if (g.spatial) {
    B200SD_CHECK_ARG(g.NB >= 1 && g.H >= 1 && g.W >= 1, "gemm: spatial dims NB=%d H=%d W=%d", g.NB, g.H, g.W);
    // choose (bw, bh, bn) powers of two with product 128 maximising the fraction of useful rows
    double best = -1.0;
    int best_w = 0, best_h = 0, best_n = 0;
    for (int lw = 7; lw >= 0; --lw) {
      for (int lh = 7 - lw; lh >= 0; --lh) {
        int ln = 7 - lw - lh;
        long long tw = (g.W + (1 << lw) - 1) >> lw, th = (g.H + (1 << lh) - 1) >> lh, tn = (g.NB + (1 << ln) - 1) >> ln;
        double util = static_cast<double>(g.W) * g.H * g.NB / (static_cast<double>(tw * th * tn) * 128.0);
        if (util > best + 1e-9) { best = util; best_w = lw; best_h = lh; best_n = ln; }
      }
    }
    t->bw_log2 = best_w; t->bh_log2 = best_h; t->bn_log2 = best_n;
    t->tiles_w = (g.W + (1 << best_w) - 1) >> best_w;
    t->tiles_h = (g.H + (1 << best_h) - 1) >> best_h;
    t->tiles_nb = (g.NB + (1 << best_n) - 1) >> best_n;
    t->num_m_tiles = t->tiles_w * t->tiles_h * t->tiles_nb;
  } else {
    B200SD_CHECK_ARG(g.M >= 1, "gemm: M=%d", g.M);
    t->bw_log2 = t->bh_log2 = t->bn_log2 = 0;
    t->tiles_w = t->tiles_h = t->tiles_nb = 0;
    t->num_m_tiles = (g.M + BM - 1) / BM;
  }
  // ---- block_n x CTA pair x split-K, chosen together by an estimate of the kernel's cycles --------------------------
  // Measured on B200 (tools/bench_gemm_sweep.py): an SM ingests operands at ~2.6 clk per 128-byte row (~49 B/clk — the
  // same ceiling cuBLAS sits on), so one k-block costs max(MMA cycles = 2*bn, 2.6 * (128 rows of A + bn/ncta rows of
  // W)); wide tiles and CTA pairs (tcgen05 cta_group::2: each CTA stages half of W) buy flops per ingested row, narrow
  // ones buy fuller waves on the small-M levels.  Per work item: ~1500 clk of pipeline fill + exposed epilogue.
  B200SD_CHECK_ARG(g.cta_pair >= 0 && g.cta_pair <= 2, "gemm: cta_pair=%d (0 auto, 1, 2)", g.cta_pair);
  B200SD_CHECK_ARG(g.block_n == 0 || g.block_n == 32 || g.block_n == 64 || g.block_n == 128 || g.block_n == 160 ||
                   g.block_n == 256, "gemm: block_n=%d unsupported", g.block_n);
  if (geglu && g.block_n) {
    B200SD_CHECK_ARG(g.block_n % 64 == 0 && g.N % g.block_n == 0,
                     "gemm: GEGLU needs N %% block_n == 0 and block_n %% 64 == 0 (N=%d, bn=%d)", g.N, g.block_n);
  }
  // Weight-resident mode (linears with one short-K segment): a unit keeps its N tile, the tile's weight rows stay in
  // shared memory for all its M tiles and only activations stream — a k-block then costs max(2*bn, 2.6 * 128).  The
  // L2 -> SM stream is a chip-wide ceiling (~6300 B/clk over all SMs), so the weight re-reads it removes (39 % of the
  // ingested rows at 256x160 pair tiles, 50 % at 256x256) are what bounded the K = 320 / 640 linears.
  static const bool wres_enabled = [] { const char* e = getenv("B200SD_GEMM_WRES"); return !(e && e[0] == '0'); }();
  const int bn_cand[5] = {256, 160, 128, 64, 32};
  double best_cost = 1e300;
  int best_bn = 0, best_ncta = 1, best_s = 1, best_wres = 0, best_wst = 0, best_upn = 0;
  // One search over (block_n, CTA pair, split-K[, weight-resident]) for a problem of `m_tiles` 128-row tiles;
  // `force_splits` > 0 pins the split factor.
  auto search = [&](int m_tiles, int force_splits) {
    const bool wres_shape = wres_enabled && !g.spatial && g.nseg == 1 && g.seg[0].taps == 1 && force_splits <= 1;
    best_cost = 1e300; best_bn = 0; best_ncta = 1; best_s = 1; best_wres = 0; best_wst = 0; best_upn = 0;
    double wres_cost[10];
    int wres_cfg[10][4], wres_n = 0;
    for (int bi = 0; bi < 5; ++bi) {
      const int bn = bn_cand[bi];
      if (g.block_n ? bn != g.block_n : ((bn == 64 && g.N > 64) || (bn == 32 && g.N > 32) || (geglu && (bn % 64 || g.N % bn))))
        continue;  // the narrow tiles only serve narrow outputs; a GEGLU tile holds whole h|g column pairs
      const int n_tiles = (g.N + bn - 1) / bn;
      for (int ncta = 1; ncta <= 2; ++ncta) {
        if (g.cta_pair && ncta != g.cta_pair) continue;
        if (ncta == 2 && (bn % 32 != 0 || (!g.cta_pair && m_tiles < 2))) continue;
        const int num_pm = (m_tiles + ncta - 1) / ncta;
        const int units = std::max(1, sm_count() / ncta);
        const double t_kb = std::max(2.0 * bn, 2.6 * (128 + bn / ncta));
        const int max_s = force_splits > 0 ? force_splits : std::min(32, std::max(1, kb_total / 4));
        for (int sp = force_splits > 0 ? force_splits : 1; sp <= max_s; ++sp) {
          const int per = (kb_total + sp - 1) / sp;
          if (force_splits <= 0 && (kb_total + per - 1) / per != sp) continue;
          const long long waves = (static_cast<long long>(num_pm) * n_tiles * sp + units - 1) / units;
          double cost = static_cast<double>(waves) * (per * t_kb + 1500.0);
          if (sp > 1)  // fp32 partials written (row-strided 64-byte pieces) and re-read by the reduce kernel (+ its launch);
                       // calibrated on tools/sweep_small_gemm.py: the 24 B/clk/SM first assumed made 4096-row convs split
                       // 4 ways (135 us) when 256x160 pair tiles unsplit run them in 101
            cost += 4000.0 + static_cast<double>(m_tiles) * BM * n_tiles * bn * 4.0 * (sp + 1) / (148.0 * 9.6);
          if (cost < best_cost * (1.0 - 1e-9)) { best_cost = cost; best_bn = bn; best_ncta = ncta; best_s = sp; best_wres = 0; }
        }
        if (wres_shape && n_tiles <= units) {
          const int b_bytes = kb_total * (bn / ncta) * 128;
          const int wst = std::min(MAX_STAGES, (gemm_region_bytes(bn) - b_bytes) / A_STAGE_BYTES);
          const int upn = std::min(units / n_tiles, num_pm);  // work units per N tile
          if (b_bytes < gemm_region_bytes(bn) && wst >= WRES_MIN_STAGES && num_pm >= 3 * upn) {
            const long long rounds = (num_pm + upn - 1) / upn;
            const double cost = 2.6 * kb_total * (bn / ncta) + static_cast<double>(rounds) * (kb_total * std::max(2.0 * bn, 2.6 * 128) + 1500.0);
            wres_cost[wres_n] = cost; wres_cfg[wres_n][0] = bn; wres_cfg[wres_n][1] = ncta; wres_cfg[wres_n][2] = wst; wres_cfg[wres_n][3] = upn;
            ++wres_n;
          }
        }
      }
    }
    // the weight-resident form only where the estimate says it clearly pays (measured: 65536x320x320 -5 %, but 16384x640x640
    // as 256x128 resident tiles 22.2 us vs 20.9 for plain 256x160 pair tiles, at an estimated 7 % advantage)
    for (int i = 0; i < wres_n; ++i) {
      if (wres_cost[i] < 0.9 * best_cost && (!best_wres || wres_cost[i] < wres_cost[best_wres - 1])) best_wres = i + 1;
    }
    if (best_wres) {
      const int* c = wres_cfg[best_wres - 1];
      best_cost = wres_cost[best_wres - 1]; best_bn = c[0]; best_ncta = c[1]; best_s = 1; best_wst = c[2]; best_upn = c[3];
      best_wres = 1;
    }
  };
  int force_splits = g.splits > 0 ? g.splits : 0;
  if (force_splits == 0 && batch_invariant()) {
    // Batch-invariant mode: split-K changes the order of the K summation, so its factor is decided for a canonical batch
    // of kCanonBatch images of this op's per-image shape (H*W pixels, or rows_per_sample rows) and then pinned; block_n,
    // CTA pairs and the weight-resident form only change which CTA computes a row, not how, and stay free.
    long long rows_ps = g.spatial ? static_cast<long long>(g.H) * g.W : static_cast<long long>(g.rows_per_sample);
    if (rows_ps > 0) {
      search(static_cast<int>((rows_ps * kCanonBatch + BM - 1) / BM), 0);
      force_splits = best_s;
    }
  }
  search(t->num_m_tiles, force_splits);
  B200SD_CHECK_ARG(best_bn != 0, "gemm: no tile shape fits N=%d block_n=%d cta_pair=%d", g.N, g.block_n, g.cta_pair);
  const int bn = best_bn;
  t->block_n = bn;
  t->num_n_tiles = (g.N + bn - 1) / bn;
  t->ncta = best_ncta;
  t->num_pm = (t->num_m_tiles + best_ncta - 1) / best_ncta;
  int splits = best_s;
  B200SD_CHECK_ARG(splits >= 1 && splits <= kb_total, "gemm: splits=%d invalid for %d k-blocks", splits, kb_total);
  int per = (kb_total + splits - 1) / splits;
  splits = (kb_total + per - 1) / per;
  t->splits = splits;
  t->kb_per_split = per;
  t->wres = best_wres; t->wres_stages = best_wst; t->units_per_nt = best_upn;
  static const bool debug = [] { const char* e = getenv("B200SD_GEMM_DEBUG"); return e && e[0] == '1'; }();
  if (debug)
    fprintf(stderr, "gemm tiling: M-tiles %d N %d kb %d -> bn %d ncta %d splits %d wres %d (stages %d, units/nt %d) est %.0f clk\n",
            t->num_m_tiles, g.N, kb_total, bn, best_ncta, splits, best_wres, best_wst, best_upn, best_cost);
  return 0;
}

int plan_gemm(const b200sd_gemm& g, PlannedOp* out) {
  GemmTiling t;
  int rc = choose_gemm_tiling(g, &t);
  if (rc) return rc;
  GemmPlan* pl = new (std::nothrow) GemmPlan();
  if (!pl) return fail(B200SD_ERR_CUDA, "out of host memory");
  GemmKParams& kp = pl->kp;
  memset(&kp, 0, sizeof(kp));
  const int bn = t.block_n;
  B200SD_CHECK_ARG(g.out != nullptr, "gemm: out is NULL");
  B200SD_CHECK_ARG(g.ldw % 8 == 0, "gemm: ldw=%lld must be a multiple of 8", (long long)g.ldw);
  kp.nseg = g.nseg;
  int kb = 0;
  for (int s = 0; s < g.nseg; ++s) {
    const b200sd_gemm_seg& sg = g.seg[s];
    B200SD_CHECK_ARG(sg.a && sg.w, "gemm: segment %d has NULL pointers", s);
    B200SD_CHECK_ARG(sg.lda % 8 == 0 && sg.lda >= sg.channels, "gemm: segment %d lda=%lld invalid", s, (long long)sg.lda);
    B200SD_CHECK_ARG((reinterpret_cast<uintptr_t>(sg.a) & 15) == 0 && (reinterpret_cast<uintptr_t>(sg.w) & 15) == 0,
                     "gemm: segment %d pointers must be 16-byte aligned", s);
    kp.seg_cb[s] = (sg.channels + BK - 1) / BK;
    kp.seg_taps[s] = sg.taps;
    kp.seg_kb_start[s] = kb;
    kb += kp.seg_cb[s] * sg.taps;
    for (int i = 0; i < sg.taps; ++i) {
      kp.tap_phase[s][i] = sg.tap_phase[i];
      kp.tap_dh[s][i] = sg.tap_dh[i];
      kp.tap_dw[s][i] = sg.tap_dw[i];
    }
    // activation map
    if (g.spatial) {
      B200SD_CHECK_ARG(sg.Hs >= 1 && sg.Ws >= 1 && sg.phases >= 1, "gemm: segment %d spatial source dims invalid", s);
      uint64_t dims[5] = {(uint64_t)sg.channels, (uint64_t)sg.Ws, (uint64_t)sg.Hs, (uint64_t)g.NB, (uint64_t)sg.phases};
      uint64_t st[4] = {(uint64_t)sg.lda * 2, (uint64_t)sg.Ws * sg.lda * 2, (uint64_t)sg.Hs * sg.Ws * sg.lda * 2,
                        (uint64_t)g.NB * sg.Hs * sg.Ws * sg.lda * 2};
      uint32_t box[5] = {BK, 1u << t.bw_log2, 1u << t.bh_log2, 1u << t.bn_log2, 1};
      rc = encode_tensor_map_bf16(&kp.tmA[s], sg.a, 5, dims, st, box);
    } else {
      uint64_t dims[2] = {(uint64_t)sg.channels, (uint64_t)g.M};
      uint64_t st[1] = {(uint64_t)sg.lda * 2};
      uint32_t box[2] = {BK, BM};
      rc = encode_tensor_map_bf16(&kp.tmA[s], sg.a, 2, dims, st, box);
    }
    if (rc) { delete pl; return rc; }
    // weight map: [N][taps][channels], row pitch ldw
    {
      uint64_t dims[3] = {(uint64_t)sg.channels, (uint64_t)sg.taps, (uint64_t)g.N};
      uint64_t st[2] = {(uint64_t)sg.channels * 2, (uint64_t)g.ldw * 2};
      uint32_t box[3] = {BK, 1, (uint32_t)(bn / t.ncta)};  // a CTA pair stages half of the N tile per CTA
      rc = encode_tensor_map_bf16(&kp.tmB[s], sg.w, 3, dims, st, box);
      if (rc) { delete pl; return rc; }
    }
  }
  kp.seg_kb_start[g.nseg] = kb;
  kp.kb_total = t.kb_total;
  kp.kb_per_split = t.kb_per_split;
  kp.splits = t.splits;
  kp.spatial = g.spatial;
  kp.N = g.N;
  kp.NB = g.NB; kp.H = g.H; kp.W = g.W;
  kp.M = g.spatial ? g.NB * g.H * g.W : g.M;
  kp.bw_log2 = t.bw_log2; kp.bh_log2 = t.bh_log2; kp.bn_log2 = t.bn_log2;
  kp.tiles_w = std::max(t.tiles_w, 1); kp.tiles_h = std::max(t.tiles_h, 1); kp.tiles_nb = std::max(t.tiles_nb, 1);
  kp.num_m_tiles = t.num_m_tiles; kp.num_n_tiles = t.num_n_tiles;
  kp.ncta = t.ncta; kp.num_pm = t.num_pm;
  kp.wres = t.wres; kp.wres_stages = t.wres_stages; kp.units_per_nt = t.units_per_nt;
  {
    long long act_bytes = 0, w_bytes = 0;
    for (int s = 0; s < g.nseg; ++s) {
      act_bytes += static_cast<long long>(kp.M) * g.seg[s].channels * 2;
      w_bytes += static_cast<long long>(g.N) * g.seg[s].channels * g.seg[s].taps * 2;
    }
    kp.n_fastest = act_bytes > w_bytes ? 1 : 0;
  }
  kp.out = g.out; kp.ldc = g.ldc;
  if (g.spatial) {
    kp.OH = g.OH > 0 ? g.OH : g.H; kp.OW = g.OW > 0 ? g.OW : g.W;
    kp.sy = g.sy > 0 ? g.sy : 1; kp.sx = g.sx > 0 ? g.sx : 1;
    kp.oy = g.oy; kp.ox = g.ox;
  }
  kp.bias = g.bias; kp.rowvec = g.rowvec; kp.ld_rowvec = g.ld_rowvec; kp.rows_per_group = g.rows_per_group;
  kp.residual = static_cast<const bf16*>(g.residual); kp.ldr = g.ldr;
  kp.alpha = g.alpha == 0.0f ? 1.0f : g.alpha;
  kp.flags = g.flags;
  if ((g.flags & (B200SD_GEMM_BIAS | B200SD_GEMM_BIAS_ROW)) && !g.bias) { delete pl; return fail(B200SD_ERR_INVALID, "gemm: bias flag without bias pointer"); }
  if ((g.flags & B200SD_GEMM_ROWVEC) && !g.rowvec) { delete pl; return fail(B200SD_ERR_INVALID, "gemm: rowvec flag without pointer"); }
  if ((g.flags & B200SD_GEMM_RESIDUAL) && !g.residual) { delete pl; return fail(B200SD_ERR_INVALID, "gemm: residual flag without pointer"); }
  if ((g.flags & B200SD_GEMM_GEGLU) && (g.flags & (B200SD_GEMM_RESIDUAL | B200SD_GEMM_ROWVEC | B200SD_GEMM_SILU | B200SD_GEMM_QUICK_GELU | B200SD_GEMM_GELU | B200SD_GEMM_OUT_F32 | B200SD_GEMM_BIAS_ROW))) {
    delete pl; return fail(B200SD_ERR_INVALID, "gemm: GEGLU epilogue only combines with BIAS");
  }
  if ((g.flags & B200SD_GEMM_GEGLU) && (g.ldc % 8 != 0)) { delete pl; return fail(B200SD_ERR_INVALID, "gemm: GEGLU needs ldc %% 8 == 0"); }
  if (t.splits > 1) {
    long long need = static_cast<long long>(t.splits) * t.num_m_tiles * t.num_n_tiles * BM * bn * 4;
    if (!g.workspace || g.workspace_bytes < need) {
      delete pl;
      return fail(B200SD_ERR_INVALID, "gemm: split-K=%d needs %lld workspace bytes, got %lld", t.splits, need,
                  (long long)g.workspace_bytes);
    }
    kp.partial = static_cast<float*>(g.workspace);
  }
  // ---- staged (TMA-store) epilogue when the output is a plain bf16 tile grid --------------------------------
  {
    const bool geglu = (g.flags & B200SD_GEMM_GEGLU) != 0;
    const int n_out = geglu ? g.N / 2 : g.N;
    const int bn_out = geglu ? bn / 2 : bn;
    // A strided output lattice (out_scatter: the phase convs of a folded nearest-2x upsample write every second pixel of
    // every second row) is still a plain tile grid for TMA: the tensor map carries the pixel / row strides and the base
    // pointer carries the phase offset, coordinates stay those of the H x W grid this GEMM computes.
    const bool res = (g.flags & B200SD_GEMM_RESIDUAL) != 0;
    bool ok = t.splits == 1 && !(g.flags & B200SD_GEMM_OUT_F32) && bn >= 64 && bn_out % 32 == 0 && g.ldc % 8 == 0 &&
              (reinterpret_cast<uintptr_t>(g.out) & 15) == 0 && n_out >= 32;
    if (g.spatial)
      ok = ok && kp.oy >= 0 && kp.ox >= 0 && kp.oy < kp.sy && kp.ox < kp.sx && (g.H - 1) * kp.sy + kp.oy < kp.OH &&
           (g.W - 1) * kp.sx + kp.ox < kp.OW;
    if (res) ok = ok && g.ldr % 8 == 0 && (reinterpret_cast<uintptr_t>(g.residual) & 15) == 0;
    kp.staged = 0;
    if (ok) {
      auto make = [&](CUtensorMap* m, const void* base, long long ld) {
        if (g.spatial) {
          const char* b0 = static_cast<const char*>(base) + (static_cast<long long>(kp.oy) * kp.OW + kp.ox) * ld * 2;
          uint64_t dims[4] = {(uint64_t)n_out, (uint64_t)g.W, (uint64_t)g.H, (uint64_t)g.NB};
          uint64_t st[3] = {(uint64_t)kp.sx * ld * 2, (uint64_t)kp.sy * kp.OW * ld * 2, (uint64_t)kp.OH * kp.OW * ld * 2};
          uint32_t box[4] = {32, 1u << t.bw_log2, 1u << t.bh_log2, 1u << t.bn_log2};
          return encode_tensor_map_bf16(m, b0, 4, dims, st, box, 64);
        }
        uint64_t dims[2] = {(uint64_t)n_out, (uint64_t)g.M};
        uint64_t st[1] = {(uint64_t)ld * 2};
        uint32_t box[2] = {32, BM};
        return encode_tensor_map_bf16(m, base, 2, dims, st, box, 64);
      };
      rc = make(&kp.tmC, g.out, g.ldc);
      if (!rc) rc = make(&kp.tmR, res ? g.residual : g.out, res ? g.ldr : g.ldc);
      if (rc) { delete pl; return rc; }
      kp.staged = 1;
    }
  }
  pl->block_n = bn;
  pl->ncta = t.ncta;
  pl->epi = epilogue_family(bn, kp.flags, kp.staged, kp.splits);
  rc = prepare_gemm(bn, t.ncta, pl->epi);
  if (rc) { delete pl; return rc; }
  long long total_work = static_cast<long long>(t.num_pm) * t.num_n_tiles * t.splits;
  pl->grid = static_cast<int>(std::min<long long>(total_work, std::max(1, sm_count() / t.ncta))) * t.ncta;
  if (t.wres) pl->grid = t.units_per_nt * t.num_n_tiles * t.ncta;  // unit u: N tile u % num_n_tiles, M slots u / num_n_tiles + k * units_per_nt
  pl->reduce_grid_x = t.num_m_tiles * t.num_n_tiles;
  {
    const int rpb = splitk_rows_per_block(bn, (g.flags & B200SD_GEMM_GEGLU) != 0);
    pl->reduce_grid_y = (BM + rpb - 1) / rpb;
  }
  out->kind = B200SD_OP_GEMM;
  out->plan = pl;
  out->launch = launch_gemm;
  out->destroy = destroy_plan<GemmPlan>;
  out->launches = t.splits > 1 ? 2 : 1;
  (void)ilog2;
  return 0;
}

}  // namespace b200sd
