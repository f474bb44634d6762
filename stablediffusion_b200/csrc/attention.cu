// Fused flash-style attention on tcgen05 (sm_100a):  O = softmax(scale * Q K^T) V  per (batch, head).
// Replaces F.scaled_dot_product_attention behind diffusers' AttnProcessor2_0 (self- and cross-attention of
// every BasicTransformerBlock).  No mask, no dropout, scale = d^-1/2 (given).
//
// One CTA = NQ query tiles of 128 rows of one head (NQ = 4 at d <= 64, 2 above), KV blocks of 64 keys, one CTA per
// SM.  Warp roles: warp 0 TMA producer, warp 1 MMA issuer, then one softmax warpgroup per query tile (softmax +
// correction + epilogue, one query row per thread: the TMEM lane IS the row, so row max / row sum need no shuffles).
//   S_j = Q K_j^T          tcgen05.mma 128x64x16, K-major operands, into the tile's 64 TMEM columns
//   P_j = exp2(S_j - m)    registers -> bf16 pairs -> tcgen05.st over columns [0, 32) of S_j (tensor memory)
//   O  += P_j V_j          tcgen05.mma 128 x round16(d) x 16 with P as the TMEM A operand ("TS" form), V consumed
//                          MN-major straight from its [keys, d] row layout (no transpose anywhere), O accumulates in TMEM.
// (Until session 12 P went through a swizzled shared-memory tile — the -DB200SD_ATT_SS=1 build; measured on one box:
//  d = 64 N = 4096: 276 -> 241 us, d = 160: 22.5 -> 19.4 us, d = 40: 653 -> 650 us, the latter being XU-bound.)
// Measured on B200 (diagnostic builds, N = 4096, d = 40): with one 128-row tile per CTA and four CTAs per SM the
// kernel sat on TWO equal ceilings — the softmax chain (710 us with the K/V loads removed) and the L2 -> SM traffic
// of the K/V blocks (750 us with the softmax removed; 374 us with both removed): every 128 query rows re-read all of
// K and V, 4.3 GB per launch.  Sharing each K/V block between the NQ query tiles of a CTA divides that traffic by NQ
// and frees the shared memory for a 3-deep K and V ring.  The query tiles run as independent pipelines through the
// one MMA warp (it polls their barriers round-robin), so while one tile waits for its PV_j / S_{j+1} round trip
// through the tensor core the softmax warpgroups of the others own the SM's four MUFU units.  ONE pass over S per
// block (lazy running max, below) keeps the MUFU work at its minimum of 8192 ex2 per 128x64 block.
// Head dims that are not a multiple of 64 (SD1.5: 40, 80, 160) are handled by TMA itself: the box is 64 wide
// but the tensor-map extent is d, so columns >= d arrive as zeros and never touch HBM.
#include <algorithm>
#include <cstdlib>
#include <new>
#include <type_traits>

#include "common.cuh"
#include "internal.h"

namespace b200sd {
#ifdef B200SD_ATT_TRACE
// Tuning aid (not part of the ABI, compiled only with -DB200SD_ATT_TRACE): clock stamps of one CTA's roles (0 TMA, 1 MMA, 2.. softmax of tile 0..).
#ifndef B200SD_ATT_TRACE_Z
#define B200SD_ATT_TRACE_Z 0
#endif
__device__ unsigned long long g_att_trace[6 * 1024];
#define ATT_TR(role, id)                                                                                      \
  do {                                                                                                        \
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == B200SD_ATT_TRACE_Z && tr_n < 1024)                                  \
      g_att_trace[(role) * 1024 + tr_n++] = (static_cast<unsigned long long>(id) << 48) | (clock64() & 0xFFFFFFFFFFFFull); \
  } while (0)
extern "C" __attribute__((visibility("default"))) int b200sd_debug_att_trace(unsigned long long* dst) {
  return static_cast<int>(cudaMemcpyFromSymbol(dst, g_att_trace, sizeof(g_att_trace)));
}
#else
#define ATT_TR(role, id) do {} while (0)
#endif
namespace {

// P_j as the TMEM A operand of the PV product (tcgen05.mma "TS" form: P is written with tcgen05.st over the first 32
// columns of the S_j it was computed from) instead of a swizzled shared-memory tile.  -DB200SD_ATT_SS=1 builds the
// older shared-memory form (A/B measurements).
#ifndef B200SD_ATT_SS
#define B200SD_ATT_SS 0
#endif
constexpr bool ATT_TS = (B200SD_ATT_SS == 0);

constexpr int BLK = 128;           // rows of a query tile
constexpr int KB = 64;             // keys per KV block
constexpr int TILE16K = BLK * 128; // [128 rows x 64 bf16] swizzled tile (Q chunk, P)
constexpr int TILE8K = KB * 128;   // [64 rows x 64 bf16] swizzled tile (K / V chunk)

struct alignas(64) AttnParams {
  CUtensorMap tmQ;       // 4D (d, heads, rows, B), box (64, 1, 128, 1)
  CUtensorMap tmK, tmV;  // 4D (d, heads, rows, B), box (64, 1, 64, 1)
  bf16* out;
  long long ldo;
  int B, heads, Nq, Nkv, d;
  int kv_blocks;
  int n_pv;        // round16(d): UMMA N of the PV product
  int ksteps_qk;   // ceil(d/16)
  float scale_log2;
  int causal;      // key j visible to query i only when j <= i
};

// NQ = query tiles (softmax warpgroups) per CTA.  The wide configurations (4 tiles at d <= 64, 2 above) own an SM;
// the single-tile ones keep the small footprint (4 / 2 / 1 CTAs per SM) for grids the wide CTAs would quantize badly.
template <int DPAD, int NQ>
struct AttnCfg {
  static constexpr int DB = DPAD / 64;
  static constexpr int NMMA = NQ >= 4 ? 2 : 1;              // MMA warps; each serves NQ / NMMA tiles
  static constexpr int THREADS = 64 + 128 * NQ + (NMMA - 1) * 32;  // warp 0 TMA, warp 1 MMA, NQ softmax warpgroups, [last warp MMA]
  static_assert(NQ % NMMA == 0, "tiles per MMA warp");
  static constexpr int OCC = NQ > 1 ? 1 : (DPAD == 64 ? 4 : (DPAD == 128 ? 2 : 1));  // CTAs per SM
  static constexpr int TILE_COLS = KB + DPAD;               // TMEM columns of a tile: 64 of S, then O
  static constexpr int TMEM_COLS = NQ * TILE_COLS <= 128 ? 128 : (NQ * TILE_COLS <= 256 ? 256 : 512);
  // ring depths: K_j is free once S_j has completed (a whole softmax before K_{j+1} is needed), V_j once PV_j has
  static constexpr int NSTK = NQ > 1 ? (DPAD == 192 ? 2 : 3) : 1;
  static constexpr int NSTV = NQ > 1 ? (DPAD == 192 ? 2 : 3) : (DPAD == 64 ? 1 : 2);
  static constexpr int QT_BYTES = DB * TILE16K;             // one query tile
  static constexpr int KV_BYTES = DB * TILE8K;              // one K (or V) stage
  static constexpr int P_BYTES = ATT_TS ? 0 : TILE16K;      // shared-memory P tile (SS form only)
  static constexpr int SMEM = NQ * (QT_BYTES + P_BYTES) + (NSTK + NSTV) * KV_BYTES + 1024 + 256;
  static_assert(OCC * (SMEM + 1024) <= 233472, "shared memory budget");
  static_assert(OCC * TMEM_COLS <= 512 && NQ * TILE_COLS <= TMEM_COLS, "TMEM budget");
};

__device__ __forceinline__ float ex2_approx(float x) {
#ifdef B200SD_ATT_NOEXP  // diagnostic build only: how fast is everything but the MUFU work?
  return x * 0.001f + 0.01f;
#else
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
#endif
}
// Packed fp32 pairs (FFMA2 / FADD2: two results per issue slot).  The softmax warps share their schedulers' issue slots
// with the MUFU stream, so every instruction saved raises the MUFU density.
__device__ __forceinline__ uint64_t pack_f32x2(uint32_t lo, uint32_t hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ void unpack_f32x2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma_f32x2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t add_f32x2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
// 2^x for a PAIR of scores on the FMA pipe (no MUFU): round-to-nearest split x = n + f with the 1.5 * 2^23 trick, cubic
// minimax for 2^f on [-0.5, 0.5] (relative error 7.5e-5, a 50th of a bf16 ulp of P), n added into the exponent field.
// The softmax of the d = 40 heads is bound by the XU pipe (16 ex2 per clock and SM); with the packed-pair arithmetic the
// warps' issue slots are only ~40 % used, so every B200SD_ATT_POLY_EVERY-th pair takes this path instead (0 = none).
// Measured (tools/ab_attention_poly.sh, N = 4096, d = 40, same box): none 637.8 us, every 8th 613.4, every 4th 640.0, every
// 3rd 654.3, every 2nd 697.3 — a pair costs ~10 issue slots on this path, so beyond one in eight the issue slots become the
// bound the MUFU pipe was.
#ifndef B200SD_ATT_POLY_EVERY
#define B200SD_ATT_POLY_EVERY 8
#endif
__device__ __forceinline__ void ex2_poly_f32x2(float x0, float x1, float& p0, float& p1) {
  const uint64_t X = pack_f32x2(__float_as_uint(fmaxf(x0, -125.0f)), __float_as_uint(fmaxf(x1, -125.0f)));
  const uint64_t MAGIC = pack_f32x2(0x4B400000u, 0x4B400000u);    // 12582912.0f
  const uint64_t NMAGIC = pack_f32x2(0xCB400000u, 0xCB400000u);
  const uint64_t XR = add_f32x2(X, MAGIC);                          // low mantissa bits = round(x)
  const uint64_t N = add_f32x2(XR, NMAGIC);
  const uint64_t F = fma_f32x2(N, pack_f32x2(0xBF800000u, 0xBF800000u), X);  // x - n
  uint64_t P = fma_f32x2(pack_f32x2(__float_as_uint(0.05517115f), __float_as_uint(0.05517115f)), F,
                         pack_f32x2(__float_as_uint(0.24260987f), __float_as_uint(0.24260987f)));
  P = fma_f32x2(P, F, pack_f32x2(__float_as_uint(0.69326097f), __float_as_uint(0.69326097f)));
  P = fma_f32x2(P, F, pack_f32x2(__float_as_uint(0.99992818f), __float_as_uint(0.99992818f)));
  const uint32_t lo = static_cast<uint32_t>(P) + (static_cast<uint32_t>(XR) << 23);
  const uint32_t hi = static_cast<uint32_t>(P >> 32) + (static_cast<uint32_t>(XR >> 32) << 23);
  p0 = __uint_as_float(lo);
  p1 = __uint_as_float(hi);
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// A shared-memory matrix descriptor `bytes` further on (the start address is the low 14 bits, in 16-byte units; shared
// memory is < 256 KB, so the sum never carries out of the field).
__device__ __forceinline__ uint64_t desc_advance(uint64_t desc, uint32_t bytes) { return desc + (bytes >> 4); }

template <int DPAD, int NQ>
__global__ void __launch_bounds__(AttnCfg<DPAD, NQ>::THREADS, AttnCfg<DPAD, NQ>::OCC)
attention_kernel(const __grid_constant__ AttnParams p) {
  using Cfg = AttnCfg<DPAD, NQ>;
  constexpr int DB = Cfg::DB, NSTK = Cfg::NSTK, NSTV = Cfg::NSTV;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                              // [NQ][QT_BYTES]
  uint8_t* sP = sQ + NQ * Cfg::QT_BYTES;           // [NQ][P_BYTES]
  uint8_t* sK = sP + NQ * Cfg::P_BYTES;            // [NSTK][KV_BYTES]
  uint8_t* sV = sK + NSTK * Cfg::KV_BYTES;         // [NSTV][KV_BYTES]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + NSTV * Cfg::KV_BYTES);
  uint64_t* q_full = bars;             // 1
  uint64_t* k_full = bars + 1;         // NSTK
  uint64_t* k_empty = k_full + NSTK;   // NSTK  (one arrival per live tile: its S_j has completed)
  uint64_t* v_full = k_empty + NSTK;   // NSTV
  uint64_t* v_empty = v_full + NSTV;   // NSTV  (one arrival per live tile: its PV_j has completed)
  uint64_t* s_full = v_empty + NSTV;   // NQ
  uint64_t* p_full = s_full + NQ;      // NQ    (128 arrivals: the tile's softmax threads)
  uint64_t* pv_done = p_full + NQ;     // NQ
  uint64_t* s_free = pv_done + NQ;     // NQ    (128 arrivals: S_j has been read out)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_free + NQ);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * (BLK * NQ);
  const int head = blockIdx.y;
  const int batch = blockIdx.z;
  const int nkv = p.kv_blocks;
  const int live = min(NQ, (p.Nq - q0 + BLK - 1) / BLK);  // query tiles of this CTA that exist (>= 1)

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < NSTK; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], live); }
    for (int i = 0; i < NSTV; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], live); }
    for (int t = 0; t < NQ; ++t) {
      mbar_init(&s_full[t], 1); mbar_init(&p_full[t], 128); mbar_init(&pv_done[t], 1); mbar_init(&s_free[t], 128);
    }
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish(); }
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&p.tmQ); tma_prefetch_desc(&p.tmK); tma_prefetch_desc(&p.tmV); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // the prologue above overlapped the previous kernel's tail

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      mbar_arrive_expect_tx(q_full, live * Cfg::QT_BYTES);
      for (int t = 0; t < live; ++t)
        for (int db = 0; db < DB; ++db)
          tma_load_4d(sQ + t * Cfg::QT_BYTES + db * TILE16K, &p.tmQ, q_full, db * 64, head, q0 + t * BLK, batch);
      int tr_n = 0; (void)tr_n;
      for (int j = 0; j < nkv; ++j) {
        const int sk = j % NSTK, sv = j % NSTV;
        mbar_wait(&k_empty[sk], ((j / NSTK) & 1) ^ 1);
        ATT_TR(0, 9);
#ifdef B200SD_ATT_NOKV  // diagnostic build only: K/V blocks after the first ring fill are not loaded
        if (j >= NSTK) { mbar_arrive(&k_full[sk]); } else
#endif
        {
          mbar_arrive_expect_tx(&k_full[sk], Cfg::KV_BYTES);
          for (int db = 0; db < DB; ++db)
            tma_load_4d(sK + sk * Cfg::KV_BYTES + db * TILE8K, &p.tmK, &k_full[sk], db * 64, head, j * KB, batch);
        }
        mbar_wait(&v_empty[sv], ((j / NSTV) & 1) ^ 1);
        ATT_TR(0, 10);
#ifdef B200SD_ATT_NOKV
        if (j >= NSTV) { mbar_arrive(&v_full[sv]); } else
#endif
        {
          mbar_arrive_expect_tx(&v_full[sv], Cfg::KV_BYTES);
          for (int db = 0; db < DB; ++db)
            tma_load_4d(sV + sv * Cfg::KV_BYTES + db * TILE8K, &p.tmV, &v_full[sv], db * 64, head, j * KB, batch);
        }
      }
    }
  } else if (warp == 1 || warp == 2 + 4 * NQ) {
    // ===================== MMA issuers =====================
    // Per tile and block j there are two hand-offs: the tile's softmax warpgroup has read S_j out -> issue S_{j+1};
    // it has written P_j -> issue PV_j.  Such a pair of services costs ~500 cycles of the issuing warp
    // (it shares its SM sub-partition with softmax warps), so four tiles get two issuer warps; each serves its tiles
    // in a FIXED round-robin order with blocking waits.  The tiles settle into a staggered steady state (a tile's
    // softmax runs while the next one is served).  Measured alternatives at N = 4096, d = 40: one warp for all four
    // tiles 809 us, one warp polling the barriers of four tiles 1022 us (a probe is a serialized shared-memory round
    // trip), one warp per tile 700 us, this 670 us.  In-order service is deadlock-free: block j of every tile of a warp
    // is issued before block j+1 of any, and the two warps depend on each other only through the K/V ring, which the
    // producer refills as soon as the LAST tile's MMA has released a slot.  All shared-memory descriptors are built
    // once, up front; a tile or ring slot is a constant added to the start-address field.
    constexpr int TPW = NQ / Cfg::NMMA;            // tiles per MMA warp
    const int t_first = warp == 1 ? 0 : TPW;       // this warp's tiles: [t_first, t_first + TPW)
    constexpr int KS_MAX = DPAD / 16;
    const uint32_t idesc_s = umma_idesc_bf16(BLK, KB, 0);
    const uint32_t idesc_o = umma_idesc_bf16(BLK, p.n_pv, 1);
    uint64_t dq[KS_MAX], dk[KS_MAX], dp[KB / 16], dv[KB / 16];
    {
      const uint32_t q_addr = smem_u32(sQ), k_addr = smem_u32(sK), p_addr = smem_u32(sP), v_addr = smem_u32(sV);
#pragma unroll
      for (int ks = 0; ks < KS_MAX; ++ks) {
        dq[ks] = umma_smem_desc_sw128(q_addr + (ks >> 2) * TILE16K + (ks & 3) * 32, 16, 1024);
        dk[ks] = umma_smem_desc_sw128(k_addr + (ks >> 2) * TILE8K + (ks & 3) * 32, 16, 1024);
      }
#pragma unroll
      for (int ks = 0; ks < KB / 16; ++ks) {
        dp[ks] = umma_smem_desc_sw128(p_addr + ks * 32, 16, 1024);
        // V: MN-major; 16 keys = 2 groups of 8 rows (SBO 1024), 64-wide d chunks TILE8K apart (LBO)
        dv[ks] = umma_smem_desc_sw128(v_addr + ks * 2048, TILE8K, 1024);
      }
    }
    int tr_n = 0; (void)tr_n;
    // S_t(jn) = Q_t K_jn^T (one elected lane); the K slot gets this tile's release when the product completes
    auto issue_S = [&](int t, int jn) {
      const uint32_t q_off = t * Cfg::QT_BYTES, k_off = (jn % NSTK) * Cfg::KV_BYTES;
      const uint32_t tmem_S = tmem_base + t * Cfg::TILE_COLS;
#pragma unroll
      for (int ks = 0; ks < KS_MAX; ++ks)
        if (ks < p.ksteps_qk)
          umma_bf16_ss(tmem_S, desc_advance(dq[ks], q_off), desc_advance(dk[ks], k_off), idesc_s, ks != 0 ? 1u : 0u);
      umma_commit(&s_full[t]);
      umma_commit(&k_empty[jn % NSTK]);
    };
    mbar_wait(q_full, 0);
    mbar_wait(&k_full[0], 0);
    tc_fence_after();
    if (elect_one()) {
      for (int t = t_first; t < min(live, t_first + TPW); ++t) issue_S(t, 0);
    }
    __syncwarp();
    if constexpr (ATT_TS) {
      // One service per tile and block: P_j (in TMEM, over S_j's first 32 columns) is ready -> PV_j, then S_{j+1}
      // into the same columns.  The tensor core executes a thread's MMAs in issue order, so S_{j+1} cannot overwrite
      // P_j before PV_j has consumed it, and s_full(j+1) also tells the softmax warps that PV_j is complete.
      for (int j = 0; j < nkv; ++j) {
        const int st = j % NSTV;
        mbar_wait(&v_full[st], (j / NSTV) & 1);
        if (j + 1 < nkv) mbar_wait(&k_full[(j + 1) % NSTK], ((j + 1) / NSTK) & 1);
#pragma unroll
        for (int tt = 0; tt < TPW; ++tt) {
          const int t = t_first + tt;
          if (t < live) {
            mbar_wait(&p_full[t], j & 1);  // P_j of tile t written
            tc_fence_after();
            if (elect_one()) {
              if (t == 0) ATT_TR(1, 7);
              const uint32_t v_off = st * Cfg::KV_BYTES;
              const uint32_t tmem_P = tmem_base + t * Cfg::TILE_COLS;
              const uint32_t tmem_O = tmem_P + KB;
#pragma unroll
              for (int ks = 0; ks < KB / 16; ++ks)
                umma_bf16_ts(tmem_O, tmem_P + ks * 8, desc_advance(dv[ks], v_off), idesc_o, (j | ks) != 0 ? 1u : 0u);
              umma_commit(&v_empty[st]);
              if (j + 1 < nkv) issue_S(t, j + 1);
              else umma_commit(&pv_done[t]);
              if (t == 0) ATT_TR(1, 8);
            }
            __syncwarp();
          }
        }
      }
    } else {
    for (int j = 0; j < nkv; ++j) {
      if (j + 1 < nkv) {
        mbar_wait(&k_full[(j + 1) % NSTK], ((j + 1) / NSTK) & 1);  // K_{j+1}: loaded a block or more ago
#pragma unroll
        for (int tt = 0; tt < TPW; ++tt) {
          const int t = t_first + tt;
          if (t < live) {
            mbar_wait(&s_free[t], j & 1);  // S_j of tile t read out (half a block before P_j is complete)
            tc_fence_after();
            if (elect_one()) {
              if (t == 0) ATT_TR(1, 4);
              issue_S(t, j + 1);
              if (t == 0) ATT_TR(1, 6);
            }
            __syncwarp();
          }
        }
      }
      const int st = j % NSTV;
      mbar_wait(&v_full[st], (j / NSTV) & 1);  // V_j: likewise
#pragma unroll
      for (int tt = 0; tt < TPW; ++tt) {
        const int t = t_first + tt;
        if (t < live) {
          mbar_wait(&p_full[t], j & 1);  // P_j of tile t written
          tc_fence_after();
          if (elect_one()) {
            if (t == 0) ATT_TR(1, 7);
            const uint32_t p_off = t * TILE16K, v_off = st * Cfg::KV_BYTES;
            const uint32_t tmem_O = tmem_base + t * Cfg::TILE_COLS + KB;
#pragma unroll
            for (int ks = 0; ks < KB / 16; ++ks)
              umma_bf16_ss(tmem_O, desc_advance(dp[ks], p_off), desc_advance(dv[ks], v_off), idesc_o, (j | ks) != 0 ? 1u : 0u);
            umma_commit(&pv_done[t]);
            umma_commit(&v_empty[st]);
            if (t == 0) ATT_TR(1, 8);
          }
          __syncwarp();
        }
      }
    }
    }
  } else {
    // ===================== softmax / correction / epilogue: thread <-> query row =====================
    const int t = (warp - 2) >> 2;  // query tile of this warpgroup
    const int q = warp & 3;         // TMEM lane quarter this warp may touch
    const int r = q * 32 + lane;
    const uint32_t lane_base = static_cast<uint32_t>(q * 32) << 16;
    const uint32_t s_addr = tmem_base + t * Cfg::TILE_COLS + lane_base;
    const uint32_t o_addr = s_addr + KB;
    const uint32_t prow = smem_u32(sP) + t * TILE16K + r * 128;
    uint64_t* const my_s_full = &s_full[t];
    uint64_t* const my_p_full = &p_full[t];
    uint64_t* const my_pv_done = &pv_done[t];
    uint64_t* const my_s_free = &s_free[t];
    const int nblk = t < live ? nkv : 0;  // a warpgroup whose tile lies beyond Nq has nothing to do
    const int rsw = r & 7;
    float m_run = -INFINITY;  // softmax reference point of this row, scaled log2 domain
    float l_run = 0.f;
    int tr_n = 0; (void)tr_n;
    // Lazy running max (single pass over S on every block but the first): probabilities are taken relative to the
    // max of the keys BEFORE this block, and the block's own max is tracked in the same pass.  Softmax is invariant
    // to the reference point; bf16 P and the fp32 row sum keep their relative precision at any magnitude, so the
    // reference only has to move (and O be rescaled) when a row's scores outgrow it by more than 2^LAZY_TAU — then
    // the warp redoes its block with the exact max (S_j is still in TMEM: the warp has not released it yet).
    // S_j is released to the MMA warp as soon as its second half sits in registers (and its max is known to be
    // harmless), i.e. half a block BEFORE P_j is complete: S_{j+1} is computed under the second half's exponentials,
    // so the warpgroup goes from block to block without waiting for the tensor core.
    constexpr float LAZY_TAU = 8.0f;
    for (int j = 0; j < nblk; ++j) {
      const int valid = min(KB, p.Nkv - j * KB);  // keys of this block that exist
      // keys of this block this row may see (causal: those up to its own position; <= 0 when the block is all future)
      const int lim = p.causal ? min(valid, q0 + t * BLK + r + 1 - j * KB) : valid;
      mbar_wait(my_s_full, j & 1);
      tc_fence_after();
      bool p_busy = j > 0;  // PV_{j-1} may still be reading P when S_j lands
      if (q == 2 && lane == 0) ATT_TR(2 + t, 1);
      // max of the valid raw scores of 32-key chunk c
      auto raw_max = [&](auto masked_tag, int c, const uint32_t (&u)[32]) {
        constexpr bool MASKED = decltype(masked_tag)::value;
        float mx = -INFINITY;
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (!MASKED || c * 32 + i < lim) mx = fmaxf(mx, __uint_as_float(u[i]));
        return mx;
      };
      // chunk c: probabilities relative to m_ref -> bf16 -> swizzled smem (K-major, 128-byte rows, 16-byte piece ^=
      // row % 8); returns the chunk's row sum
      auto exp_store = [&](auto masked_tag, int c, const uint32_t (&u)[32], const float m_ref) {
        constexpr bool MASKED = decltype(masked_tag)::value;
        float rsum = 0.f;
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float p0 = ex2_approx(fmaf(__uint_as_float(u[2 * i]), p.scale_log2, -m_ref));
          float p1 = ex2_approx(fmaf(__uint_as_float(u[2 * i + 1]), p.scale_log2, -m_ref));
          if (MASKED) {
            if (c * 32 + 2 * i >= lim) p0 = 0.f;
            if (c * 32 + 2 * i + 1 >= lim) p1 = 0.f;
          }
          rsum += p0 + p1;
          pk[i] = pack_bf16x2(p0, p1);
        }
        if (p_busy) {  // (complete long before the first chunk's exponentials are)
          mbar_wait(my_pv_done, (j - 1) & 1);
          p_busy = false;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
          sts128(prow + (((c * 4 + i) ^ rsw) << 4), pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
        return rsum;
      };
      // The body is instantiated twice: the mask-free version runs on every full block (no per-element predicates).
      auto block_body = [&](auto masked_tag) {
        uint32_t u[32];
        if (j == 0) {  // very first block: exact max (nothing to rescale yet)
          float mx = -INFINITY;
#pragma unroll
          for (int c = 0; c < KB / 32; ++c) {
            tmem_ld32(s_addr + c * 32, u);
            tmem_ld_wait();
            mx = fmaxf(mx, raw_max(masked_tag, c, u));
          }
          m_run = mx * p.scale_log2;
        }
        tmem_ld32(s_addr, u);
        tmem_ld_wait();
        float mx = raw_max(masked_tag, 0, u);
        float rsum = exp_store(masked_tag, 0, u, m_run);
        tmem_ld32(s_addr + 32, u);
        tmem_ld_wait();
        mx = fmaxf(mx, raw_max(masked_tag, 1, u));
        const float m_blk = mx * p.scale_log2;
        const bool outgrown = m_blk > m_run + LAZY_TAU;
        if (!__any_sync(0xffffffffu, outgrown)) {
          tc_fence_before();
          mbar_arrive(my_s_free);  // S_j is in registers: S_{j+1} may overwrite it
          rsum += exp_store(masked_tag, 1, u, m_run);
        } else {  // rare; warp-uniform because tcgen05.ld/st are .sync.aligned
          float alpha = 1.0f;
          if (outgrown) {  // this row moves its reference to the exact max; the other rows just repeat
            alpha = ex2_approx(m_run - m_blk);
            m_run = m_blk;
          }
          rsum = exp_store(masked_tag, 1, u, m_run);
          tmem_ld32(s_addr, u);  // first half again, against the new reference
          tmem_ld_wait();
          tc_fence_before();
          mbar_arrive(my_s_free);
          rsum += exp_store(masked_tag, 0, u, m_run);
          l_run *= alpha;
          // O *= alpha once PV_{j-1} (the latest PV issued) has landed
          if (j > 0) mbar_wait(my_pv_done, (j - 1) & 1);
          tc_fence_after();
#pragma unroll 1
          for (int c = 0; c < p.n_pv / 16; ++c) {
            uint32_t o[16];
            tmem_ld16(o_addr + c * 16, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st16(o_addr + c * 16, o);
          }
          tmem_st_wait();
        }
        l_run += rsum;
      };
      // chunk c: probabilities relative to m_ref -> packed bf16 pairs in registers; returns the chunk's row sum
      auto exp_pack = [&](auto masked_tag, int c, const uint32_t (&u)[32], const float m_ref, uint32_t (&pk)[16]) {
        constexpr bool MASKED = decltype(masked_tag)::value;
        const uint64_t sc2 = pack_f32x2(__float_as_uint(p.scale_log2), __float_as_uint(p.scale_log2));
        const uint64_t nm2 = pack_f32x2(__float_as_uint(-m_ref), __float_as_uint(-m_ref));
        uint64_t acc = 0;  // two partial row sums
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float x0, x1;
          unpack_f32x2(fma_f32x2(pack_f32x2(u[2 * i], u[2 * i + 1]), sc2, nm2), x0, x1);
          float p0, p1;
          constexpr int PE = B200SD_ATT_POLY_EVERY > 0 ? B200SD_ATT_POLY_EVERY : 1;
          if (B200SD_ATT_POLY_EVERY > 0 && DPAD == 64 && i % PE == PE - 1) {  // (d > 64: two tiles per CTA, not XU-bound: 93.2 vs 91.1 us)
            ex2_poly_f32x2(x0, x1, p0, p1);
          } else {
            p0 = ex2_approx(x0);
            p1 = ex2_approx(x1);
          }
          if (MASKED) {
            if (c * 32 + 2 * i >= lim) p0 = 0.f;
            if (c * 32 + 2 * i + 1 >= lim) p1 = 0.f;
          }
          acc = add_f32x2(acc, pack_f32x2(__float_as_uint(p0), __float_as_uint(p1)));
          pk[i] = pack_bf16x2(p0, p1);
        }
        float r0, r1;
        unpack_f32x2(acc, r0, r1);
        return r0 + r1;
      };
      // TS form.  P_j (64 keys = 32 packed columns) overwrites columns [0, 32) of S_j: the first half is stored only
      // after the second half of S_j sits in registers and the lazy-max vote has passed, so the rare redo still finds
      // S_j intact.  s_full(j) implies PV_{j-1} has completed (same issuing thread, in order): no other wait is needed
      // before P_{j-1} / O are touched.
      auto block_body_ts = [&](auto masked_tag) {
        uint32_t u[32], pk0[16], pk1[16];
        if (j == 0) {  // very first block: exact max (nothing to rescale yet)
          float mx = -INFINITY;
#pragma unroll
          for (int c = 0; c < KB / 32; ++c) {
            tmem_ld32(s_addr + c * 32, u);
            tmem_ld_wait();
            mx = fmaxf(mx, raw_max(masked_tag, c, u));
          }
          m_run = mx * p.scale_log2;
        }
        tmem_ld32(s_addr, u);
        tmem_ld_wait();
        float mx = raw_max(masked_tag, 0, u);
        float rsum = exp_pack(masked_tag, 0, u, m_run, pk0);
        tmem_ld32(s_addr + 32, u);
        tmem_ld_wait();
        mx = fmaxf(mx, raw_max(masked_tag, 1, u));
        const float m_blk = mx * p.scale_log2;
        const bool outgrown = m_blk > m_run + LAZY_TAU;
        if (!__any_sync(0xffffffffu, outgrown)) {
          tmem_st16(s_addr, pk0);
          rsum += exp_pack(masked_tag, 1, u, m_run, pk1);
          tmem_st16(s_addr + 16, pk1);
        } else {  // rare; warp-uniform because tcgen05.ld/st are .sync.aligned
          float alpha = 1.0f;
          if (outgrown) {  // this row moves its reference to the exact max; the other rows just repeat
            alpha = ex2_approx(m_run - m_blk);
            m_run = m_blk;
          }
          rsum = exp_pack(masked_tag, 1, u, m_run, pk1);
          tmem_ld32(s_addr, u);  // first half again (not yet overwritten), against the new reference
          tmem_ld_wait();
          rsum += exp_pack(masked_tag, 0, u, m_run, pk0);
          tmem_st16(s_addr, pk0);
          tmem_st16(s_addr + 16, pk1);
          l_run *= alpha;
#pragma unroll 1
          for (int c = 0; c < p.n_pv / 16; ++c) {  // O *= alpha (PV_{j-1} has landed, PV_j is not issued before p_full)
            uint32_t o[16];
            tmem_ld16(o_addr + c * 16, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st16(o_addr + c * 16, o);
          }
        }
        tmem_st_wait();
        l_run += rsum;
      };
      if constexpr (ATT_TS) {
        if (valid == KB && !p.causal) block_body_ts(std::false_type{});
        else block_body_ts(std::true_type{});
      } else {
        if (valid == KB && !p.causal) block_body(std::false_type{});
        else block_body(std::true_type{});
        fence_proxy_async_smem();  // P (generic-proxy stores) -> visible to the MMA's async-proxy reads
      }
      if (q == 2 && lane == 0) ATT_TR(2 + t, 2);
      tc_fence_before();
      mbar_arrive(my_p_full);
      if (q == 2 && lane == 0) ATT_TR(2 + t, 3);
    }
    // ---- epilogue: O / l -> bf16 -> global ----
    if (nblk > 0) mbar_wait(my_pv_done, ATT_TS ? 0 : (nkv - 1) & 1);  // (TS form: committed once, after the last PV)
    tc_fence_after();
    const float inv_l = 1.0f / l_run;
    const int qrow = nblk > 0 ? q0 + t * BLK + r : p.Nq;
    bf16* orow = p.out + (static_cast<long long>(batch) * p.Nq + qrow) * p.ldo + head * p.d;
#pragma unroll 1
    for (int c = 0; c < p.n_pv / 16; ++c) {
      uint32_t u[16];
      tmem_ld16(o_addr + c * 16, u);
      tmem_ld_wait();
      if (qrow < p.Nq) {
        uint32_t o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i)
          o[i] = pack_bf16x2(__uint_as_float(u[2 * i]) * inv_l, __uint_as_float(u[2 * i + 1]) * inv_l);
        if (c * 16 + 8 <= p.d) *reinterpret_cast<uint4*>(orow + c * 16) = make_uint4(o[0], o[1], o[2], o[3]);
        if (c * 16 + 16 <= p.d) *reinterpret_cast<uint4*>(orow + c * 16 + 8) = make_uint4(o[4], o[5], o[6], o[7]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}


// ======================================================================================================================
// Short-KV attention (Nkv <= 128, no mask): the cross-attention of every BasicTransformerBlock (77 prompt tokens).
// The general kernel spends such a launch on per-CTA latency: one 128-row tile per CTA means barrier init, TMEM
// allocation, a cold Q/K/V load, two S -> softmax -> PV round trips and the teardown in series, ~8 us per CTA for 32 KB of
// traffic (in-graph, 16x8x4096x77, d = 40: 55 us for 84 MB, 1.5 TB/s).  Here CTAs are persistent over a contiguous range
// of (batch, head, query tile) items: TMEM and barriers are set up once, K/V of a (batch, head) are loaded once per run of
// tiles, the next tile's Q is in flight while this one is computed (two Q slots), both key blocks are multiplied up front
// (S is 128 columns) so the softmax is a single exact pass (max, then exponentials), P goes back over S in tensor memory
// and PV_i is followed at once by S_{i+1}.
// ======================================================================================================================
template <int DPAD>
struct ShortCfg {
  static constexpr int DB = DPAD / 64;
  static constexpr int THREADS = 192;                          // warp 0 TMA, warp 1 MMA, warps 2-5 softmax + epilogue
  static constexpr int KVSLOTS = DPAD == 64 ? 2 : 1;
  static constexpr int OCC = DPAD == 64 ? 2 : 1;
  static constexpr int QT_BYTES = DB * TILE16K;
  static constexpr int KV_BYTES = 2 * DB * TILE8K;             // K (or V) of one (batch, head): two 64-key blocks
  static constexpr int TMEM_COLS = 128 + DPAD <= 256 ? 256 : 512;  // S (128) then O
  static constexpr int SMEM = 2 * QT_BYTES + KVSLOTS * 2 * KV_BYTES + 1024 + 256;
  static_assert(OCC * (SMEM + 1024) <= 233472 && OCC * TMEM_COLS <= 512, "budget");
};

template <int DPAD>
__global__ void __launch_bounds__(ShortCfg<DPAD>::THREADS, ShortCfg<DPAD>::OCC)
attention_shortkv_kernel(const __grid_constant__ AttnParams p) {
  using Cfg = ShortCfg<DPAD>;
  constexpr int DB = Cfg::DB, KVS = Cfg::KVSLOTS;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                               // [2][QT_BYTES]
  uint8_t* sK = sQ + 2 * Cfg::QT_BYTES;             // [KVS][KV_BYTES]
  uint8_t* sV = sK + KVS * Cfg::KV_BYTES;           // [KVS][KV_BYTES]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + KVS * Cfg::KV_BYTES);
  uint64_t* q_full = bars;            // 2
  uint64_t* q_empty = bars + 2;       // 2
  uint64_t* kv_full = bars + 4;       // KVS
  uint64_t* kv_empty = bars + 6;      // KVS
  uint64_t* s_full = bars + 8;        // S_i complete
  uint64_t* p_full = bars + 9;        // P_i written (128 arrivals)
  uint64_t* o_full = bars + 10;       // PV_i complete
  uint64_t* o_free = bars + 11;       // O_i read out (128 arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles = (p.Nq + BLK - 1) / BLK;
  const int total = tiles * p.heads * p.B;
  const int per_cta = (total + gridDim.x - 1) / gridDim.x;
  const int item0 = blockIdx.x * per_cta;
  const int n_items = max(0, min(per_cta, total - item0));
  const int nkb = p.kv_blocks;                        // 1 or 2
  const int n1 = nkb > 1 ? ((p.Nkv - KB + 15) / 16) * 16 : 0;  // UMMA N of the second key block

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); }
    for (int i = 0; i < KVS; ++i) { mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], 1); }
    mbar_init(s_full, 1); mbar_init(p_full, 128); mbar_init(o_full, 1); mbar_init(o_free, 128);
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish(); }
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&p.tmQ); tma_prefetch_desc(&p.tmK); tma_prefetch_desc(&p.tmV); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int run = -1, prev_bh = -1;
      for (int it = 0; it < n_items; ++it) {
        const int item = item0 + it;
        const int bh = item / tiles, qt = item - bh * tiles;
        const int batch = bh / p.heads, head = bh - batch * p.heads;
        if (bh != prev_bh) {  // new (batch, head): its K and V, once
          prev_bh = bh; ++run;
          const int sl = run % KVS;
          mbar_wait(&kv_empty[sl], ((run / KVS) & 1) ^ 1);
          mbar_arrive_expect_tx(&kv_full[sl], 2 * nkb * DB * TILE8K);
          for (int b = 0; b < nkb; ++b)
            for (int db = 0; db < DB; ++db) {
              tma_load_4d(sK + sl * Cfg::KV_BYTES + (b * DB + db) * TILE8K, &p.tmK, &kv_full[sl], db * 64, head, b * KB, batch);
              tma_load_4d(sV + sl * Cfg::KV_BYTES + (b * DB + db) * TILE8K, &p.tmV, &kv_full[sl], db * 64, head, b * KB, batch);
            }
        }
        const int qs = it & 1;
        mbar_wait(&q_empty[qs], ((it >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&q_full[qs], Cfg::QT_BYTES);
        for (int db = 0; db < DB; ++db)
          tma_load_4d(sQ + qs * Cfg::QT_BYTES + db * TILE16K, &p.tmQ, &q_full[qs], db * 64, head, qt * BLK, batch);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr int KS_MAX = DPAD / 16;
    const uint32_t idesc_s0 = umma_idesc_bf16(BLK, KB, 0);
    const uint32_t idesc_s1 = umma_idesc_bf16(BLK, n1 > 0 ? n1 : 16, 0);
    const uint32_t idesc_o = umma_idesc_bf16(BLK, p.n_pv, 1);
    uint64_t dq[KS_MAX], dk[KS_MAX];
    const uint32_t q_addr = smem_u32(sQ), k_addr = smem_u32(sK), v_addr = smem_u32(sV);
#pragma unroll
    for (int ks = 0; ks < KS_MAX; ++ks) {
      dq[ks] = umma_smem_desc_sw128(q_addr + (ks >> 2) * TILE16K + (ks & 3) * 32, 16, 1024);
      dk[ks] = umma_smem_desc_sw128(k_addr + (ks >> 2) * TILE8K + (ks & 3) * 32, 16, 1024);
    }
    const uint64_t dv0 = umma_smem_desc_sw128(v_addr, TILE8K, 1024);  // MN-major; 16 keys = 2048 bytes further on
    const int pv_steps = (p.Nkv + 15) / 16;
    int run = -1, prev_bh = -1;
    auto item_bh = [&](int it) { return (item0 + it) / tiles; };
    // S_i = Q_i [K_0 | K_1]^T into columns [0, 64) and [64, 64 + n1)
    auto issue_S = [&](int it, int sl) {
      const uint32_t q_off = (it & 1) * Cfg::QT_BYTES, k_off = sl * Cfg::KV_BYTES;
#pragma unroll
      for (int ks = 0; ks < KS_MAX; ++ks)
        if (ks < p.ksteps_qk)
          umma_bf16_ss(tmem_base, desc_advance(dq[ks], q_off), desc_advance(dk[ks], k_off), idesc_s0, ks != 0 ? 1u : 0u);
      if (n1 > 0) {
#pragma unroll
        for (int ks = 0; ks < KS_MAX; ++ks)
          if (ks < p.ksteps_qk)
            umma_bf16_ss(tmem_base + KB, desc_advance(dq[ks], q_off), desc_advance(dk[ks], k_off + DB * TILE8K), idesc_s1,
                         ks != 0 ? 1u : 0u);
      }
      umma_commit(s_full);
    };
    auto wait_inputs = [&](int it) {  // Q of item `it` (and K/V when it opens a new run); returns the K/V slot
      const int bh = item_bh(it);
      if (bh != prev_bh) {
        prev_bh = bh; ++run;
        mbar_wait(&kv_full[run % KVS], (run / KVS) & 1);
      }
      mbar_wait(&q_full[it & 1], (it >> 1) & 1);
      tc_fence_after();
      return run % KVS;
    };
    int sl = 0;
    if (n_items > 0) {
      sl = wait_inputs(0);
      if (elect_one()) issue_S(0, sl);
      __syncwarp();
    }
    for (int it = 0; it < n_items; ++it) {
      mbar_wait(p_full, it & 1);                      // P_it sits over S_it
      if (it > 0) mbar_wait(o_free, (it - 1) & 1);    // O_{it-1} has been read out
      tc_fence_after();
      const bool last_of_run = it + 1 >= n_items || item_bh(it + 1) != item_bh(it);
      if (elect_one()) {
        const uint32_t v_off = sl * Cfg::KV_BYTES;
        for (int ks = 0; ks < pv_steps; ++ks) {
          // key group ks: block ks / 4 (KV_BYTES / 2 apart), 16 keys = 2048 bytes within the block
          const uint32_t off = v_off + (ks >> 2) * (DB * TILE8K) + (ks & 3) * 2048;
          umma_bf16_ts(tmem_base + 128, tmem_base + ks * 8, desc_advance(dv0, off), idesc_o, ks != 0 ? 1u : 0u);
        }
        umma_commit(o_full);
        umma_commit(&q_empty[it & 1]);
        if (last_of_run) umma_commit(&kv_empty[sl]);
      }
      __syncwarp();
      if (it + 1 < n_items) {
        sl = wait_inputs(it + 1);
        if (elect_one()) issue_S(it + 1, sl);   // after PV_it in issue order: it cannot overwrite P_it early
        __syncwarp();
      }
    }
  } else {
    // ===================== softmax + epilogue: thread <-> query row =====================
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t s_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t o_addr = s_addr + 128;
    const int nch = (p.Nkv + 31) / 32;  // 32-key chunks that hold valid keys
    for (int it = 0; it < n_items; ++it) {
      const int item = item0 + it;
      const int bh = item / tiles, qt = item - bh * tiles;
      const int batch = bh / p.heads, head = bh - batch * p.heads;
      mbar_wait(s_full, it & 1);
      tc_fence_after();
      uint32_t u[32];
      float mx = -INFINITY;
      for (int c = 0; c < nch; ++c) {
        tmem_ld32(s_addr + c * 32, u);
        tmem_ld_wait();
        if (c * 32 + 32 <= p.Nkv) {
#pragma unroll
          for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(u[i]));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) if (c * 32 + i < p.Nkv) mx = fmaxf(mx, __uint_as_float(u[i]));
        }
      }
      const float m_ref = mx * p.scale_log2;
      const uint64_t sc2 = pack_f32x2(__float_as_uint(p.scale_log2), __float_as_uint(p.scale_log2));
      const uint64_t nm2 = pack_f32x2(__float_as_uint(-m_ref), __float_as_uint(-m_ref));
      uint64_t acc = 0;
      for (int c = 0; c < nch; ++c) {
        tmem_ld32(s_addr + c * 32, u);
        tmem_ld_wait();
        uint32_t pk[16];
        const bool full = c * 32 + 32 <= p.Nkv;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float x0, x1;
          unpack_f32x2(fma_f32x2(pack_f32x2(u[2 * i], u[2 * i + 1]), sc2, nm2), x0, x1);
          float p0 = ex2_approx(x0), p1 = ex2_approx(x1);
          if (!full) {
            if (c * 32 + 2 * i >= p.Nkv) p0 = 0.f;
            if (c * 32 + 2 * i + 1 >= p.Nkv) p1 = 0.f;
          }
          acc = add_f32x2(acc, pack_f32x2(__float_as_uint(p0), __float_as_uint(p1)));
          pk[i] = pack_bf16x2(p0, p1);
        }
        tmem_st16(s_addr + c * 16, pk);  // P chunk c over columns already consumed (16c + 16 <= 32c + 32)
      }
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(p_full);
      float l0, l1;
      unpack_f32x2(acc, l0, l1);
      const float inv_l = 1.0f / (l0 + l1);
      // ---- epilogue: O / l -> bf16 -> global ----
      mbar_wait(o_full, it & 1);
      tc_fence_after();
      const int qrow = qt * BLK + r;
      bf16* orow = p.out + (static_cast<long long>(batch) * p.Nq + qrow) * p.ldo + head * p.d;
#pragma unroll 1
      for (int c = 0; c < p.n_pv / 16; ++c) {
        uint32_t o[16];
        tmem_ld16(o_addr + c * 16, o);
        tmem_ld_wait();
        if (c == p.n_pv / 16 - 1) {  // O is in registers: PV of the next item may overwrite it
          tc_fence_before();
          mbar_arrive(o_free);
        }
        if (qrow < p.Nq) {
          uint32_t w[8];
#pragma unroll
          for (int i = 0; i < 8; ++i)
            w[i] = pack_bf16x2(__uint_as_float(o[2 * i]) * inv_l, __uint_as_float(o[2 * i + 1]) * inv_l);
          if (c * 16 + 8 <= p.d) *reinterpret_cast<uint4*>(orow + c * 16) = make_uint4(w[0], w[1], w[2], w[3]);
          if (c * 16 + 16 <= p.d) *reinterpret_cast<uint4*>(orow + c * 16 + 8) = make_uint4(w[4], w[5], w[6], w[7]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

struct AttnPlan {
  AttnParams kp;
  int dpad, nq;  // nq == 0: the short-KV kernel
  dim3 grid;
};

template <int DPAD, int NQ>
int prepare_attention_t() {  // plan time, outside any stream capture
  static bool attr_set = false;
  if (!attr_set) {
    B200SD_CUDA(cudaFuncSetAttribute(attention_kernel<DPAD, NQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnCfg<DPAD, NQ>::SMEM));
    attr_set = true;
  }
  return 0;
}
template <int DPAD, int NQ>
int launch_attention_t(const AttnPlan& pl, cudaStream_t s) {
  using Cfg = AttnCfg<DPAD, NQ>;
  B200SD_CUDA(launch_kernel(attention_kernel<DPAD, NQ>, pl.grid, dim3(Cfg::THREADS), Cfg::SMEM, s, 1, pl.kp));
  return 0;
}
// One switch for both uses: LAUNCH = false sets the kernel attribute, true launches.
template <bool LAUNCH>
int dispatch_attention(int dpad, int nq, const AttnPlan* pl, cudaStream_t s) {
#define B200SD_ATT_CASE(D, Q) \
  if (dpad == D && nq == Q) { if constexpr (LAUNCH) return launch_attention_t<D, Q>(*pl, s); else return prepare_attention_t<D, Q>(); }
  B200SD_ATT_CASE(64, 4)
  B200SD_ATT_CASE(64, 1)
  B200SD_ATT_CASE(128, 2)
  B200SD_ATT_CASE(128, 1)
  B200SD_ATT_CASE(192, 2)
  B200SD_ATT_CASE(192, 1)
#undef B200SD_ATT_CASE
  return fail(B200SD_ERR_UNSUPPORTED, "attention: padded head dim %d with %d query tiles per CTA unsupported", dpad, nq);
}
template <int DPAD>
int prepare_shortkv_t() {
  static bool attr_set = false;
  if (!attr_set) {
    B200SD_CUDA(cudaFuncSetAttribute(attention_shortkv_kernel<DPAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, ShortCfg<DPAD>::SMEM));
    attr_set = true;
  }
  return 0;
}
template <bool LAUNCH>
int dispatch_shortkv(int dpad, const AttnPlan* pl, cudaStream_t s) {
#define B200SD_ATT_SHORT_CASE(D)                                                                                        \
  if (dpad == D) {                                                                                                      \
    if constexpr (LAUNCH) {                                                                                             \
      B200SD_CUDA(launch_kernel(attention_shortkv_kernel<D>, pl->grid, dim3(ShortCfg<D>::THREADS), ShortCfg<D>::SMEM, s, 1, pl->kp)); \
      return 0;                                                                                                         \
    } else {                                                                                                            \
      return prepare_shortkv_t<D>();                                                                                    \
    }                                                                                                                   \
  }
  B200SD_ATT_SHORT_CASE(64)
  B200SD_ATT_SHORT_CASE(128)
  B200SD_ATT_SHORT_CASE(192)
#undef B200SD_ATT_SHORT_CASE
  return fail(B200SD_ERR_UNSUPPORTED, "attention: padded head dim %d unsupported", dpad);
}
int shortkv_occupancy(int dpad) { return dpad == 64 ? 2 : 1; }
int launch_attention(const void* plan, cudaStream_t s) {
  const AttnPlan& pl = *static_cast<const AttnPlan*>(plan);
  if (pl.nq == 0) return dispatch_shortkv<true>(pl.dpad, &pl, s);
  return dispatch_attention<true>(pl.dpad, pl.nq, &pl, s);
}
int wide_tiles(int dpad) { return dpad == 64 ? 4 : 2; }
int occupancy(int dpad, int nq) { return nq > 1 ? 1 : (dpad == 64 ? 4 : (dpad == 128 ? 2 : 1)); }

// Query tiles per CTA for a problem: the wide CTA (K/V blocks shared by its tiles, deep rings) unless its grid
// quantizes worse on the 148 SMs than the single-tile one (B200SD_ATT_NQ overrides, for tuning).
int choose_tiles(int dpad, int B, int heads, int Nq, int kv_blocks) {
  if (const char* e = std::getenv("B200SD_ATT_NQ")) {
    const int v = std::atoi(e);
    if (v == 1 || v == wide_tiles(dpad)) return v;
  }
  if (kv_blocks <= 2) return 1;  // cross-attention (77 keys): the run is all prologue, which the small CTA keeps short
  const int sms = sm_count();
  const long long tiles = static_cast<long long>((Nq + BLK - 1) / BLK);
  auto cost = [&](int nq) {
    const long long ctas = ((tiles + nq - 1) / nq) * heads * B;
    const double per_sm = static_cast<double>((ctas + sms - 1) / sms);  // CTAs the busiest SM works through
    // per-CTA time in units of one tile running alone on an SM share: a single-tile CTA runs `occ` to an SM; a wide one
    // owns the SM for its nq tiles, each ~12% cheaper for sharing K/V (measured 0.77-0.93 across the SD shapes)
    const double per_cta = (nq == 1 ? 1.0 : 0.88 * nq) / static_cast<double>(occupancy(dpad, 1));
    return per_sm * per_cta;
  };
  const int wide = wide_tiles(dpad);
  return cost(wide) < cost(1) ? wide : 1;
}

}  // namespace

int plan_attention(const b200sd_attention& a, PlannedOp* out) {
  B200SD_CHECK_ARG(a.q && a.k && a.v && a.out, "attention: NULL pointer");
  B200SD_CHECK_ARG(a.B >= 1 && a.heads >= 1 && a.Nq >= 1 && a.Nkv >= 1, "attention: bad shape");
  B200SD_CHECK_ARG(!a.causal || a.Nq == a.Nkv, "attention: the causal mask needs Nq == Nkv");
  B200SD_CHECK_ARG(a.d >= 8 && a.d % 8 == 0 && a.d <= 192, "attention: head dim %d unsupported (multiple of 8, <= 192)", a.d);
  B200SD_CHECK_ARG(a.ldq % 8 == 0 && a.ldk % 8 == 0 && a.ldv % 8 == 0 && a.ldo % 8 == 0, "attention: pitches must be multiples of 8");
  B200SD_CHECK_ARG(a.ldq >= a.heads * a.d && a.ldk >= a.heads * a.d && a.ldv >= a.heads * a.d && a.ldo >= a.heads * a.d,
                   "attention: pitch smaller than heads*d");
  for (const void* ptr : {a.q, a.k, a.v, (const void*)a.out})
    B200SD_CHECK_ARG((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "attention: pointers must be 16-byte aligned");
  AttnPlan* pl = new (std::nothrow) AttnPlan();
  if (!pl) return fail(B200SD_ERR_CUDA, "out of host memory");
  AttnParams& kp = pl->kp;
  memset(&kp, 0, sizeof(kp));
  auto make_map = [&](CUtensorMap* m, const void* base, long long ld, int rows, int box_rows) {
    uint64_t dims[4] = {(uint64_t)a.d, (uint64_t)a.heads, (uint64_t)rows, (uint64_t)a.B};
    uint64_t st[3] = {(uint64_t)a.d * 2, (uint64_t)ld * 2, (uint64_t)rows * ld * 2};
    uint32_t box[4] = {64, 1, (uint32_t)box_rows, 1};
    return encode_tensor_map_bf16(m, base, 4, dims, st, box);
  };
  int rc = make_map(&kp.tmQ, a.q, a.ldq, a.Nq, BLK);
  if (!rc) rc = make_map(&kp.tmK, a.k, a.ldk, a.Nkv, KB);
  if (!rc) rc = make_map(&kp.tmV, a.v, a.ldv, a.Nkv, KB);
  if (rc) { delete pl; return rc; }
  kp.out = static_cast<bf16*>(a.out);
  kp.ldo = a.ldo;
  kp.B = a.B; kp.heads = a.heads; kp.Nq = a.Nq; kp.Nkv = a.Nkv; kp.d = a.d;
  kp.kv_blocks = (a.Nkv + KB - 1) / KB;
  kp.n_pv = (a.d + 15) / 16 * 16;
  kp.ksteps_qk = (a.d + 15) / 16;
  kp.scale_log2 = a.scale * 1.4426950408889634f;
  kp.causal = a.causal ? 1 : 0;
  pl->dpad = (a.d + 63) / 64 * 64;
  static const bool shortkv_on = [] { const char* e = std::getenv("B200SD_ATT_SHORTKV"); return !(e && e[0] == '0'); }();
  if (shortkv_on && !a.causal && a.Nkv <= 2 * KB) {
    pl->nq = 0;
    const long long items = static_cast<long long>((a.Nq + BLK - 1) / BLK) * a.heads * a.B;
    pl->grid = dim3(static_cast<unsigned>(std::min<long long>(items, static_cast<long long>(sm_count()) * shortkv_occupancy(pl->dpad))));
    rc = dispatch_shortkv<false>(pl->dpad, nullptr, nullptr);
  } else {
    pl->nq = choose_tiles(pl->dpad, a.B, a.heads, a.Nq, kp.kv_blocks);
    const int rows_per_cta = BLK * pl->nq;
    pl->grid = dim3((a.Nq + rows_per_cta - 1) / rows_per_cta, a.heads, a.B);
    rc = dispatch_attention<false>(pl->dpad, pl->nq, nullptr, nullptr);
  }
  if (rc) { delete pl; return rc; }
  out->kind = B200SD_OP_ATTENTION;
  out->plan = pl;
  out->launch = launch_attention;
  out->destroy = destroy_plan<AttnPlan>;
  out->launches = 1;
  return 0;
}

}  // namespace b200sd
