// Fused flash-style attention on tcgen05 (sm_100a):  O = softmax(scale * Q K^T) V  per (batch, head).
// Replaces F.scaled_dot_product_attention behind diffusers' AttnProcessor2_0 (self- and cross-attention of
// every BasicTransformerBlock).  No mask, no dropout, scale = d^-1/2 (given).
//
// One CTA = 128 query rows of one head, KV blocks of 64 keys.  Warp roles: warp 0 TMA producer, warp 1 MMA issuer,
// warps 2-5 softmax + correction + epilogue (one query row per thread: the TMEM lane IS the row, so row max / row
// sum need no shuffles).
//   S_j = Q K_j^T          tcgen05.mma 128x64x16, K-major operands, into 64 TMEM columns
//   P_j = exp2(S_j - m)    registers -> bf16 -> shared memory in the K-major SWIZZLE_128B layout the MMA expects
//   O  += P_j V_j          tcgen05.mma 128 x round16(d) x 16, V consumed MN-major straight from its [keys, d]
//                          row layout (no transpose anywhere), O accumulates in TMEM.
// The kernel is bound by the MUFU pipe (8192 ex2 per 128x64 block = 512 SM cycles vs 192 MMA cycles at d=40), so the
// design is about keeping MUFU issue slots full: ONE pass over S per block (lazy running max, below) and a footprint
// small enough (64 KB smem, 128 TMEM columns, <= 112 registers at d <= 64) for THREE CTAs per SM — while one CTA
// waits for its PV_j / S_{j+1} round trip through the tensor core, the softmax warps of the other two own the SM's
// four MUFU units.
// Head dims that are not a multiple of 64 (SD1.5: 40, 80, 160) are handled by TMA itself: the box is 64 wide
// but the tensor-map extent is d, so columns >= d arrive as zeros and never touch HBM.
#include <algorithm>
#include <new>
#include <type_traits>

#include "common.cuh"
#include "internal.h"

namespace b200sd {
#ifdef B200SD_ATT_TRACE
// Tuning aid (not part of the ABI, compiled only with -DB200SD_ATT_TRACE): clock stamps of CTA (0,0,0)'s roles.
__device__ unsigned long long g_att_trace[3 * 1024];
#define ATT_TR(role, id)                                                                                      \
  do {                                                                                                        \
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && tr_n < 1024)                                  \
      g_att_trace[(role) * 1024 + tr_n++] = (static_cast<unsigned long long>(id) << 48) | (clock64() & 0xFFFFFFFFFFFFull); \
  } while (0)
extern "C" __attribute__((visibility("default"))) int b200sd_debug_att_trace(unsigned long long* dst) {
  return static_cast<int>(cudaMemcpyFromSymbol(dst, g_att_trace, sizeof(g_att_trace)));
}
#else
#define ATT_TR(role, id) do {} while (0)
#endif
namespace {

constexpr int ATT_THREADS = 192;
constexpr int BLK = 128;           // query rows per CTA
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
};

template <int DPAD>
struct AttnCfg {
  static constexpr int DB = DPAD / 64;
  static constexpr int OCC = DPAD == 64 ? 4 : (DPAD == 128 ? 2 : 1);  // CTAs per SM (see the header comment)
  static constexpr int NSTK = 1;  // K stages: K_j is free as soon as S_j has completed, a whole softmax before K_{j+1} is needed
  static constexpr int NSTV = DPAD == 64 ? 1 : 2;  // V stages: V_j is free when PV_j has completed, a softmax before V_{j+1} is needed
  static constexpr int TMEM_COLS = DPAD == 64 ? 128 : 256;            // 64 columns of S, then O (<= DPAD columns)
  static constexpr int Q_BYTES = DB * TILE16K;
  static constexpr int KV_BYTES = DB * TILE8K;                        // one K (or V) stage
  static constexpr int P_BYTES = TILE16K;                             // 128 rows x 64 keys bf16
  static constexpr int SMEM = Q_BYTES + (NSTK + NSTV) * KV_BYTES + P_BYTES + 1024 + 256;
};
static_assert(AttnCfg<192>::SMEM <= 232448 && 4 * (AttnCfg<64>::SMEM + 1024) <= 233472 &&
              2 * (AttnCfg<128>::SMEM + 1024) <= 233472, "shared memory budget");

__device__ __forceinline__ float ex2_approx(float x) {
#ifdef B200SD_ATT_NOEXP  // diagnostic build only: how fast is everything but the MUFU work?
  return x * 0.001f + 0.01f;
#else
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
#endif
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

template <int DPAD>
__global__ void __launch_bounds__(ATT_THREADS, AttnCfg<DPAD>::OCC) attention_kernel(const __grid_constant__ AttnParams p) {
  using Cfg = AttnCfg<DPAD>;
  constexpr int DB = Cfg::DB;
  constexpr int NSTK = Cfg::NSTK, NSTV = Cfg::NSTV;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + Cfg::Q_BYTES;                 // [NSTK][KV_BYTES]
  uint8_t* sV = sK + NSTK * Cfg::KV_BYTES;         // [NSTV][KV_BYTES]
  uint8_t* sP = sV + NSTV * Cfg::KV_BYTES;         // [TILE16K]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + Cfg::P_BYTES);
  uint64_t* q_full = bars;             // 1
  uint64_t* k_full = bars + 1;         // NSTK
  uint64_t* k_empty = k_full + NSTK;   // NSTK
  uint64_t* v_full = k_empty + NSTK;   // NSTV
  uint64_t* v_empty = v_full + NSTV;   // NSTV
  uint64_t* s_full = v_empty + NSTV;   // 1
  uint64_t* p_full = s_full + 1;       // 1
  uint64_t* pv_done = p_full + 1;      // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_done + 1);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * BLK;
  const int head = blockIdx.y;
  const int batch = blockIdx.z;
  const int nkv = p.kv_blocks;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < NSTK; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); }
    for (int i = 0; i < NSTV; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
    mbar_init(s_full, 1);
    mbar_init(p_full, 128);
    mbar_init(pv_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish(); }
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&p.tmQ); tma_prefetch_desc(&p.tmK); tma_prefetch_desc(&p.tmV); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // the prologue above overlapped the previous kernel's tail
  const uint32_t tmem_S = tmem_base;       // 64 columns
  const uint32_t tmem_O = tmem_base + KB;  // round16(d) columns

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      mbar_arrive_expect_tx(q_full, Cfg::Q_BYTES);
      for (int db = 0; db < DB; ++db) tma_load_4d(sQ + db * TILE16K, &p.tmQ, q_full, db * 64, head, q0, batch);
      int tr_n = 0; (void)tr_n;
      for (int j = 0; j < nkv; ++j) {
        const int sk = j % NSTK, sv = j % NSTV;
        mbar_wait(&k_empty[sk], ((j / NSTK) & 1) ^ 1);
        ATT_TR(0, 9);
        mbar_arrive_expect_tx(&k_full[sk], Cfg::KV_BYTES);
        for (int db = 0; db < DB; ++db)
          tma_load_4d(sK + sk * Cfg::KV_BYTES + db * TILE8K, &p.tmK, &k_full[sk], db * 64, head, j * KB, batch);
        mbar_wait(&v_empty[sv], ((j / NSTV) & 1) ^ 1);
        ATT_TR(0, 10);
        mbar_arrive_expect_tx(&v_full[sv], Cfg::KV_BYTES);
        for (int db = 0; db < DB; ++db)
          tma_load_4d(sV + sv * Cfg::KV_BYTES + db * TILE8K, &p.tmV, &v_full[sv], db * 64, head, j * KB, batch);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // This warp shares its SM sub-partition with the softmax warps of up to three CTAs, so every instruction between
    // "P_j is ready" and "S_{j+1} is issued" sits on the critical path of the whole CTA at a fraction of the issue
    // rate: all smem descriptors are built once, up front, and the operand barriers (V_j, K_{j+1}: both land long
    // before P_j) are waited for BEFORE the P barrier — what remains after it is 7 tcgen05.mma and 4 commits.
    constexpr int KS_MAX = DPAD / 16;
    const uint32_t idesc_s = umma_idesc_bf16(BLK, KB, 0);
    const uint32_t idesc_o = umma_idesc_bf16(BLK, p.n_pv, 1);
    uint64_t dq[KS_MAX], dk[KS_MAX], dp[KB / 16], dv0[KB / 16], dv1[KB / 16];
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
        dv0[ks] = umma_smem_desc_sw128(v_addr + ks * 2048, TILE8K, 1024);
        dv1[ks] = umma_smem_desc_sw128(v_addr + (NSTV - 1) * Cfg::KV_BYTES + ks * 2048, TILE8K, 1024);
      }
    }
    static_assert(NSTK == 1 && NSTV <= 2, "descriptor tables assume one K stage and at most two V stages");
    int tr_n = 0; (void)tr_n;
    auto issue_S = [&]() {  // S = Q K^T (lane 0); the K stage is released when it completes
#pragma unroll
      for (int ks = 0; ks < KS_MAX; ++ks)
        if (ks < p.ksteps_qk) umma_bf16_ss(tmem_S, dq[ks], dk[ks], idesc_s, ks != 0 ? 1u : 0u);
      umma_commit(s_full);
      umma_commit(&k_empty[0]);
    };
    mbar_wait(q_full, 0);
    mbar_wait(&k_full[0], 0);
    tc_fence_after();
    if (elect_one()) issue_S();
    __syncwarp();
    for (int j = 0; j < nkv; ++j) {
      const int st = j % NSTV;
      mbar_wait(&v_full[st], (j / NSTV) & 1);             // V_j: loaded a block ago
      if (j + 1 < nkv) mbar_wait(&k_full[0], (j + 1) & 1);  // K_{j+1}: loaded while the softmax of block j ran
      mbar_wait(p_full, j & 1);                            // P_j written and S_j fully read
      tc_fence_after();
      if (elect_one()) {
        ATT_TR(1, 4);
        if (j + 1 < nkv) issue_S();  // S_{j+1} first: it is what the softmax warps are waiting for
        ATT_TR(1, 6);
#pragma unroll
        for (int ks = 0; ks < KB / 16; ++ks)
          umma_bf16_ss(tmem_O, dp[ks], st ? dv1[ks] : dv0[ks], idesc_o, (j | ks) != 0 ? 1u : 0u);
        umma_commit(pv_done);
        umma_commit(&v_empty[st]);
        ATT_TR(1, 8);
      }
      __syncwarp();
    }
  } else {
    // ===================== softmax / correction / epilogue: thread <-> query row =====================
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t lane_base = static_cast<uint32_t>(q * 32) << 16;
    const uint32_t s_addr = tmem_S + lane_base;
    const uint32_t o_addr = tmem_O + lane_base;
    const uint32_t prow = smem_u32(sP) + r * 128;
    const int rsw = r & 7;
    float m_run = -INFINITY;  // softmax reference point of this row, scaled log2 domain
    float l_run = 0.f;
    int tr_n = 0; (void)tr_n;
    // Lazy running max (single pass over S on every block but the first): probabilities are taken relative to the
    // max of the keys BEFORE this block, and the block's own max is tracked in the same pass.  Softmax is invariant
    // to the reference point; bf16 P and the fp32 row sum keep their relative precision at any magnitude, so the
    // reference only has to move (and O be rescaled) when a row's scores outgrow it by more than 2^LAZY_TAU — then
    // the warp redoes its block with the exact max (S_j is still in TMEM).
    constexpr float LAZY_TAU = 8.0f;
    for (int j = 0; j < nkv; ++j) {
      const int valid = min(KB, p.Nkv - j * KB);  // keys of this block that exist
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      bool p_busy = j > 0;  // PV_{j-1} is issued right AFTER S_j: it may still be reading P when S_j lands
      if (warp == 2 && lane == 0) ATT_TR(2, 1);
      // probabilities -> bf16 -> swizzled smem (K-major, 128-byte rows, 16-byte chunk ^= row % 8); returns the row sum
      auto exp_pass = [&](auto masked_tag, const float m_ref, float& mx_out) {
        constexpr bool MASKED = decltype(masked_tag)::value;
        float rsum = 0.f, mx = -INFINITY;
#pragma unroll
        for (int c = 0; c < KB / 32; ++c) {
          uint32_t u[32];
          tmem_ld32(s_addr + c * 32, u);
          tmem_ld_wait();
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float s0 = __uint_as_float(u[2 * i]), s1 = __uint_as_float(u[2 * i + 1]);
            float p0 = ex2_approx(fmaf(s0, p.scale_log2, -m_ref));
            float p1 = ex2_approx(fmaf(s1, p.scale_log2, -m_ref));
            if (MASKED) {
              if (c * 32 + 2 * i >= valid) p0 = 0.f; else mx = fmaxf(mx, s0);
              if (c * 32 + 2 * i + 1 >= valid) p1 = 0.f; else mx = fmaxf(mx, s1);
            } else {
              mx = fmaxf(mx, fmaxf(s0, s1));
            }
            rsum += p0 + p1;
            pk[i] = pack_bf16x2(p0, p1);
          }
          if (p_busy) {  // (complete long before the first chunk's exponentials are)
            mbar_wait(pv_done, (j - 1) & 1);
            p_busy = false;
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
#ifdef B200SD_ATT_NOSTS  // diagnostic build only
            if (pk[4 * i] == 0x12345678u)
#endif
            sts128(prow + (((c * 4 + i) ^ rsw) << 4), pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
          }
        }
        mx_out = mx * p.scale_log2;
        return rsum;
      };
      // The body is instantiated twice: the mask-free version runs on every full block (no per-element predicates).
      auto block_body = [&](auto masked_tag) {
        constexpr bool MASKED = decltype(masked_tag)::value;
        if (j == 0) {  // very first block: exact max (nothing to rescale yet)
          float mx = -INFINITY;
#pragma unroll
          for (int c = 0; c < KB / 32; ++c) {
            uint32_t u[32];
            tmem_ld32(s_addr + c * 32, u);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (!MASKED || c * 32 + i < valid) mx = fmaxf(mx, __uint_as_float(u[i]));
          }
          m_run = mx * p.scale_log2;
        }
        float m_blk;
        float rsum = exp_pass(masked_tag, m_run, m_blk);
        const bool outgrown = m_blk > m_run + LAZY_TAU;
        if (__any_sync(0xffffffffu, outgrown)) {  // rare; warp-uniform because tcgen05.ld/st are .sync.aligned
          float alpha = 1.0f;
          if (outgrown) {  // this row moves its reference to the exact max; the other rows just repeat
            alpha = ex2_approx(m_run - m_blk);
            m_run = m_blk;
          }
          rsum = exp_pass(masked_tag, m_run, m_blk);
          l_run *= alpha;
          // O *= alpha once PV_{j-1} (the latest PV issued) has landed
          mbar_wait(pv_done, (j - 1) & 1);
          tc_fence_after();
#pragma unroll 1
          for (int c = 0; c < p.n_pv / 16; ++c) {
            uint32_t u[16];
            tmem_ld16(o_addr + c * 16, u);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) u[i] = __float_as_uint(__uint_as_float(u[i]) * alpha);
            tmem_st16(o_addr + c * 16, u);
          }
          tmem_st_wait();
        }
        l_run += rsum;
      };
      if (valid == KB) block_body(std::false_type{});
      else block_body(std::true_type{});
      if (warp == 2 && lane == 0) ATT_TR(2, 2);
      fence_proxy_async_smem();  // P (generic-proxy stores) -> visible to the MMA's async-proxy reads
      tc_fence_before();
      mbar_arrive(p_full);
      if (warp == 2 && lane == 0) ATT_TR(2, 3);
    }
    // ---- epilogue: O / l -> bf16 -> global ----
    mbar_wait(pv_done, (nkv - 1) & 1);
    tc_fence_after();
    const float inv_l = 1.0f / l_run;
    const int qrow = q0 + r;
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

struct AttnPlan {
  AttnParams kp;
  int dpad;
  dim3 grid;
};

template <int DPAD>
int prepare_attention_t() {  // plan time, outside any stream capture
  static bool attr_set = false;
  if (!attr_set) {
    B200SD_CUDA(cudaFuncSetAttribute(attention_kernel<DPAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnCfg<DPAD>::SMEM));
    attr_set = true;
  }
  return 0;
}
int prepare_attention(int dpad) {
  switch (dpad) {
    case 64: return prepare_attention_t<64>();
    case 128: return prepare_attention_t<128>();
    case 192: return prepare_attention_t<192>();
  }
  return fail(B200SD_ERR_UNSUPPORTED, "attention: padded head dim %d unsupported", dpad);
}

template <int DPAD>
int launch_attention_t(const AttnPlan& pl, cudaStream_t s) {
  B200SD_CUDA(launch_kernel(attention_kernel<DPAD>, pl.grid, dim3(ATT_THREADS), AttnCfg<DPAD>::SMEM, s, 1, pl.kp));
  return 0;
}
int launch_attention(const void* plan, cudaStream_t s) {
  const AttnPlan& pl = *static_cast<const AttnPlan*>(plan);
  switch (pl.dpad) {
    case 64: return launch_attention_t<64>(pl, s);
    case 128: return launch_attention_t<128>(pl, s);
    case 192: return launch_attention_t<192>(pl, s);
  }
  return fail(B200SD_ERR_UNSUPPORTED, "attention: padded head dim %d unsupported", pl.dpad);
}

}  // namespace

int plan_attention(const b200sd_attention& a, PlannedOp* out) {
  B200SD_CHECK_ARG(a.q && a.k && a.v && a.out, "attention: NULL pointer");
  B200SD_CHECK_ARG(a.B >= 1 && a.heads >= 1 && a.Nq >= 1 && a.Nkv >= 1, "attention: bad shape");
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
  pl->dpad = (a.d + 63) / 64 * 64;
  pl->grid = dim3((a.Nq + BLK - 1) / BLK, a.heads, a.B);
  rc = prepare_attention(pl->dpad);
  if (rc) { delete pl; return rc; }
  out->kind = B200SD_OP_ATTENTION;
  out->plan = pl;
  out->launch = launch_attention;
  out->destroy = destroy_plan<AttnPlan>;
  out->launches = 1;
  return 0;
}

}  // namespace b200sd
