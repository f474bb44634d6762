// Host-side internals shared by the translation units of libb200sd.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <utility>
#include <vector>

#include "../../include/b200sd.h"

namespace b200sd {

// ---- error plumbing (thread-local message; C ABI returns the status code) -----------------------------
std::string& last_error_ref();
int fail(int status, const char* fmt, ...);

#define B200SD_CHECK_ARG(cond, ...)                                  \
  do {                                                               \
    if (!(cond)) return ::b200sd::fail(B200SD_ERR_INVALID, __VA_ARGS__); \
  } while (0)

#define B200SD_CUDA(call)                                                                       \
  do {                                                                                          \
    cudaError_t e__ = (call);                                                                   \
    if (e__ != cudaSuccess)                                                                     \
      return ::b200sd::fail(B200SD_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                            __FILE__, __LINE__);                                                \
  } while (0)

// ---- TMA tensor-map encoding through the driver entry point (no link-time libcuda dependency) --------
// dims/strides innermost first; strides_bytes has rank-1 entries (dim 0 is contiguous).
int encode_tensor_map_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                           const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes = 128);

// ---- planned ops -------------------------------------------------------------------------------------
struct PlannedOp {
  int kind = 0;
  int tag = 0;
  int launches = 0;
  // type-erased plan owned by the op's translation unit
  void* plan = nullptr;
  int (*launch)(const void* plan, cudaStream_t s) = nullptr;
  void (*destroy)(void* plan) = nullptr;
};

int plan_gemm(const b200sd_gemm& g, PlannedOp* out);
int plan_attention(const b200sd_attention& a, PlannedOp* out);
int plan_groupnorm(const b200sd_groupnorm& g, PlannedOp* out);
int plan_layernorm(const b200sd_layernorm& g, PlannedOp* out);
int plan_timestep_embed(const b200sd_timestep_embed& g, PlannedOp* out);
int plan_prep_latents(const b200sd_prep_latents& g, PlannedOp* out);
int plan_upsample2x(const b200sd_upsample2x& g, PlannedOp* out);
int plan_phase_split(const b200sd_phase_split& g, PlannedOp* out);
int plan_softmax_rows(const b200sd_softmax_rows& g, PlannedOp* out);
int plan_cfg_step(const b200sd_cfg_step& g, PlannedOp* out);
int plan_axpby(const b200sd_axpby& g, PlannedOp* out);
int plan_image_out(const b200sd_image_out& g, PlannedOp* out);
int plan_step_advance(const b200sd_step_advance& g, PlannedOp* out);
int plan_image_u8(const b200sd_image_u8& g, PlannedOp* out);
int plan_embed(const b200sd_embed& g, PlannedOp* out);
int plan_gauss_sample(const b200sd_gauss_sample& g, PlannedOp* out);
int plan_pair_exchange(const b200sd_pair_exchange& g, PlannedOp* out);

// host-only helpers used by sizing entry points
struct GemmTiling {
  int block_n, splits, kb_total, kb_per_split;
  int num_m_tiles, num_n_tiles;
  int ncta, num_pm;  // CTAs per tile row block (1, or 2 = cta_group::2 pair) and ceil(num_m_tiles / ncta)
  int bw_log2, bh_log2, bn_log2, tiles_w, tiles_h, tiles_nb;
  int wres, wres_stages, units_per_nt;  // weight-resident mode (gemm.cu): A-only stages, work units per N tile
};
int choose_gemm_tiling(const b200sd_gemm& g, GemmTiling* t);
int groupnorm_chunks(int NB, int HW, int C);
int sm_count();  // of the current device; 148 when no device is visible (planning arithmetic only)
// Batch-invariant mode (b200sd_set_batch_invariant): reduction-order choices are made for a canonical batch of
// kCanonBatch images of the op's per-image shape, whatever the actual batch is.
bool batch_invariant();
constexpr int kCanonBatch = 8;

// Every kernel launch of the library: programmatic stream serialization (see pdl_wait() in common.cuh; B200SD_PDL=0 in
// the environment turns it off) and, for the CTA-pair GEMM, a cluster dimension.  Works under stream capture.
bool pdl_enabled();
template <class... KArgs, class... Args>
cudaError_t launch_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, int cluster_x,
                          Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (cluster_x > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster_x;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

template <class T>
void destroy_plan(void* p) {
  delete static_cast<T*>(p);
}

}  // namespace b200sd
