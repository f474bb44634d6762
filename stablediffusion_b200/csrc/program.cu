// C ABI of libb200sd.so: op programs (validate + plan once, replay on a stream / as a CUDA graph).
#include <mutex>
#include <new>

#include <stdlib.h>

#include "internal.h"

namespace b200sd {

std::string& last_error_ref() {
  static thread_local std::string msg;
  return msg;
}

int fail(int status, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  last_error_ref() = buf;
  return status;
}

bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("B200SD_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}

static int g_batch_invariant = -1;  // -1: not decided yet (environment read at first use)
bool batch_invariant() {
  if (g_batch_invariant < 0) {
    const char* e = getenv("B200SD_BATCH_INVARIANT");
    g_batch_invariant = (e && e[0] == '1') ? 1 : 0;
  }
  return g_batch_invariant != 0;
}
int set_batch_invariant(int on) {
  const int prev = batch_invariant() ? 1 : 0;
  g_batch_invariant = on ? 1 : 0;
  return prev;
}

int sm_count() {
  static int cached = -1;
  if (cached > 0) return cached;
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) == cudaSuccess &&
      cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) {
    cached = n;
    return n;
  }
  (void)cudaGetLastError();
  return 148;  // B200; used for planning arithmetic when no device is visible
}

// cuTensorMapEncodeTiled is fetched through the runtime so the library has no link-time libcuda dependency
// (it must load on a machine without a driver for the symbol / layout checks).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
    else
      (void)cudaGetLastError();
  });
  return fn;
}

int encode_tensor_map_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                           const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes) {
  EncodeTiledFn fn = get_encode_fn();
  CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  if (!fn) return fail(B200SD_ERR_NO_DEVICE, "cuTensorMapEncodeTiled unavailable (no CUDA driver): there is no CPU fallback");
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (box[i] < 1 || box[i] > 256) return fail(B200SD_ERR_INVALID, "tensor map: box[%d]=%u out of range", i, box[i]);
  }
  for (int i = 0; i + 1 < rank; ++i) {
    gstr[i] = strides_bytes[i];
    if (strides_bytes[i] % 16 != 0) return fail(B200SD_ERR_INVALID, "tensor map: stride[%d]=%llu not a multiple of 16 bytes", i, (unsigned long long)strides_bytes[i]);
  }
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, static_cast<cuuint32_t>(rank), const_cast<void*>(base), gdim, gstr,
                  bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    std::string d;
    for (int i = 0; i < rank; ++i) d += " " + std::to_string(dims[i]) + "/" + std::to_string(box[i]);
    return fail(B200SD_ERR_CUDA, "cuTensorMapEncodeTiled failed (CUresult %d) rank=%d dims/box:%s", (int)r, rank, d.c_str());
  }
  return 0;
}

}  // namespace b200sd

using namespace b200sd;

struct b200sd_program {
  int device = 0;
  std::vector<PlannedOp> ops;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t graph_exec = nullptr;
  std::vector<cudaEvent_t> events;
  ~b200sd_program() {
    for (auto& op : ops)
      if (op.plan && op.destroy) op.destroy(op.plan);
    if (graph_exec) cudaGraphExecDestroy(graph_exec);
    if (graph) cudaGraphDestroy(graph);
    for (auto e : events) cudaEventDestroy(e);
  }
};

extern "C" {

int b200sd_abi_version(void) { return B200SD_ABI_VERSION; }

const char* b200sd_last_error(void) { return last_error_ref().c_str(); }

size_t b200sd_sizeof_op(void) { return sizeof(b200sd_op); }

size_t b200sd_sizeof_struct(int32_t kind) {
  switch (kind) {
    case B200SD_OP_GEMM: return sizeof(b200sd_gemm);
    case B200SD_OP_GROUPNORM: return sizeof(b200sd_groupnorm);
    case B200SD_OP_LAYERNORM: return sizeof(b200sd_layernorm);
    case B200SD_OP_ATTENTION: return sizeof(b200sd_attention);
    case B200SD_OP_TIMESTEP_EMBED: return sizeof(b200sd_timestep_embed);
    case B200SD_OP_PREP_LATENTS: return sizeof(b200sd_prep_latents);
    case B200SD_OP_UPSAMPLE2X: return sizeof(b200sd_upsample2x);
    case B200SD_OP_PHASE_SPLIT: return sizeof(b200sd_phase_split);
    case B200SD_OP_SOFTMAX_ROWS: return sizeof(b200sd_softmax_rows);
    case B200SD_OP_CFG_STEP: return sizeof(b200sd_cfg_step);
    case B200SD_OP_AXPBY: return sizeof(b200sd_axpby);
    case B200SD_OP_IMAGE_OUT: return sizeof(b200sd_image_out);
    case B200SD_OP_STEP_ADVANCE: return sizeof(b200sd_step_advance);
    case B200SD_OP_IMAGE_U8: return sizeof(b200sd_image_u8);
    case B200SD_OP_EMBED: return sizeof(b200sd_embed);
    case B200SD_OP_GAUSS_SAMPLE: return sizeof(b200sd_gauss_sample);
    case B200SD_OP_PAIR_EXCHANGE: return sizeof(b200sd_pair_exchange);
    case 100: return sizeof(b200sd_gemm_seg);
  }
  return 0;
}

int b200sd_device_check(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) {
    (void)cudaGetLastError();
    return fail(B200SD_ERR_NO_DEVICE, "no CUDA device visible: libb200sd has no CPU fallback");
  }
  if (device < 0 || device >= n) return fail(B200SD_ERR_INVALID, "device %d out of range (%d devices)", device, n);
  int major = 0, minor = 0;
  B200SD_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  B200SD_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
  if (major != 10) return fail(B200SD_ERR_NO_DEVICE, "device %d is sm_%d%d; this library only contains sm_100a code", device, major, minor);
  return 0;
}

int b200sd_set_batch_invariant(int on) { return set_batch_invariant(on); }
int b200sd_get_batch_invariant(void) { return batch_invariant() ? 1 : 0; }

int b200sd_groupnorm_chunks(int32_t NB, int32_t HW, int32_t C) { return groupnorm_chunks(NB, HW, C); }

int64_t b200sd_gemm_workspace_bytes(const b200sd_gemm* g) {
  if (!g) return fail(B200SD_ERR_INVALID, "NULL gemm");
  GemmTiling t;
  int rc = choose_gemm_tiling(*g, &t);
  if (rc) return rc;
  if (t.splits <= 1) return 0;
  return static_cast<int64_t>(t.splits) * t.num_m_tiles * t.num_n_tiles * 128 * t.block_n * 4;
}

int b200sd_program_create(b200sd_program** out, int device) {
  if (!out) return fail(B200SD_ERR_INVALID, "NULL out");
  int rc = b200sd_device_check(device);
  if (rc) return rc;
  B200SD_CUDA(cudaSetDevice(device));
  b200sd_program* p = new (std::nothrow) b200sd_program();
  if (!p) return fail(B200SD_ERR_CUDA, "out of host memory");
  p->device = device;
  *out = p;
  return 0;
}

int b200sd_program_add(b200sd_program* p, const b200sd_op* op) {
  if (!p || !op) return fail(B200SD_ERR_INVALID, "NULL program/op");
  if (p->graph_exec) return fail(B200SD_ERR_INVALID, "program already captured into a graph");
  B200SD_CUDA(cudaSetDevice(p->device));
  PlannedOp po;
  int rc = 0;
  switch (op->kind) {
    case B200SD_OP_GEMM: rc = plan_gemm(op->u.gemm, &po); break;
    case B200SD_OP_GROUPNORM: rc = plan_groupnorm(op->u.groupnorm, &po); break;
    case B200SD_OP_LAYERNORM: rc = plan_layernorm(op->u.layernorm, &po); break;
    case B200SD_OP_ATTENTION: rc = plan_attention(op->u.attention, &po); break;
    case B200SD_OP_TIMESTEP_EMBED: rc = plan_timestep_embed(op->u.timestep_embed, &po); break;
    case B200SD_OP_PREP_LATENTS: rc = plan_prep_latents(op->u.prep_latents, &po); break;
    case B200SD_OP_UPSAMPLE2X: rc = plan_upsample2x(op->u.upsample2x, &po); break;
    case B200SD_OP_PHASE_SPLIT: rc = plan_phase_split(op->u.phase_split, &po); break;
    case B200SD_OP_SOFTMAX_ROWS: rc = plan_softmax_rows(op->u.softmax_rows, &po); break;
    case B200SD_OP_CFG_STEP: rc = plan_cfg_step(op->u.cfg_step, &po); break;
    case B200SD_OP_AXPBY: rc = plan_axpby(op->u.axpby, &po); break;
    case B200SD_OP_IMAGE_OUT: rc = plan_image_out(op->u.image_out, &po); break;
    case B200SD_OP_STEP_ADVANCE: rc = plan_step_advance(op->u.step_advance, &po); break;
    case B200SD_OP_IMAGE_U8: rc = plan_image_u8(op->u.image_u8, &po); break;
    case B200SD_OP_EMBED: rc = plan_embed(op->u.embed, &po); break;
    case B200SD_OP_GAUSS_SAMPLE: rc = plan_gauss_sample(op->u.gauss_sample, &po); break;
    case B200SD_OP_PAIR_EXCHANGE: rc = plan_pair_exchange(op->u.pair_exchange, &po); break;
    default: return fail(B200SD_ERR_INVALID, "unknown op kind %d", op->kind);
  }
  if (rc) return rc;
  po.tag = op->tag;
  p->ops.push_back(po);
  return static_cast<int>(p->ops.size()) - 1;
}

int b200sd_program_num_ops(const b200sd_program* p) { return p ? static_cast<int>(p->ops.size()) : 0; }

int b200sd_program_num_launches(const b200sd_program* p) {
  if (!p) return 0;
  int n = 0;
  for (const auto& op : p->ops) n += op.launches;
  return n;
}

int b200sd_program_run_range(b200sd_program* p, int first, int last, void* cuda_stream) {
  if (!p) return fail(B200SD_ERR_INVALID, "NULL program");
  if (first < 0 || last > static_cast<int>(p->ops.size()) || first > last)
    return fail(B200SD_ERR_INVALID, "op range [%d,%d) invalid for %zu ops", first, last, p->ops.size());
  cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
  for (int i = first; i < last; ++i) {
    int rc = p->ops[i].launch(p->ops[i].plan, s);
    if (rc) return rc;
  }
  return 0;
}

int b200sd_program_run(b200sd_program* p, void* cuda_stream) {
  if (!p) return fail(B200SD_ERR_INVALID, "NULL program");
  return b200sd_program_run_range(p, 0, static_cast<int>(p->ops.size()), cuda_stream);
}

int b200sd_program_run_timed(b200sd_program* p, void* cuda_stream, float* ms, int n) {
  if (!p || !ms) return fail(B200SD_ERR_INVALID, "NULL argument");
  const int nops = static_cast<int>(p->ops.size());
  if (n < nops) return fail(B200SD_ERR_INVALID, "ms buffer holds %d entries, program has %d ops", n, nops);
  cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
  while (static_cast<int>(p->events.size()) < nops + 1) {
    cudaEvent_t e;
    B200SD_CUDA(cudaEventCreate(&e));
    p->events.push_back(e);
  }
  B200SD_CUDA(cudaEventRecord(p->events[0], s));
  for (int i = 0; i < nops; ++i) {
    int rc = p->ops[i].launch(p->ops[i].plan, s);
    if (rc) return rc;
    B200SD_CUDA(cudaEventRecord(p->events[i + 1], s));
  }
  B200SD_CUDA(cudaStreamSynchronize(s));
  for (int i = 0; i < nops; ++i) B200SD_CUDA(cudaEventElapsedTime(&ms[i], p->events[i], p->events[i + 1]));
  return 0;
}

int b200sd_program_graph_capture(b200sd_program* p, void* cuda_stream) {
  if (!p) return fail(B200SD_ERR_INVALID, "NULL program");
  if (p->graph_exec) return 0;
  cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
  // capture only records: nothing executes here (kernel attributes were set at plan time)
  B200SD_CUDA(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
  int rc = b200sd_program_run(p, s);
  cudaGraph_t g = nullptr;
  cudaError_t e = cudaStreamEndCapture(s, &g);
  if (rc) {
    if (g) cudaGraphDestroy(g);
    return rc;
  }
  if (e != cudaSuccess) return fail(B200SD_ERR_CUDA, "cudaStreamEndCapture failed: %s", cudaGetErrorString(e));
  p->graph = g;
  B200SD_CUDA(cudaGraphInstantiate(&p->graph_exec, g, 0));
  return 0;
}

int b200sd_program_graph_launch(b200sd_program* p, void* cuda_stream) {
  if (!p || !p->graph_exec) return fail(B200SD_ERR_INVALID, "program has no captured graph");
  B200SD_CUDA(cudaGraphLaunch(p->graph_exec, static_cast<cudaStream_t>(cuda_stream)));
  return 0;
}

int b200sd_program_destroy(b200sd_program* p) {
  delete p;
  return 0;
}

}  // extern "C"
