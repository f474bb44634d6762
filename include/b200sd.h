/*
 * b200sd.h — C ABI of libb200sd.so: the B200 (sm_100a) kernels behind the Stable Diffusion denoising
 * hot path (UNet forward per sampler step, CFG combine + scheduler update, VAE decode).
 *
 * What this boundary replaces.  The reference (GrafikXxxxxxxYyyyyyyyyyy/StableDiffusion) has no FFI of its own: its "operator API"
 * for this path is Python duck typing on three attributes of SDModelWrapper (models/stable_diffusion.py:40)
 * consumed by StableDiffusionUnifiedPipeline.__call__ (pipelines/sd_unified_pipeline.py:137):
 *
 *   model.base(sample, t, ctx, ...)            pipelines/sd_unified_pipeline.py:475-482   -> a program of ops below
 *   scheduler.scale_model_input / .step        pipelines/sd_unified_pipeline.py:472,489   -> PREP_LATENTS / CFG_STEP
 *   guidance combine                           pipelines/sd_unified_pipeline.py:484-486   -> CFG_STEP (fused)
 *   model.vae.decode(latents / sf)             pipelines/sd_unified_pipeline.py:511-523   -> a program of ops below
 *
 * The library executes *op programs*: the host (stablediffusion_b200/, Python, mirroring the reference's
 * objects) lowers a UNet / VAE / scheduler configuration to a flat list of fused ops with resolved device
 * pointers; the library validates each op, builds TMA tensor maps and launch plans once, and then replays
 * the whole list on a CUDA stream (optionally as one CUDA graph).  Plain pointers and sizes only; no
 * torch types; every entry point returns 0 on success or a negative b200sd_status and never throws.
 * There is no CPU fallback: without a CUDA device every compute entry point fails with
 * B200SD_ERR_NO_DEVICE.
 *
 * Layout conventions: activations are bf16, NHWC / token-major ([rows, channels], row pitch given in
 * elements); weights are bf16 [N_out, K] with K contiguous (3x3 conv weights are [C_out, kh, kw, C_in]);
 * bias / norm affine parameters are fp32.  Accumulation and all normalisation statistics are fp32.
 */
#ifndef B200SD_H_
#define B200SD_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200SD_API __attribute__((visibility("default")))
#else
#define B200SD_API
#endif

#define B200SD_ABI_VERSION 3
#define B200SD_MAX_SEG 3
#define B200SD_MAX_TAPS 9

typedef enum b200sd_status {
  B200SD_OK = 0,
  B200SD_ERR_INVALID = -1,   /* bad argument (maps to ValueError on the Python side) */
  B200SD_ERR_CUDA = -2,      /* CUDA runtime / driver error                           */
  B200SD_ERR_NO_DEVICE = -3, /* no sm_100 device: there is no fallback path           */
  B200SD_ERR_UNSUPPORTED = -4
} b200sd_status;

typedef enum b200sd_op_kind {
  B200SD_OP_GEMM = 1,          /* tcgen05 implicit GEMM: linear / 1x1 / 3x3 conv (+fused epilogue)        */
  B200SD_OP_GROUPNORM = 2,     /* GroupNorm (+SiLU), one or two channel-concatenated sources              */
  B200SD_OP_LAYERNORM = 3,
  B200SD_OP_ATTENTION = 4,     /* fused flash-style attention on tcgen05                                  */
  B200SD_OP_TIMESTEP_EMBED = 5,
  B200SD_OP_PREP_LATENTS = 6,  /* CFG duplicate + scale_model_input + NCHW->NHWC8 (pipeline :467-472)      */
  B200SD_OP_UPSAMPLE2X = 7,    /* nearest 2x, NHWC                                                        */
  B200SD_OP_PHASE_SPLIT = 8,   /* space-to-depth: [NB,H,W,C] -> [4,NB,H/2,W/2,C] for stride-2 convs        */
  B200SD_OP_SOFTMAX_ROWS = 9,
  B200SD_OP_CFG_STEP = 10,     /* eps = u + s(c-u); scheduler update (Euler / DDIM / DPM++2M) (:484-489)   */
  B200SD_OP_AXPBY = 11,        /* out = a*x + b*y  (add_noise, scale_model_input, latent un-scale)         */
  B200SD_OP_IMAGE_OUT = 12,    /* NHWC bf16 -> NCHW image tensor                                          */
  B200SD_OP_STEP_ADVANCE = 13, /* ++(*step_counter)                                                       */
  B200SD_OP_IMAGE_U8 = 14,     /* NCHW image in [-1,1] -> uint8 HWC (runpod-worker/handler_logic.py:21-29) */
  B200SD_OP_EMBED = 15,        /* token + position embedding lookup of the CLIP text encoders             */
  B200SD_OP_GAUSS_SAMPLE = 16, /* VAE posterior sample: mean + exp(0.5*clamp(logvar)) * noise (pipeline :1027)  */
  B200SD_OP_PAIR_EXCHANGE = 17 /* cond/uncond GPU-pair split: this rank's eps half -> the peer's window, in-graph    */
} b200sd_op_kind;

/* element types of host-visible tensors (latents, images) */
typedef enum b200sd_dtype { B200SD_BF16 = 0, B200SD_F32 = 1, B200SD_F16 = 2 } b200sd_dtype;

/* ---- GEMM ------------------------------------------------------------------------------------------ */
enum {
  B200SD_GEMM_BIAS = 1 << 0,      /* += bias[n]                                                          */
  B200SD_GEMM_BIAS_ROW = 1 << 1,  /* += bias[m] (row bias; used when operands are swapped)                */
  B200SD_GEMM_ROWVEC = 1 << 2,    /* += rowvec[group(m)][n]  (resnet time-embedding add)                  */
  B200SD_GEMM_RESIDUAL = 1 << 3,  /* += residual[m][n] (bf16)                                             */
  B200SD_GEMM_SILU = 1 << 4,      /* v = silu(v) after bias                                               */
  B200SD_GEMM_GEGLU = 1 << 5,     /* out[m][j] = h*gelu(g); weight rows pre-interleaved per N tile        */
  B200SD_GEMM_OUT_F32 = 1 << 6,   /* store fp32 instead of bf16                                           */
  B200SD_GEMM_QUICK_GELU = 1 << 7, /* v = v * sigmoid(1.702 v) after bias (CLIP-L MLP)                     */
  B200SD_GEMM_GELU = 1 << 8       /* v = gelu(v) (erf form) after bias (OpenCLIP bigG MLP)                 */
};

/* One K segment of an implicit GEMM.  The A operand of the whole GEMM is the K-concatenation of the
 * segments; segment i contributes taps*channels columns and pairs with weight columns starting at w. */
typedef struct b200sd_gemm_seg {
  const void* a;     /* bf16 activations. linear: [M, lda]; spatial: [phases, NB, Hs, Ws, lda] (first `channels` used) */
  const void* w;     /* bf16 weights for this segment: [N, taps, channels] with row pitch ldw              */
  int64_t lda;       /* row / pixel pitch of a, elements                                                   */
  int32_t channels;  /* C_i (multiple of 8)                                                                */
  int32_t taps;      /* 1 (linear, 1x1) .. 9 (3x3)                                                         */
  int32_t Hs, Ws;    /* spatial mode: source height/width per phase                                        */
  int32_t phases;    /* spatial mode: 1, or 4 after PHASE_SPLIT                                            */
  int8_t tap_phase[B200SD_MAX_TAPS];
  int8_t tap_dh[B200SD_MAX_TAPS]; /* source row = oh + tap_dh (zero outside [0,Hs))                         */
  int8_t tap_dw[B200SD_MAX_TAPS];
  int8_t pad_[1];
} b200sd_gemm_seg;

typedef struct b200sd_gemm {
  int32_t spatial; /* 0: rows are linear; 1: rows are pixels of an [NB,H,W] output grid                  */
  int32_t nseg;
  b200sd_gemm_seg seg[B200SD_MAX_SEG];
  int64_t ldw;     /* weight row pitch (total K), elements                                               */
  int32_t M, N;    /* output rows (linear mode; spatial: NB*H*W is used) and columns                      */
  int32_t NB, H, W;
  void* out;
  int64_t ldc;     /* output row / pixel pitch, elements                                                  */
  /* spatial scatter: pixel(n,oh,ow) -> (n*OH + oh*sy+oy)*OW + ow*sx+ox ; set OH=H,OW=W,sy=sx=1,oy=ox=0 for identity */
  int32_t OH, OW, sy, sx, oy, ox;
  const float* bias;
  const float* rowvec;
  int64_t ld_rowvec;
  int32_t rows_per_group; /* linear mode: group = m / rows_per_group; spatial mode: group = image index   */
  int32_t flags;
  const void* residual;
  int64_t ldr;
  float alpha;            /* accumulator scale applied before bias (1.0 default)                          */
  int32_t splits;         /* split-K factor; 0 = choose automatically                                     */
  void* workspace;        /* fp32 split-K partials (may be NULL when splits==1)                           */
  int64_t workspace_bytes;
  int32_t block_n;        /* 0 = auto; else one of 32, 64, 128, 160, 256                                  */
  int32_t cta_pair;       /* 0 = auto; 1 = one CTA per 128-row tile; 2 = CTA pair (cta_group::2) per 256 rows */
  int32_t rows_per_sample; /* linear mode: rows that belong to one image (0 = unknown).  Only read in the batch-invariant
                              mode (b200sd_set_batch_invariant), where split-K is chosen from the per-image shape      */
  int32_t pad2_;
} b200sd_gemm;

/* ---- GroupNorm (+SiLU) ---------------------------------------------------------------------------- */
typedef struct b200sd_groupnorm {
  const void* x1;   /* bf16 [NB, HW, ld1], first C1 channels used                                        */
  const void* x2;   /* optional second source concatenated along channels (skip connection)              */
  int64_t ld1, ld2;
  int32_t C1, C2;
  int32_t NB, HW, groups, silu;
  float eps;
  int32_t pad_;
  const float* gamma; /* [C1+C2] */
  const float* beta;
  void* out;          /* bf16 [NB, HW, C1+C2]                                                             */
  float* partials;    /* workspace: NB * chunks * groups * 2 floats, chunks = b200sd_groupnorm_chunks()   */
  int64_t partials_bytes;
} b200sd_groupnorm;

typedef struct b200sd_layernorm {
  const void* x; /* bf16 [rows, ldx] */
  void* out;     /* bf16 [rows, ldo] */
  int64_t ldx, ldo;
  int32_t rows, C;
  float eps;
  int32_t pad_;
  const float* gamma;
  const float* beta;
} b200sd_layernorm;

/* ---- fused attention ------------------------------------------------------------------------------- */
typedef struct b200sd_attention {
  const void* q; /* bf16 [B, Nq, ldq]; head h occupies columns [h*d, (h+1)*d)                             */
  const void* k; /* bf16 [B, Nkv, ldk]                                                                    */
  const void* v; /* bf16 [B, Nkv, ldv]                                                                    */
  void* out;     /* bf16 [B, Nq, ldo]                                                                     */
  int64_t ldq, ldk, ldv, ldo;
  int32_t B, heads, Nq, Nkv, d;
  float scale;   /* softmax(scale * q k^T)                                                                */
  int32_t causal; /* 1: key j is visible to query i only when j <= i (CLIP text encoders)                  */
  int32_t pad_;
} b200sd_attention;

/* ---- small bandwidth ops --------------------------------------------------------------------------- */
/* Per-step scalars live in a device table of rows of 16 floats indexed by a device-side step counter, so
 * a captured CUDA graph can be replayed for every sampler step without re-baking kernel arguments.      */
#define B200SD_STEP_ROW 16
enum { /* columns of one step-table row */
  B200SD_ST_TIMESTEP = 0,    /* value fed to the sinusoidal embedding (float32 image of the timestep)     */
  B200SD_ST_IN_SCALE = 1,    /* scale_model_input factor (Euler: 1/sqrt(sigma^2+1); else 1)               */
  B200SD_ST_K1 = 2,          /* x0 = (x - K1*eps) * K2                                                    */
  B200SD_ST_K2 = 3,
  B200SD_ST_CX = 4,          /* x' = CX*x + CE*eps + C0*x0 + C1*(x0 - x0_prev)                            */
  B200SD_ST_CE = 5,
  B200SD_ST_C0 = 6,
  B200SD_ST_C1 = 7,
  B200SD_ST_GUIDANCE = 8,    /* guidance scale (read when cfg_step.gs_from_table != 0)                     */
  B200SD_ST_CN = 9,          /* x' += CN * noise[step]            (ancestral / SDE samplers)                      */
  B200SD_ST_H1 = 10,         /* x' += H1*e[-1] + H2*e[-2] + H3*e[-3]   (linear multistep on eps: PNDM / PLMS)     */
  B200SD_ST_H2 = 11,
  B200SD_ST_H3 = 12,
  B200SD_ST_FLAGS = 13,      /* float image of B200SD_STF_* bits                                               */
  B200SD_ST_BA = 14,         /* inpaint blend: x' = (1-m)*(BA*image_latents + BS*noise) + m*x'   (pipeline :492-506)  */
  B200SD_ST_BS = 15
};
enum {
  B200SD_STF_PUSH_EPS = 1,   /* after the update: e[-3] <- e[-2] <- e[-1] <- eps                                */
  B200SD_STF_SAVE_X = 2,     /* remember the sample the step started from (PNDM `cur_sample`)                   */
  B200SD_STF_USE_SAVED_X = 4, /* update the remembered sample instead of the current one (PNDM counter == 1)    */
  /* UniPC (predictor-corrector, order <= 2).  With this bit the row is read as
   *   x0 = (x - K1*eps)*K2
   *   xc = CORRECT ? H1*x_saved + H2*x0_prev + H3*(eps_hist[0] - x0_prev) + CE*(x0 - x0_prev) : x      (UniC)
   *   x' = CX*xc + C0*x0 + C1*(x0 - x0_prev)                                                            (UniP)
   *   x_saved <- xc ; eps_hist[0] <- x0_prev ; x0_prev <- x0          (eps_hist[0] holds the x0 before last) */
  B200SD_STF_UNIPC = 8,
  B200SD_STF_UNIPC_CORRECT = 16
};

typedef struct b200sd_timestep_embed {
  const float* t;          /* explicit timesteps [rows] (t_stride 1) or [1] (t_stride 0); NULL -> step table */
  const float* step_table; /* [nsteps][16]                                                                */
  const int32_t* step_counter;
  int32_t t_stride;
  int32_t rows, dim, flip_sin_to_cos;
  float freq_shift, max_period;
  void* out; /* bf16 [rows, dim] */
} b200sd_timestep_embed;

typedef struct b200sd_prep_latents {
  const void* latents; /* [B, C, h, w] NCHW, dtype in_dtype                                              */
  void* out;           /* bf16 [rep*B, h*w, Cpad]                                                        */
  const float* step_table;
  const int32_t* step_counter; /* NULL -> scale = fixed_scale                                             */
  float fixed_scale;
  int32_t in_dtype, B, C, hw, Cpad, rep;
} b200sd_prep_latents;

typedef struct b200sd_upsample2x {
  const void* x; /* bf16 [NB, H, W, C] */
  void* out;     /* bf16 [NB, 2H, 2W, C] */
  int32_t NB, H, W, C;
} b200sd_upsample2x;

typedef struct b200sd_phase_split {
  const void* x; /* bf16 [NB, H, W, C], H and W even */
  void* out;     /* bf16 [4, NB, H/2, W/2, C]; phase = (row parity)*2 + (col parity) */
  int32_t NB, H, W, C;
} b200sd_phase_split;

typedef struct b200sd_softmax_rows {
  const void* x; /* bf16 [rows, ldx] */
  void* out;
  int64_t ldx, ldo;
  int32_t rows, cols;
  float scale;
  int32_t pad_;
} b200sd_softmax_rows;

typedef struct b200sd_cfg_step {
  const void* eps;  /* bf16 [(cfg?2:1)*B, hw, ld_eps] token-major; rows [0,B) uncond, [B,2B) cond          */
  int64_t ld_eps;
  const void* x;    /* latents [B, C, h, w] NCHW, dtype x_dtype                                           */
  void* x_out;      /* may alias x                                                                        */
  float* x0_prev;   /* fp32 [B, C, hw] history for 2nd-order multistep (NULL if unused)                    */
  void* next_in;    /* optional: bf16 [rep*B, hw, Cpad] UNet input for the NEXT step (already scaled)      */
  const float* step_table;
  const int32_t* step_counter;
  float guidance_scale;
  int32_t cfg;      /* 1: eps holds 2B rows and is combined; 0: eps is used as is                          */
  int32_t x_dtype, B, C, hw, Cpad, rep;
  int32_t nsteps;   /* rows in the table (next_in uses row step+1 if it exists)                            */
  int32_t gs_from_table; /* 1: take the guidance scale from the step table (column B200SD_ST_GUIDANCE)      */
  const float* noise;    /* fp32 [nsteps_of_noise, B, C, hw]: row `step` is this step's standard-normal draw (NULL: CN ignored) */
  float* eps_hist;       /* fp32 [3, B, C, hw] eps history for the H1..H3 terms (NULL if unused)            */
  float* x_saved;        /* fp32 [B, C, hw] for B200SD_STF_SAVE_X / USE_SAVED_X (NULL if unused)             */
  const float* blend_mask;  /* fp32 [B, hw], 1 = repaint; NULL = no inpaint blend                            */
  const float* blend_image; /* fp32 [B, C, hw] latents of the original image                                 */
  const float* blend_noise; /* fp32 [B, C, hw] the noise the trajectory started from                         */
} b200sd_cfg_step;

typedef struct b200sd_axpby {
  const void* x;
  const void* y; /* may be NULL */
  void* out;
  float a, b;
  int32_t x_dtype, y_dtype, out_dtype;
  int32_t pad_;
  int64_t n;
  const float* coef; /* optional DEVICE pointer to {a, b}: read at run time instead of the baked a, b, so one planned op
                        serves every sampler step (scheduler.scale_model_input / add_noise called the reference's way) */
} b200sd_axpby;

typedef struct b200sd_image_out {
  const void* x; /* bf16 [B, HW, ldx], first C channels */
  void* out;     /* [B, C, HW] dtype out_dtype */
  int64_t ldx;
  int32_t B, C, HW, out_dtype;
} b200sd_image_out;

/* The serving output stage, `convert_pt_to_numpy` (runpod-worker/handler_logic.py:21-29), as one kernel:
 *   out[b, pix, c] = uint8( clamp(x[b, c, pix] / 2 + 0.5, 0, 1) * 255 )      (truncation, like Tensor.to(torch.uint8))
 * Every intermediate is rounded to the element type of x, as torch's eager elementwise ops do, so the bytes are the
 * reference's bytes for the same image tensor. */
typedef struct b200sd_image_u8 {
  const void* x; /* [B, C, HW] dtype x_dtype */
  uint8_t* out;  /* [B, HW, C] */
  int32_t B, C, HW, x_dtype;
} b200sd_image_u8;

/* CLIPTextEmbeddings: out[r] = token_embedding[ids[r]] + position_embedding[r % L]  (pipeline :587-589 calls the
 * encoder; the module is transformers' CLIPTextEmbeddings). */
typedef struct b200sd_embed {
  const int32_t* ids; /* [rows] token ids, 0 <= id < vocab (checked on the device: an id out of range traps)   */
  const void* tok;    /* bf16 [vocab, D]                                                                      */
  const void* pos;    /* bf16 [L, D]                                                                          */
  void* out;          /* bf16 [rows, D]                                                                       */
  int32_t rows, L, D, vocab;
} b200sd_embed;

/* DiagonalGaussianDistribution.sample of the VAE encoder output (diffusers; called at pipeline :98-106, :1027):
 *   out[b,c,p] = scale * (mom[b,c,p] + exp(0.5 * clamp(mom[b,C+c,p], -30, 20)) * noise[b,c,p])      c < C
 * moments are NCHW [B, 2C, hw] (mean | logvar); arithmetic in fp32. */
typedef struct b200sd_gauss_sample {
  const void* moments; /* [B, 2C, hw] dtype m_dtype */
  const void* noise;   /* [B, C, hw] dtype n_dtype: standard-normal draws                                */
  void* out;           /* [B, C, hw] dtype out_dtype                                                       */
  float scale;         /* 1.0, or vae.config.scaling_factor when the caller fuses pipeline :1039           */
  int32_t m_dtype, n_dtype, out_dtype;
  int32_t B, C, hw, pad_;
} b200sd_gauss_sample;

/* The per-step exchange of the cond/uncond GPU-pair split (BASELINE config 5), inside the captured step graph: copy this
 * rank's eps half into the same slot of the PEER's window (peer-mapped memory, b200sd_peer_open), release the peer's
 * flag, then wait for the peer's half to land in the local window.  `sync_*` are int32[2] {flag, ack} inside the
 * windows, `state` is a LOCAL int32[4] {seq, err, arrival counter, -} (zero-initialised; err != 0 after a spin timed out). */
typedef struct b200sd_pair_exchange {
  const void* src;      /* local window: this rank's eps half (16-byte aligned)                                  */
  void* dst_remote;     /* the same slot of the peer's window                                                    */
  int64_t bytes;        /* multiple of 16                                                                        */
  const int32_t* sync_local;
  int32_t* sync_remote;
  int32_t* state;
  int32_t timeout_ms;   /* 0 = 20000                                                                             */
  int32_t pad_;
} b200sd_pair_exchange;

typedef struct b200sd_step_advance {
  int32_t* step_counter;
  /* pair split only (else NULL): after the step, acknowledge to the peer that its eps half has been consumed
   * (sync_remote[1] = seq) and advance the local sequence number (state[0] = seq) */
  int32_t* ack_remote;  /* &sync_remote[1] of the peer's window */
  int32_t* state;
} b200sd_step_advance;

typedef struct b200sd_op {
  int32_t kind;   /* b200sd_op_kind */
  int32_t tag;    /* free for the caller (shows up in timing reports) */
  union {
    b200sd_gemm gemm;
    b200sd_groupnorm groupnorm;
    b200sd_layernorm layernorm;
    b200sd_attention attention;
    b200sd_timestep_embed timestep_embed;
    b200sd_prep_latents prep_latents;
    b200sd_upsample2x upsample2x;
    b200sd_phase_split phase_split;
    b200sd_softmax_rows softmax_rows;
    b200sd_cfg_step cfg_step;
    b200sd_axpby axpby;
    b200sd_image_out image_out;
    b200sd_step_advance step_advance;
    b200sd_image_u8 image_u8;
    b200sd_embed embed;
    b200sd_gauss_sample gauss_sample;
    b200sd_pair_exchange pair_exchange;
  } u;
} b200sd_op;

typedef struct b200sd_program b200sd_program;

/* ---- entry points ---------------------------------------------------------------------------------- */
B200SD_API int b200sd_abi_version(void);
B200SD_API const char* b200sd_last_error(void);          /* thread-local message of the last failing call            */
B200SD_API size_t b200sd_sizeof_op(void);                /* layout self-check for foreign-language bindings          */
B200SD_API size_t b200sd_sizeof_struct(int32_t kind);    /* size of the payload struct of `kind`                     */
B200SD_API int b200sd_device_check(int device);          /* 0 iff `device` exists and is sm_100                      */

/* Batch-invariant mode (process-wide switch; default off, or 1 when B200SD_BATCH_INVARIANT=1 is in the environment at
 * first use).  When on, every choice that changes the ORDER of a floating-point reduction — the split-K factor of a GEMM,
 * the chunking of the GroupNorm statistics — is made from the per-image shape only (N, K, H*W, rows_per_sample), never
 * from the batch size, so image i of a batch-B run is bit-identical to image i computed in any other batch (the
 * data-parallel shard == full-batch guarantee of SURVEY section 4).  Ops planned before a change keep their plan. */
B200SD_API int b200sd_set_batch_invariant(int on);       /* returns the previous setting                              */
B200SD_API int b200sd_get_batch_invariant(void);

/* Peer windows: device buffers another PROCESS on the same node can map and store into over NVLink (cudaIpc* under the
 * hood; one process per GPU).  `peer_alloc` returns a zero-filled buffer and its 64-byte handle; the handle travels to the
 * peer by any host channel (torch.distributed here), which maps it with `peer_open`. */
#define B200SD_IPC_HANDLE_BYTES 64
B200SD_API int b200sd_peer_alloc(int device, size_t bytes, void** ptr, unsigned char* handle);
B200SD_API int b200sd_peer_open(int device, const unsigned char* handle, void** remote_ptr);
B200SD_API int b200sd_peer_close(void* remote_ptr);
B200SD_API int b200sd_peer_free(void* ptr);

/* GroupNorm (+SiLU) without any allocation, for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.
 * Same arithmetic and kernels as B200SD_OP_GROUPNORM (two-kernel form), launched directly on `cuda_stream`; the partial
 * sums live in a file-scope `static __device__` array of B200SD_GN_STATIC_SCRATCH_BYTES inside the library (held by every
 * GPU that loads it until the process exits), so `g->partials` must be NULL and `partials_bytes` is ignored.  The caller
 * still passes inputs and outputs.  Shapes that need more scratch (NB * chunks * groups * 8 bytes, chunks =
 * b200sd_groupnorm_chunks()) are rejected with B200SD_ERR_INVALID.  Calls no cudaMalloc / cudaFree / allocating routine.
 * The array is shared by all calls and keeps the last call's contents (nothing needs to start at zero: every slot read is
 * written first); calls that OVERLAP on the GPU race on it — keep them on one stream or order them with events. */
#define B200SD_GN_STATIC_SCRATCH_BYTES (1 << 20)
B200SD_API int b200sd_groupnorm_static_scratch(int device, const b200sd_groupnorm* g, void* cuda_stream);

/* workspace sizing helpers (pure host arithmetic; callable without a GPU) */
B200SD_API int b200sd_groupnorm_chunks(int32_t NB, int32_t HW, int32_t C);
B200SD_API int64_t b200sd_gemm_workspace_bytes(const b200sd_gemm* g); /* bytes needed for the split-K it would choose */

B200SD_API int b200sd_program_create(b200sd_program** out, int device);
B200SD_API int b200sd_program_add(b200sd_program* p, const b200sd_op* op); /* validate + plan; returns op index >= 0 */
B200SD_API int b200sd_program_num_ops(const b200sd_program* p);
B200SD_API int b200sd_program_num_launches(const b200sd_program* p);       /* kernel launches per run                  */
B200SD_API int b200sd_program_run(b200sd_program* p, void* cuda_stream);
B200SD_API int b200sd_program_run_range(b200sd_program* p, int first, int last, void* cuda_stream);
/* run with a CUDA event pair around every op; ms[i] = device time of op i (for profiling only) */
B200SD_API int b200sd_program_run_timed(b200sd_program* p, void* cuda_stream, float* ms, int n);
B200SD_API int b200sd_program_graph_capture(b200sd_program* p, void* cuda_stream); /* capture run() into a CUDA graph */
B200SD_API int b200sd_program_graph_launch(b200sd_program* p, void* cuda_stream);
B200SD_API int b200sd_program_destroy(b200sd_program* p);

#ifdef __cplusplus
}
#endif
#endif /* B200SD_H_ */
