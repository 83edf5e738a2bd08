/* C ABI of libb200ops.so — the sm_100a kernel library of llm_in_practise_b200.
 *
 * Every entry point takes raw device pointers plus a cudaStream_t (passed as void*), launches
 * asynchronously on that stream and returns 0 on success, a cudaError_t (< 1000) or a
 * B200_ERR_* code (>= 10001). Nothing here synchronises the device. Entry points that need
 * temporary global memory take it from the caller (`scratch` / `workspace` arguments).
 *
 * Each declaration cites the reference call site (path:line under /root/reference) whose
 * library kernel it replaces; the reference itself contains no native code.
 */
#ifndef B200_OPS_H_
#define B200_OPS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
  B200_DT_BF16 = 0,
  B200_DT_FP16 = 1,
  B200_DT_FP32 = 2, /* as GEMM input: multiplied on the tensor cores as TF32 */
};

#ifndef B200_STATUS_CODES
#define B200_STATUS_CODES
enum B200Status {
  B200_OK = 0,
  B200_ERR_BAD_SHAPE = 10001,   /* shape / size argument out of the supported domain          */
  B200_ERR_BAD_ALIGN = 10002,   /* pointer or stride not a multiple of 16 bytes               */
  B200_ERR_NO_DRIVER = 10003,   /* cuTensorMapEncodeTiled could not be resolved               */
  B200_ERR_TENSORMAP = 10004,   /* the driver rejected a TMA descriptor                       */
  B200_ERR_TOO_LARGE = 10005,   /* size beyond a kernel's stated maximum                      */
  B200_ERR_UNSUPPORTED = 10006  /* dtype / flag combination not implemented                   */
};
#endif

/* ------------------------------------------------------------------------------------------
 * K1 / K3: tcgen05 GEMM.  D[b] = alpha * A[b] x B[b / b_div]^T (+ A2 x B2^T) (+ bias) + beta * D[b]
 *
 *   A is logically [M, K], B is logically [N, K].  "K-major" means the K index is contiguous
 *   (an nn.Linear weight, an activation matrix); "MN-major" (a_mn / b_mn = 1) means the operand
 *   is stored transposed, [K, M] or [K, N] row-major, so dgrad (dY x W) and wgrad (dY^T x X)
 *   read their operands in place with no transpose copy (16-bit inputs only).
 *   *_rs is the stride in elements between consecutive rows of the stored matrix;
 *   *_bs0 / *_bs1 are batch strides (batch index b -> (b % nb0, b / nb0)); B uses b / b_div
 *   (grouped-query attention shares one K/V head among b_div query heads).
 *   A2 [M, K2], B2 [N, K2] (both K-major, K2 <= 64) add a low-rank term on the tensor cores as
 *   one more k-block: the LoRA update  y += (s * x A^T) B^T  of peft's lora.Linear.
 *   Replaces: cuBLAS behind every nn.Linear (LLM_Distributed_Trainning/PyTorch/ddp_basics/
 *   ddp_gpt_wikitext2.py:86-165) and peft LoRA (Fine-Tuning/qwen3-8b-lora-dist.py:118-125).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const void* A;
  const void* B;
  void* D;
  const void* bias; /* [N], dtype bias_dtype, may be NULL */
  const void* A2;   /* optional low-rank term, may be NULL */
  const void* B2;
  int32_t M, N, K, K2;
  int32_t batch, nb0, b_div, bias_dtype;
  int64_t a_rs, a_bs0, a_bs1;
  int64_t b_rs, b_bs0, b_bs1;
  int64_t d_rs, d_bs0, d_bs1;
  int64_t a2_rs, b2_rs;
  int32_t a_mn, b_mn, in_dtype, out_dtype;
  float alpha, beta;
  /* Optional stream-K scratch (zero-initialised once by the caller, then owned by one stream):
   * at least b200_gemm_workspace_bytes() bytes. With it, problems whose tile count does not fill
   * whole SM waves are split evenly along K across all SMs; without it (NULL) every tile is
   * computed by one CTA / CTA pair. */
  void* workspace;
  int64_t workspace_bytes;
  /* Optional per-row beta (float[M], batch 1 only): D[m, :] = alpha * acc + beta_rows[m] * D[m, :]. The running
   * rescale of the online-softmax lm_head + cross-entropy (flash-attention style accumulation over vocabulary chunks). */
  const float* beta_rows;
  /* Optional: all-reduce D over tensor-parallel ranks inside the same launch (see b200_gemm_allreduce below); b200_gemm only. */
  const struct b200_gemm_allreduce* ar;
  /* Optional grouped GEMM (K16, mixture-of-experts FFN — DeepSeekLike_spare_MoE_wikitext2.py:300-329): b_map is a DEVICE
   * array of `batch` ints; batch entry b multiplies with B matrix b_map[b] (B + b_map[b] * b_bs0, bias + b_map[b] *
   * bias_bs) and is skipped when b_map[b] < 0. b_count = number of B matrices. The caller sorts tokens by expert into
   * row tiles (one batch entry = one tile of M rows) on the device; nothing about the grouping is read by the host. */
  const int32_t* b_map;
  int32_t b_count;
  int64_t bias_bs;
} b200_gemm_args;

/* K-C5: row-parallel projection + all-reduce in ONE launch (y = sum over ranks of x_r W_r^T — o_proj / down_proj under
 * vLLM's --tensor-parallel-size, Quantization/README.md:175; Deployment/Ray/serve_deploy_examples/*.yaml). Every rank calls
 * b200_gemm with D = its copy of one symmetric bf16 buffer (same offset everywhere, batch 1, 16-byte aligned rows) and
 * this descriptor. Tiles are announced to their owner rank as they finish; CTAs that have run out of GEMM tiles sum the
 * owned tiles in place inside NVSwitch (multimem.ld_reduce / multimem.st on d_mc, fp32 accumulation) while other CTAs
 * are still multiplying (between two of their own tiles, too); the launch returns with the full sum in every rank's D. All ranks must launch it (bounded
 * waits trap otherwise). tile_flag_ptrs[r]: uint32 [max_tiles * 16] on rank r (zero-initialised once); sync_flag_ptrs[r]:
 * uint32 [channels][8] epoch flags on rank r; ticket: one local zero-initialised uint32; epoch: grows by >= 1 per call. */
typedef struct b200_gemm_allreduce {
  void* d_mc;                          /* multicast address of D */
  const void* const* d_ptrs;           /* [world] peer addresses of D (optional; used instead of d_mc when world == 2) */
  const void* const* tile_flag_ptrs;   /* [world] */
  const void* const* sync_flag_ptrs;   /* [world] */
  unsigned int* ticket;
  int32_t world, rank, channel;
  uint32_t epoch;
  int64_t max_tiles;
} b200_gemm_allreduce;

long long b200_gemm_workspace_bytes(void);

int b200_gemm(const b200_gemm_args* args, void* stream);

/* The same GEMM with the B operand expanded from 4-bit codes inside the kernel: NF4 (QLoRA base,
 * Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:101-106) or GPTQ/AWQ W4A16 weights
 * (Quantization/GPTQModel/quantize_qwen3_4b_gptq.py:16-21), optionally row-sharded over up to 8 buffers that may be
 * peer-GPU memory (ZeRO-3 / FSDP parameter shards; the all-gather happens as TMA reads inside the GEMM). W is
 * [rows, kw], packed [rows, kw / 2] uint8 (even element in the high nibble), scale / zs fp32 [rows, kw / group];
 * w = lut[code] * scale - zs (lut NULL = NF4). args->B is ignored. args->b_mn = 0: D = A x W^T (forward, N = rows,
 * K = kw); b_mn = 1: D = A x W (dgrad, N = kw, K = rows). LoRA operands (A2/B2), bias, alpha/beta and the stream-K
 * workspace work as in b200_gemm. Returns B200_ERR_UNSUPPORTED_ for shapes outside its domain (M <= 128,
 * N % 256, K % 64, group % 64, shard rows % 128 (forward) / % 64 (dgrad), fp32 inputs): callers fall back to
 * b200_w4_dequant + b200_gemm. */
typedef struct {
  const void* packed[8];
  const float* scale[8];
  const float* zs[8];   /* entries may be NULL */
  const float* lut;     /* 16 floats, or NULL for the NF4 code book */
  int32_t nshards;
  int32_t rows, kw, group;
  int64_t rows_per_shard;
} b200_w4_desc;

int b200_gemm_w4(const b200_gemm_args* args, const b200_w4_desc* w, void* stream);

/* ------------------------------------------------------------------------------------------
 * K5: RMSNorm / LayerNorm (layer_norm = 0 / 1) with the residual add fused in.
 *   fwd: h = x (+ res); y = norm(h) * w (+ b). stats[rows, 2] = (mean, rstd). h_out may be NULL.
 *   bwd: dx = d norm (+ dres_in); dw / db (optional) are reduced from per-CTA partial rows held in
 *        `scratch`, which must provide 2 * b200_norm_bwd_parts(rows) * cols floats.
 *   Replaces nn.LayerNorm (ddp_gpt_wikitext2.py:99-116) and HF Qwen3RMSNorm (Fine-Tuning/*.py).
 * ------------------------------------------------------------------------------------------ */
int b200_norm_fwd(int layer_norm, const void* x, const void* res, const void* w, const void* b, void* y,
                  void* h_out, float* stats, int rows, int cols, float eps, int dtype, int wdtype, void* stream);
int b200_norm_bwd_parts(int rows);
int b200_norm_bwd(int layer_norm, const void* dy, const void* h, const void* w, const float* stats,
                  const void* dres_in, void* dx, void* dw, void* db, float* scratch, int accumulate, int rows,
                  int cols, int dtype, int wdtype, void* stream);

/* b200_norm_bwd without a scratch argument, for callers that cannot allocate (a serving process
 * whose HBM is already committed to its KV-cache pool, a host that forbids cudaMalloc after
 * start-up, a bare C harness with no allocator behind it). This entry point has no counterpart
 * in the reference. The dw/db partial rows live in an 8 MiB array built into the library (resident on every
 * GPU that loads it, until exit); limits: cols <= b200_norm_bwd_noalloc_max_cols() (8192) for bf16 / fp16
 * activations and half of that (4096) for fp32 activations, else B200_ERR_TOO_LARGE with a message on
 * stderr and nothing runs. It calls no cudaMalloc/cudaFree. The first call on each
 * device looks the array's address up (cudaGetSymbolAddress) — make that call before any stream
 * capture. All calls share the array: nothing needs clearing between calls, but calls that may
 * overlap on the GPU race on it and must be serialised by the caller. */
int b200_norm_bwd_noalloc_max_cols(void);
int b200_norm_bwd_noalloc(int layer_norm, const void* dy, const void* h, const void* w, const float* stats,
                          const void* dres_in, void* dx, void* dw, void* db, int accumulate, int rows, int cols,
                          int dtype, int wdtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * K6 / K7 / K9 / K10: fused element-wise kernels.
 *   qknorm_rope: per-head RMSNorm of q,k (weights qw/kw, may be NULL) + rotary embedding on the
 *   fused qkv projection [tokens, (n_q + n_k + n_v) * D]; v heads are copied through.
 *   interleaved = 0: rotate-half pairs (HF Qwen3); 1: (2i, 2i+1) pairs
 *   (LLM_Distributed_Trainning/PyTorch/transformer_basics/DeepSeekLike_spare_MoE_wikitext2.py:145-174).
 *   swiglu: h = silu(gate) * up on [rows, 2F] (HF Qwen3MLP); gelu: exact erf (nn.GELU);
 *   embedding: gather (+ position table) / scatter-add; dropout: Philox mask from (seed, index).
 * ------------------------------------------------------------------------------------------ */
int b200_qknorm_rope_fwd(const void* src, void* dst, const void* qw, const void* kw, const float* cos_t,
                         const float* sin_t, const int* pos, int tokens, int n_q, int n_k, int n_v, int D,
                         long long src_rs, long long dst_rs, int interleaved, float eps, float* rstd_out, int dtype,
                         void* stream);
int b200_qknorm_rope_bwd(const void* g_in, const void* raw, void* g_out, const void* qw, const void* kw,
                         const float* cos_t, const float* sin_t, const int* pos, int tokens, int n_q, int n_k, int n_v,
                         int D, long long g_rs, long long raw_rs, long long out_rs, int interleaved, float eps,
                         float* dwq, float* dwk, int dtype, void* stream);
int b200_swiglu_fwd(const void* gu, void* h, long long rows, int F, int dtype, void* stream);
int b200_swiglu_bwd(const void* dh, const void* gu, void* dgu, long long rows, int F, int dtype, void* stream);
int b200_gelu_fwd(const void* x, void* y, long long n, int dtype, void* stream);
int b200_gelu_bwd(const void* dy, const void* x, void* dx, long long n, int dtype, void* stream);
int b200_embedding_fwd(const long long* ids, const void* table, const void* pos_table, void* out, long long tokens,
                       int dim, int period, float scale, int dtype, void* stream);
int b200_embedding_bwd(const long long* ids, const void* dout, float* dtable, long long tokens, int dim, float scale,
                       int dtype, void* stream);
int b200_dropout(const void* x, void* y, long long n, unsigned long long seed, float p, int dtype, void* stream);
int b200_cast(const void* x, void* y, long long n, float scale, int in_dtype, int out_dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * K8: cross-entropy with the gradient written in place of the logits, and the masked row softmax
 *   of the tiled attention path. Replaces nn.CrossEntropyLoss (ddp_gpt_wikitext2.py:282,304), the
 *   HF ForCausalLM loss, and the L x L boolean-mask softmax of nn.MultiheadAttention (:93-95).
 * ------------------------------------------------------------------------------------------ */
int b200_cross_entropy(void* logits, const long long* targets, float* loss_rows, float* lse_rows, long long rows,
                       int vocab, long long row_stride, long long ignore_index, float grad_scale, int write_grad,
                       int dtype, void* stream);
/* same, with a per-row factor read from the device (row_scale: float[rows] or NULL) multiplied into grad_scale */
int b200_cross_entropy_rows(void* logits, const long long* targets, float* loss_rows, float* lse_rows, long long rows,
                            int vocab, long long row_stride, long long ignore_index, float grad_scale, int write_grad,
                            const float* row_scale, int dtype, void* stream);
/* K8, vocabulary-chunked (no [tokens, vocab] tensor): online-softmax statistics / probabilities of one chunk of logits,
 * the gradient of one chunk once the row log-sum-exp is known, and the closing kernel of the single-pass variant
 * (dx = (acc / s - W[target]) * scale). See csrc/loss.cu and ops/loss.py (linear_cross_entropy). */
int b200_ce_chunk_online(void* lg, long long rows, long long row_stride, int C, int c0, const long long* targets,
                         long long ignore_index, float* m_run, float* s_run, float* tgt_logit, float* alpha_out,
                         int write_p, int dtype, void* stream);
int b200_ce_chunk_grad(void* lg, long long rows, long long row_stride, int C, int c0, const long long* targets,
                       long long ignore_index, const float* m_run, const float* s_run, const float* row_scale,
                       float grad_scale, int dtype, void* stream);
int b200_ce_online_finish(const float* acc, long long acc_rs, const void* W, long long w_rs, int K,
                          const long long* targets, long long ignore_index, const float* m_run, const float* s_run,
                          const float* tgt_logit, const float* row_scale, float grad_scale, void* dx, long long dx_rs,
                          float* loss_rows, long long rows, int dtype, void* stream);
int b200_softmax_fwd(const void* S, void* P, long long batch, int rows, int cols, int causal, int window, float scale,
                     int in_dtype, int out_dtype, void* stream);
int b200_softmax_bwd(const void* P, const void* dP, void* dS, long long total_rows, int cols, float scale,
                     int p_dtype, int g_dtype, int out_dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * K11 / K12 / K13 / K15: optimizer kernels over flat ranges. dyn_scale / found_inf are optional
 *   device scalars (clip coefficient, overflow flag). Replaces torch.optim.AdamW
 *   (ddp_gpt_wikitext2.py:281), bitsandbytes paged_adamw_8bit
 *   (Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:151), clip_grad_norm_ (llm-demo/minigpt/train.py:44).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  float lr, beta1, beta2, eps, weight_decay, bias_c1, bias_c2, grad_scale;
} b200_adam_hyper;

int b200_adamw(void* p, const void* g, float* m, float* v, void* p_lowp, long long n, const b200_adam_hyper* hy,
               const float* dyn_scale, const float* found_inf, int p_dtype, int g_dtype, void* stream);
/* b200_adamw with the hyper-parameters read from device memory (8 floats, b200_adam_hyper's order) inside the kernel, so
 * that a launch captured in a CUDA graph follows the schedule of later steps; b200_set_graph_state registers the device
 * uint64 that every dropout-drawing kernel (b200_dropout, flash attention with dropout) adds to its seed argument, for the
 * same reason. Both are refreshed by the caller (a captured copy from pinned host memory) before each replay. */
int b200_adamw_graph(void* p, const void* g, float* m, float* v, void* p_lowp, long long n, const void* hyper_dev,
                     const float* dyn_scale, const float* found_inf, int p_dtype, int g_dtype, void* stream);
int b200_set_graph_state(const void* seed_offset_dev);
int b200_adamw8bit(void* p, const void* g, uint8_t* s1, uint8_t* s2, float* absmax1, float* absmax2,
                   const float* qmap1, const float* qmap2, void* p_lowp, long long n, const b200_adam_hyper* hy,
                   const float* dyn_scale, const float* found_inf, int p_dtype, int g_dtype, void* stream);
int b200_sumsq(const void* x, long long n, float* out2, float pre_scale, int dtype, void* stream);
int b200_clip_coef(const float* sumsq, float max_norm, float* coef, float* norm_out, void* stream);
int b200_scale(void* x, long long n, float scale, const float* dyn_scale, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * K2 / K14: 4-bit weights. packed[N, K/2] (K contiguous, element 2j in the high nibble),
 *   scale[N, K/group] fp32, optional zs = zero_point * scale, 16-entry code book (NULL = NF4):
 *   w = lut[q] * scale - zs. `packed/scale/zs` of b200_w4_dequant are arrays of `nshards` base
 *   pointers (row shards; peer-GPU pointers allowed: the ZeRO-3 gather fused into the dequant).
 *   Replaces bitsandbytes NF4 kernels (Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:101-106) and
 *   the GPTQ/AWQ kernels of gptqmodel / vLLM (Quantization/GPTQModel/inference_qwen3_4b_gptq.py:29).
 * ------------------------------------------------------------------------------------------ */
int b200_nf4_quantize(const void* w, uint8_t* packed, float* absmax, long long n, int dtype, void* stream);
int b200_nf4_absmax_decode(const uint8_t* q, const float* code2, const float* absmax2, float offset, float* out,
                           long long n, void* stream);
int b200_w4_dequant(const void* const* packed, const void* const* scale, const void* const* zs, int nshards,
                    long long rows_per_shard, void* out, long long N, int K, int group, const float* lut,
                    long long out_rs, int out_dtype, void* stream);
/* Transposed variant: out_t [K, N] = W^T, so the backward dgrad GEMM (dY x W) reads a K-major operand
 * (N % 64 == 0, K % 64 == 0, rows_per_shard % 64 == 0). */
int b200_w4_dequant_t(const void* const* packed, const void* const* scale, const void* const* zs, int nshards,
                      long long rows_per_shard, void* out, long long N, int K, int group, const float* lut,
                      long long out_rs, int out_dtype, void* stream);
int b200_w4_gemv(const void* x, const uint8_t* packed, const float* scale, const float* zs, const float* lut,
                 const void* bias, void* y, int M, long long N, int K, int group, long long x_rs, long long y_rs,
                 int dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * K3 (small side): skinny LoRA products with dropout fused (bf16).
 *   Replaces peft lora.Linear's dropout -> lora_A -> lora_B chain (Fine-Tuning/qwen3-8b-lora-dist.py:118-125).
 * ------------------------------------------------------------------------------------------ */
int b200_lora_down(const void* X, const void* A, void* T, int M, int K, int r, long long x_rs, long long a_rs,
                   long long t_rs, float alpha, unsigned long long seed, float p, void* stream);
int b200_lora_wgrad(const void* P, const void* X, float* G, int M, int K, int r, long long p_rs, long long x_rs,
                    long long g_rs, float alpha, unsigned long long seed, float p, void* stream);
int b200_lora_dx(const void* dT, const void* A, void* dX, int M, int K, int r, long long dt_rs, long long a_rs,
                 long long dx_rs, float alpha, unsigned long long seed, float p, void* stream);

/* ------------------------------------------------------------------------------------------
 * K-C1 / K-C2 / K-C3: collectives fused with compute over NVLink peer memory. `*_ptrs` are arrays
 *   of `world` device pointers, index = rank (symmetric memory). Flags: uint32 [channels][8] per
 *   rank; epochs grow monotonically. grid_ctr is a local uint32 counter; grid_target_base is its
 *   value before this launch. Replaces ncclAllReduce under DDP (ddp_gpt_wikitext2.py:271-279) and
 *   DeepSpeed's reduce-scatter / optimizer / all-gather step (Fine-Tuning/ds_zero3_config.json).
 * ------------------------------------------------------------------------------------------ */
int b200_allreduce(const void* const* data_ptrs, const void* const* flag_ptrs, const void* const* lowp_ptrs,
                   long long n, int world, int rank, unsigned int epoch, float scale, unsigned int* grid_ctr,
                   unsigned int grid_target_base, int grid, void* stream);
/* The same all-reduce with the reduction done by NVSwitch (NVLS): mc_ptr is the MULTICAST address of the symmetric buffer
 * (torch symmetric memory: handle.multicast_ptr + byte offset). multimem.ld_reduce / multimem.st: 2 n / W bytes per rank
 * over NVLink instead of 2 n (W - 1) / W. dtype B200_DT_FP32 (n % 4 == 0; lowp_mc_ptr = multicast address of an optional
 * bf16 copy of the result) or B200_DT_BF16 (n % 8 == 0, fp32 accumulation in the switch, lowp_mc_ptr ignored). */
int b200_allreduce_nvls(void* mc_ptr, const void* const* flag_ptrs, void* lowp_mc_ptr, long long n, int world, int rank,
                        unsigned int epoch, float scale, int dtype, unsigned int* grid_ctr, unsigned int grid_target_base,
                        int grid, void* stream);

typedef struct {
  const void* const* grad_ptrs;
  const void* const* param_ptrs;
  const void* const* flag_ptrs;
  const void* const* norm_ptrs;
  float* gred; float* m; float* v;
  uint8_t* s1; uint8_t* s2; float* absmax1; float* absmax2;
  const float* qmap1; const float* qmap2;
  float* stats_out;
  long long n;
  int world, rank, eight_bit, reduce_only;
  unsigned int epoch;
  float lr, beta1, beta2, eps, weight_decay, bias_c1, bias_c2, grad_scale, max_norm;
  unsigned int* grid_ctr;
  unsigned int grid_target_base;
  int grid;
} b200_zero_step_args;

int b200_zero_step(const b200_zero_step_args* args, void* stream);
int b200_peer_gather(const void* const* src_ptrs, long long src_off_bytes, void* out, long long shard_bytes,
                     int world, void* stream);
/* One-shot all-reduce (sum) of a small 16-bit tensor for tensor-parallel decode: y [n] local, stage_ptrs = every rank's
 * symmetric 16-bit staging buffer of 2 * half_elems elements, out [n] local. Epoch and barrier tickets live in `state`
 * (3 zero-initialised uint32 on the device), so the launch can be captured in a CUDA graph and replayed on every rank
 * (b200_allreduce takes its epoch from the host). `channel` must not be shared with any other collective. n % 8 == 0.
 * vLLM's custom all-reduce behind --tensor-parallel-size (Fine-Tuning/README.md:345-348) is the counterpart. */
int b200_allreduce_oneshot(const void* y, const void* const* stage_ptrs, const void* const* flag_ptrs, void* out,
                           long long n, long long half_elems, int world, int rank, int channel, unsigned int* state,
                           int dtype, void* stream);

/* b200_peer_gather that leaves rank `skip`'s slot of `out` untouched (ZeRO-3 layer prefetch: the local shard is read
 * in place, only the world - 1 remote shards cross NVLink). */
int b200_peer_gather_skip(const void* const* src_ptrs, long long src_off_bytes, void* out, long long shard_bytes,
                          int world, int skip, void* stream);
/* the same gather issued as one cudaMemcpyAsync per owner (copy engines instead of SMs): the ZeRO-3 shard prefetch */
int b200_peer_gather_ce(const void* const* src_ptrs, long long src_off_bytes, void* out, long long shard_bytes,
                        int world, int skip, void* stream);
/* K-C2 for FSDP / ZeRO-2: reduce-scatter of one flat gradient unit over peer memory; the cast to the wire dtype
 * (fp32 or bf16 = MixedPrecision.reduce_dtype) and the scale are fused, the staging buffer (symmetric, 2 halves of
 * half_elems elements) is written by the kernel itself. out = this rank's n / world slice (+= when accumulate).
 * Replaces reduce_scatter_tensor inside FSDP (fsdp_gpt_wikitext2.py:304-312) and DeepSpeed's ZeRO-2 reduce-scatter. */
int b200_reduce_scatter(const float* grad, const void* const* wire_ptrs, const void* const* flag_ptrs, float* out,
                        long long n, long long half_elems, int world, int rank, unsigned int epoch, float scale,
                        int accumulate, int wire_dtype, unsigned int* grid_ctr, unsigned int grid_target_base, int grid,
                        void* stream);
int b200_peer_barrier(const void* const* flag_ptrs, int channel, int world, int rank, unsigned int epoch,
                      void* stream);

/* ---- decode attention (csrc/decode_attn.cu) ----------------------------------------------------
 * One new token per sequence: appends the rotated K/V of `qkv` [B, (H + 2 Hkv) * D] at position pos[b] of the
 * caches [n_slots, Hkv, max_len, D] (row slots[b], or b when slots is NULL) and writes softmax(q K^T * scale) V over positions 0..pos[b] (optionally the last
 * `window` ones) to out [B, H * D]. The KV-cache step behind HF generate(use_cache=True)
 * (Fine-Tuning/inferences.py:50-58, Scripts/inference/01-deepseek1.5b-hf-infr.py:8-30). */
int b200_decode_attn(const void* qkv, void* cache_k, void* cache_v, const int* pos, const int* slots, void* out, int B, int H,
                     int Hkv, int D, int max_len, float scale, int window, int dtype, void* stream);

/* Split ("flash-decoding") variant for serving batches: one CTA per (sequence, kv head, split) serves every query head
 * of the group (K/V read once per group), `nsplit` CTAs share the cached positions and a combine kernel merges them.
 * scratch: B * H * nsplit * (D + 2) floats. 16-bit dtypes, D in {64, 128}, H / Hkv in {1, 2, 4, 5, 8}; otherwise
 * B200_ERR_UNSUPPORTED_ (use b200_decode_attn). */
int b200_decode_attn_split(const void* qkv, void* cache_k, void* cache_v, const int* pos, const int* slots, void* out,
                           float* scratch, int B, int H, int Hkv, int D, int max_len, float scale, int nsplit,
                           int dtype, void* stream);

/* The same two kernels over a PAGED cache (vLLM's PagedAttention layout, the engine behind every `vllm serve` deployment
 * of the reference, LLM_on_Kubernetes/Inference_Platfrom/README.md:1504-1509): cache_k / cache_v are
 * [blocks, Hkv, block_size, D]; block_table [rows, table_stride] int32 on the device, row slots[b] (or b) lists the
 * physical blocks of sequence b in logical order; block_size a power of two. Position t of a sequence lives in block
 * block_table[row][t / block_size] at offset t % block_size. */
int b200_decode_attn_paged(const void* qkv, void* cache_k, void* cache_v, const int* pos, const int* slots, void* out, int B,
                           int H, int Hkv, int D, const int* block_table, int table_stride, int block_size, float scale,
                           int window, int dtype, void* stream);
int b200_decode_attn_split_paged(const void* qkv, void* cache_k, void* cache_v, const int* pos, const int* slots, void* out,
                                 float* scratch, int B, int H, int Hkv, int D, const int* block_table, int table_stride,
                                 int block_size, float scale, int nsplit, int dtype, void* stream);

/* ---- flash attention for training / prefill (csrc/flash_attn.cu) ----------------------------------
 * softmax(Q K^T * scale [+ causal mask]) V read in place from the fused projection qkv [B*S, (H + 2 Hkv) * D]; context to
 * out [B*S, H*D], row log-sum-exp (natural log) to lse [B, H, S]. No S x S tensor, no mask tensor (the reference builds
 * an L x L boolean mask per forward, ddp_gpt_wikitext2.py:93-95; HF Qwen3 uses SDPA, Fine-Tuning/qwen3-8b-lora-dist.py:118).
 * dropout_p > 0 drops attention probabilities (nn.MultiheadAttention(dropout=0.1), ddp_gpt_wikitext2.py:92) with a Philox
 * mask that depends only on (seed, batch, head, row, key): pass the same p and seed to the backward.
 * 16-bit dtypes, D in {32, 64, 96, 128}, any S. The backward recomputes P from lse: delta [B, H, S] and dqkv32 [B*S, W] (fp32)
 * are caller-provided work buffers (dqkv32 is zeroed inside), dqkv [B*S, W] receives the 16-bit gradient. */
int b200_flash_attn_fwd(const void* qkv, void* out, float* lse, int B, int S, int H, int Hkv, int D, int causal,
                        float scale, float dropout_p, unsigned long long seed, int dtype, void* stream);
/* Packed ("varlen") forward for serving prefill: qkv [T, W] holds n_seqs sequences back to back, sequence b = rows
 * [cu_seqlens[b], cu_seqlens[b+1]) (device int32 [n_seqs + 1]); max_seqlen bounds the grid. No lse. */
int b200_flash_attn_varlen_fwd(const void* qkv, void* out, const int* cu_seqlens, int n_seqs, int max_seqlen, int H,
                               int Hkv, int D, int causal, float scale, int dtype, void* stream);
int b200_flash_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                        float* dqkv32, void* dqkv, int B, int S, int H, int Hkv, int D, int causal, float scale,
                        float dropout_p, unsigned long long seed, int dtype, void* stream);
/* The same forward on the 5th-generation tensor cores (csrc/flash_tc.cu): tcgen05.mma for Q K^T and P V with both
 * accumulators in TMEM, Q / K / V tiles by TMA straight from qkv, one softmax thread per query row. head_dim 128, 16-bit
 * dtypes, no dropout; returns B200_ERR_UNSUPPORTED for anything else (callers then use b200_flash_attn_fwd). B * S rows
 * of qkv must be addressable as one 2-D tensor map (row stride (H + 2 Hkv) * 128 elements, 16-byte aligned base). */
int b200_flash_attn_fwd_tc(const void* qkv, void* out, float* lse, int B, int S, int H, int Hkv, int D, int causal,
                           float scale, int dtype, void* stream);

/* Backward on the same tensor cores (csrc/flash_bwd_tc.cu): a dK/dV pass (K, V resident per 64-key block, query tiles
 * stream) and a dQ pass (Q, dO resident per 128-row tile, key blocks stream), every product a tcgen05.mma with TMEM
 * accumulators; no atomics and no fp32 staging buffer (delta [B, H, S] is the only work buffer). head_dim 128, 16-bit,
 * no dropout; B200_ERR_UNSUPPORTED otherwise. */
int b200_flash_attn_bwd_tc(const void* qkv, const void* out, const void* dout, const float* lse, float* delta, void* dqkv,
                           int B, int S, int H, int Hkv, int D, int causal, float scale, int dtype, void* stream);

/* ---- small-M products for decode / serving steps (csrc/gemv.cu) ---------------------------------
 * y[m, n] = sum_k x[m, k] w[n, k] (+ bias[n]) for 1 <= M <= 64 tokens with the weights read exactly once (mma.sync with
 * the weight rows as the A operand, token groups of 8 as B). K % 128 == 0; x / w 16-byte aligned, row strides
 * multiples of 8 elements. b200_gemv16: 16-bit dense weights [N, w_rs], output 16-bit or fp32 (out_f32 = 1: the
 * lm_head logits). b200_w4_gemv_batched: 4-bit weights in the b200_w4_dequant layout. What HF generate / vLLM decode
 * spend their time in (Scripts/inference/07-deepseek1.5b-api-infr.py:118-131; serving table in
 * LLM_on_Kubernetes/Inference_Platfrom/README.md:1504-1511). */
int b200_gemv16(const void* x, const void* w, const void* bias, void* y, int M, long long N, int K, long long x_rs,
                long long w_rs, long long y_rs, int dtype, int out_f32, void* stream);
int b200_w4_gemv_batched(const void* x, const uint8_t* packed, const float* scale, const float* zs, const float* lut,
                         const void* bias, void* y, int M, long long N, int K, int group, long long x_rs, long long y_rs,
                         int dtype, void* stream);
/* The same for 4-bit codes whose table is linear, lut[c] = lin_a * (c - lin_off) with an integer lin_off (RTN / GPTQ / AWQ
 * integers: lin_a 1, lin_off 8 or 0): table-free conversion (nibble OR-ed into the mantissa of 128.0 / 1024.0, one exact
 * add, one FMA), 1 <= M <= 64. NF4 (non-linear code book) keeps the table kernels. */
int b200_w4_gemv_linear(const void* x, const uint8_t* packed, const float* scale, const float* zs, const void* bias, void* y,
                        int M, long long N, int K, int group, long long x_rs, long long y_rs, float lin_a, float lin_off,
                        int dtype, void* stream);

/* K-C5 (decode flavour): a small-M projection whose launch also does what follows it in a decoder layer — the
 * tensor-parallel all-reduce of the row-parallel partial products (each rank writes its tiles into its half of a
 * symmetric 16-bit staging buffer, publishes an epoch to the peers over NVLink and sums everyone's partials), the
 * residual add and the RMSNorm that feeds the next projection. world == 1: residual + norm only (stage_ptrs[0] is a
 * local [M, N] buffer). kind: 0 dense 16-bit weights, 1 4-bit codes via table (NF4), 2 4-bit integer codes (lin_a,
 * lin_off). state: 4 zero-initialised uint32 on the device (epoch + tickets; graph-replayable). N % 8 == 0, N <= 8192,
 * M <= 64. Replaces vLLM's row-parallel linear + custom all-reduce + fused_add_rms_norm sequence that the reference
 * reaches with --tensor-parallel-size (Quantization/README.md:175). */
typedef struct {
  const void* x; const void* w; const float* scale; const float* zs; const float* lut; const void* bias;
  int kind, M; long long N; int K, group; long long x_rs, w_rs; float lin_a, lin_off; int dtype;
  int world, rank, channel; const void* const* stage_ptrs; const void* const* flag_ptrs; long long half_elems;
  unsigned int* state; const void* res_in; void* res_out; const void* norm_w; void* x_out; float eps;
} b200_gemv_fused_args;
int b200_gemv_fused(const b200_gemv_fused_args* a, void* stream);

/* ---- gradient fold (csrc/elementwise.cu) -------------------------------------------------------
 * y (fp32) += scale * x (bf16/fp16), optionally clearing x: folds the low-precision .grad autograd produced for
 * non-GEMM parameters into the fp32 main-grad arena (DeepSpeed bf16 "grad_accum_dtype": fp32 behaviour;
 * LLM-Distributed-Training/04-DeepSpeed/ds_zero2_offload.json). */
int b200_fold_grad(void* x, float* y, long long n, float scale, int clear_x, int dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_OPS_H_ */
