// K8 — fused cross-entropy (forward + gradient in one launch) and the row softmax kernels of the
// materialised-score attention path.
//
// cross_entropy: one CTA per token row. Pass 1 streams the logits row once with an online
// (max, sum-exp) reduction; pass 2 (an L2 hit for a 300 KB row) overwrites the row in place with
// d loss / d logits = (softmax - onehot) * grad_scale, so the backward of the lm_head is just the
// two GEMMs and no [tokens, vocab] fp32 tensor is ever allocated. ignore_index rows give zero loss
// and zero gradient. Per-row losses are written to loss_rows; the caller reduces them (mean over
// the non-ignored rows, as nn.CrossEntropyLoss / HF ForCausalLM do).
//
// Replaces ATen log_softmax + nll_loss behind nn.CrossEntropyLoss
// (LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:282,304) and the fp32-upcast
// loss in HF Qwen3ForCausalLM used by every Fine-Tuning/*.py script.
#include "common.cuh"
#include "../include/b200_ops.h"

namespace b200 {

constexpr int kCEThreads = 512;

template <typename T>
__global__ void __launch_bounds__(kCEThreads)
cross_entropy_kernel(T* __restrict__ logits, const long long* __restrict__ targets, float* __restrict__ loss_rows,
                     float* __restrict__ lse_rows, int vocab, long long row_stride, long long ignore_index,
                     float grad_scale, int write_grad, const float* __restrict__ row_scale) {
  constexpr int V = 16 / sizeof(T);
  __shared__ float red[32];
  const long long row = blockIdx.x;
  T* lr = logits + row * row_stride;
  const long long tgt = targets[row];
  const bool ignored = (tgt == ignore_index) || tgt < 0 || tgt >= vocab;
  const int nvec = vocab / V;

  // pass 1: online max / sum-exp
  float m = -INFINITY, s = 0.f;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    uint4 raw = *reinterpret_cast<const uint4*>(lr + (long long)i * V);
    const T* e = reinterpret_cast<const T*>(&raw);
    float v[V];
    float lm = -INFINITY;
#pragma unroll
    for (int j = 0; j < V; ++j) { v[j] = to_f32<T>(e[j]); lm = fmaxf(lm, v[j]); }
    if (lm > m) { s *= __expf(m - lm); m = lm; }
#pragma unroll
    for (int j = 0; j < V; ++j) s += __expf(v[j] - m);
  }
  for (int c = nvec * V + threadIdx.x; c < vocab; c += blockDim.x) {
    const float v = to_f32<T>(lr[c]);
    if (v > m) { s *= __expf(m - v); m = v; }
    s += __expf(v - m);
  }
  const float gm = block_max(m, red);
  s = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
  const float gs = block_sum(s, red);
  const float lse = gm + __logf(gs);
  if (threadIdx.x == 0) {
    const float tl = ignored ? 0.f : to_f32<T>(lr[tgt]);
    loss_rows[row] = ignored ? 0.f : (lse - tl);
    if (lse_rows) lse_rows[row] = lse;
  }
  if (!write_grad) return;
  __syncthreads();  // the target logit has been read before anyone overwrites the row
  // row_scale: per-row factor computed on the device (1 / (n_valid of the row's micro-batch * accumulation steps)),
  // so the host never has to read the count of non-ignored targets back
  const float gsc = ignored ? 0.f : grad_scale * (row_scale ? row_scale[row] : 1.f);
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    uint4 raw = *reinterpret_cast<const uint4*>(lr + (long long)i * V);
    const T* e = reinterpret_cast<const T*>(&raw);
    uint4 out;
    T* o = reinterpret_cast<T*>(&out);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const int c = i * V + j;
      float p = __expf(to_f32<T>(e[j]) - lse);
      if (c == tgt) p -= 1.f;
      o[j] = from_f32<T>(p * gsc);
    }
    *reinterpret_cast<uint4*>(lr + (long long)i * V) = out;
  }
  for (int c = nvec * V + threadIdx.x; c < vocab; c += blockDim.x) {
    float p = __expf(to_f32<T>(lr[c]) - lse);
    if (c == tgt) p -= 1.f;
    lr[c] = from_f32<T>(p * gsc);
  }
}

// ---------------------------------------------------------------------------------------------
// Row softmax over attention scores with an implicit causal mask (no mask tensor is ever built —
// the reference rebuilds an L x L boolean mask every forward, ddp_gpt_wikitext2.py:93-95).
//   S: [batch, rows, cols] (TI, usually fp32 scores already scaled) -> P: same shape (TO).
//   causal: column c of row r is kept iff c <= r + (cols - rows); window > 0 also drops
//   c < r + (cols - rows) - window + 1 (sliding-window / local attention).
// ---------------------------------------------------------------------------------------------
template <typename TI, typename TO>
__global__ void __launch_bounds__(256)
softmax_fwd_kernel(const TI* __restrict__ S, TO* __restrict__ P, int rows, int cols, int causal, int window,
                   float scale) {
  __shared__ float red[32];
  const long long r_glob = blockIdx.x;
  const int r = (int)(r_glob % rows);
  const TI* s = S + r_glob * cols;
  TO* p = P + r_glob * cols;
  int hi = cols - 1, lo = 0;
  if (causal) {
    hi = r + (cols - rows);
    if (hi > cols - 1) hi = cols - 1;
    if (window > 0) { lo = hi - window + 1; if (lo < 0) lo = 0; }
  }
  float m = -INFINITY;
  for (int c = lo + threadIdx.x; c <= hi; c += blockDim.x) m = fmaxf(m, to_f32<TI>(s[c]) * scale);
  m = block_max(m, red);
  float sum = 0.f;
  for (int c = lo + threadIdx.x; c <= hi; c += blockDim.x) sum += __expf(to_f32<TI>(s[c]) * scale - m);
  sum = block_sum(sum, red);
  const float inv = sum > 0.f ? 1.f / sum : 0.f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) {
    float v = 0.f;
    if (c >= lo && c <= hi) v = __expf(to_f32<TI>(s[c]) * scale - m) * inv;
    p[c] = from_f32<TO>(v);
  }
}

// dS = P * (dP - sum_c(dP * P)) * scale
template <typename TP, typename TG, typename TO>
__global__ void __launch_bounds__(256)
softmax_bwd_kernel(const TP* __restrict__ P, const TG* __restrict__ dP, TO* __restrict__ dS, int cols, float scale) {
  __shared__ float red[32];
  const long long r = blockIdx.x;
  const TP* p = P + r * cols;
  const TG* g = dP + r * cols;
  TO* o = dS + r * cols;
  float dot = 0.f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) dot += to_f32<TP>(p[c]) * to_f32<TG>(g[c]);
  dot = block_sum(dot, red);
  for (int c = threadIdx.x; c < cols; c += blockDim.x)
    o[c] = from_f32<TO>(to_f32<TP>(p[c]) * (to_f32<TG>(g[c]) - dot) * scale);
}


// ---------------------------------------------------------------------------------------------
// K8 as a vocabulary-chunked, flash-attention-style reduction: the lm_head GEMM produces one [rows, C] chunk of logits at
// a time (C = 8192 columns: 32 MB for 2048 rows, it stays in the 126 MB L2 between its GEMM and these kernels), so no
// [tokens, vocab] tensor ever exists.
//   ce_chunk_online  per row: m' = max(m, max_c l); alpha = exp(m - m'); l <- p = exp(l - m') in place; s = s * alpha + sum p;
//                    the target's logit is captured when it falls in this chunk. With write_p = 0 only the statistics are
//                    updated (first pass of the two-pass variant for a trainable head).
//                    The caller then accumulates acc = alpha[row] * acc + p @ W_chunk (GEMM with per-row beta): after the
//                    last chunk acc / s = softmax(l) @ W, i.e. dL/dx up to the one-hot term — ONE pass over the vocabulary.
//   ce_chunk_grad    second pass of the two-pass variant: l <- (exp(l - lse) - [col == target]) * scale[row] in place.
//   ce_online_finish loss[row] = m + log s - l_target;  dx[row] = (acc[row] / s - W[target]) * scale[row].
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kCEThreads)
ce_chunk_online_kernel(T* __restrict__ lg, long long row_stride, int C, int c0, const long long* __restrict__ targets,
                       long long ignore_index, float* __restrict__ m_run, float* __restrict__ s_run,
                       float* __restrict__ tgt_logit, float* __restrict__ alpha_out, int write_p) {
  constexpr int V = 16 / sizeof(T);
  __shared__ float red[32];
  const long long row = blockIdx.x;
  T* lr = lg + row * row_stride;
  const int nvec = C / V;                       // C % V == 0 (the host pads the chunk width)
  float cm = -INFINITY;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    const uint4 raw = *reinterpret_cast<const uint4*>(lr + (long long)i * V);
    const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
    for (int j = 0; j < V; ++j) cm = fmaxf(cm, to_f32<T>(e[j]));
  }
  cm = block_max(cm, red);
  const float m_old = m_run[row];
  const float m_new = fmaxf(m_old, cm);
  const long long tgt = targets[row];
  if (threadIdx.x == 0 && tgt != ignore_index && tgt >= c0 && tgt < c0 + C) tgt_logit[row] = to_f32<T>(lr[tgt - c0]);
  __syncthreads();                               // the target logit is read before the row is overwritten
  float sum = 0.f;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    const uint4 raw = *reinterpret_cast<const uint4*>(lr + (long long)i * V);
    const T* e = reinterpret_cast<const T*>(&raw);
    uint4 out;
    T* o = reinterpret_cast<T*>(&out);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const float pj = __expf(to_f32<T>(e[j]) - m_new);
      o[j] = from_f32<T>(pj);
      sum += write_p ? to_f32<T>(o[j]) : pj;     // the sum uses the rounded p that the GEMM will see
    }
    if (write_p) *reinterpret_cast<uint4*>(lr + (long long)i * V) = out;
  }
  sum = block_sum(sum, red);
  if (threadIdx.x == 0) {
    const float a = (m_old == -INFINITY) ? 0.f : __expf(m_old - m_new);
    s_run[row] = s_run[row] * a + sum;
    m_run[row] = m_new;
    if (alpha_out) alpha_out[row] = a;
  }
}

template <typename T>
__global__ void __launch_bounds__(kCEThreads)
ce_chunk_grad_kernel(T* __restrict__ lg, long long row_stride, int C, int c0, const long long* __restrict__ targets,
                     long long ignore_index, const float* __restrict__ m_run, const float* __restrict__ s_run,
                     const float* __restrict__ row_scale, float grad_scale) {
  constexpr int V = 16 / sizeof(T);
  const long long row = blockIdx.x;
  T* lr = lg + row * row_stride;
  const long long tgt = targets[row];
  const bool ignored = tgt == ignore_index;
  const float lse = m_run[row] + __logf(s_run[row]);
  const float gsc = ignored ? 0.f : grad_scale * (row_scale ? row_scale[row] : 1.f);
  const int nvec = C / V;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    const uint4 raw = *reinterpret_cast<const uint4*>(lr + (long long)i * V);
    const T* e = reinterpret_cast<const T*>(&raw);
    uint4 out;
    T* o = reinterpret_cast<T*>(&out);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float pj = __expf(to_f32<T>(e[j]) - lse);
      if ((long long)(c0 + i * V + j) == tgt) pj -= 1.f;
      o[j] = from_f32<T>(pj * gsc);
    }
    *reinterpret_cast<uint4*>(lr + (long long)i * V) = out;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
ce_online_finish_kernel(const float* __restrict__ acc, long long acc_rs, const T* __restrict__ W, long long w_rs, int K,
                        const long long* __restrict__ targets, long long ignore_index, const float* __restrict__ m_run,
                        const float* __restrict__ s_run, const float* __restrict__ tgt_logit,
                        const float* __restrict__ row_scale, float grad_scale, T* __restrict__ dx, long long dx_rs,
                        float* __restrict__ loss_rows) {
  const long long row = blockIdx.x;
  const long long tgt = targets[row];
  const bool ignored = tgt == ignore_index;
  const float s = s_run[row];
  if (threadIdx.x == 0) loss_rows[row] = ignored ? 0.f : (m_run[row] + __logf(s) - tgt_logit[row]);
  if (dx == nullptr) return;
  const float gsc = ignored ? 0.f : grad_scale * (row_scale ? row_scale[row] : 1.f);
  const float inv = 1.f / s;
  const float* ar = acc + row * acc_rs;
  const T* wr = W + (ignored ? 0 : tgt) * w_rs;
  T* dr = dx + row * dx_rs;
  for (int k = threadIdx.x; k < K; k += blockDim.x)
    dr[k] = from_f32<T>((ar[k] * inv - to_f32<T>(wr[k])) * gsc);
}

}  // namespace b200
using namespace b200;

// row_scale (may be NULL): float[rows], multiplied into grad_scale per row (device-computed normalisation).
B200_API int b200_cross_entropy_rows(void* logits, const long long* targets, float* loss_rows, float* lse_rows,
                                     long long rows, int vocab, long long row_stride, long long ignore_index,
                                     float grad_scale, int write_grad, const float* row_scale, int dtype, void* stream) {
  if (rows <= 0 || vocab <= 0) return B200_ERR_BAD_SHAPE;
  const int es = dtype == B200_DT_FP32 ? 4 : 2;
  if ((row_stride * es) % 16 || (reinterpret_cast<uintptr_t>(logits) & 15)) return B200_ERR_BAD_ALIGN;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_DT_BF16)
    cross_entropy_kernel<__nv_bfloat16><<<(unsigned)rows, kCEThreads, 0, st>>>(
        (__nv_bfloat16*)logits, targets, loss_rows, lse_rows, vocab, row_stride, ignore_index, grad_scale, write_grad, row_scale);
  else if (dtype == B200_DT_FP32)
    cross_entropy_kernel<float><<<(unsigned)rows, kCEThreads, 0, st>>>(
        (float*)logits, targets, loss_rows, lse_rows, vocab, row_stride, ignore_index, grad_scale, write_grad, row_scale);
  else if (dtype == B200_DT_FP16)
    cross_entropy_kernel<__half><<<(unsigned)rows, kCEThreads, 0, st>>>(
        (__half*)logits, targets, loss_rows, lse_rows, vocab, row_stride, ignore_index, grad_scale, write_grad, row_scale);
  else
    return B200_ERR_UNSUPPORTED;
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_cross_entropy(void* logits, const long long* targets, float* loss_rows, float* lse_rows,
                                long long rows, int vocab, long long row_stride, long long ignore_index,
                                float grad_scale, int write_grad, int dtype, void* stream) {
  return b200_cross_entropy_rows(logits, targets, loss_rows, lse_rows, rows, vocab, row_stride, ignore_index, grad_scale,
                                 write_grad, nullptr, dtype, stream);
}

B200_API int b200_softmax_fwd(const void* S, void* P, long long batch, int rows, int cols, int causal, int window,
                              float scale, int in_dtype, int out_dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = (unsigned)(batch * rows);
  if (in_dtype == B200_DT_FP32 && out_dtype == B200_DT_BF16)
    softmax_fwd_kernel<float, __nv_bfloat16><<<grid, 256, 0, st>>>((const float*)S, (__nv_bfloat16*)P, rows, cols, causal, window, scale);
  else if (in_dtype == B200_DT_FP32 && out_dtype == B200_DT_FP32)
    softmax_fwd_kernel<float, float><<<grid, 256, 0, st>>>((const float*)S, (float*)P, rows, cols, causal, window, scale);
  else if (in_dtype == B200_DT_FP32 && out_dtype == B200_DT_FP16)
    softmax_fwd_kernel<float, __half><<<grid, 256, 0, st>>>((const float*)S, (__half*)P, rows, cols, causal, window, scale);
  else if (in_dtype == B200_DT_BF16 && out_dtype == B200_DT_BF16)
    softmax_fwd_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)S, (__nv_bfloat16*)P, rows, cols, causal, window, scale);
  else
    return B200_ERR_UNSUPPORTED;
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_softmax_bwd(const void* P, const void* dP, void* dS, long long total_rows, int cols, float scale,
                              int p_dtype, int g_dtype, int out_dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = (unsigned)total_rows;
  if (p_dtype == B200_DT_BF16 && g_dtype == B200_DT_FP32 && out_dtype == B200_DT_BF16)
    softmax_bwd_kernel<__nv_bfloat16, float, __nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)P, (const float*)dP, (__nv_bfloat16*)dS, cols, scale);
  else if (p_dtype == B200_DT_FP32 && g_dtype == B200_DT_FP32 && out_dtype == B200_DT_FP32)
    softmax_bwd_kernel<float, float, float><<<grid, 256, 0, st>>>((const float*)P, (const float*)dP, (float*)dS, cols, scale);
  else if (p_dtype == B200_DT_FP16 && g_dtype == B200_DT_FP32 && out_dtype == B200_DT_FP16)
    softmax_bwd_kernel<__half, float, __half><<<grid, 256, 0, st>>>((const __half*)P, (const float*)dP, (__half*)dS, cols, scale);
  else if (p_dtype == B200_DT_BF16 && g_dtype == B200_DT_BF16 && out_dtype == B200_DT_BF16)
    softmax_bwd_kernel<__nv_bfloat16, __nv_bfloat16, __nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)P, (const __nv_bfloat16*)dP, (__nv_bfloat16*)dS, cols, scale);
  else
    return B200_ERR_UNSUPPORTED;
  B200_LAUNCH_CHECK();
  return 0;
}

// dtype: 16-bit logits chunk [rows, C] with row stride row_stride (elements); C % 8 == 0 (pad the chunk; padded columns must
// hold -inf / a large negative so that they contribute exp() = 0). m_run must start at -inf, s_run at 0.
B200_API int b200_ce_chunk_online(void* lg, long long rows, long long row_stride, int C, int c0, const long long* targets,
                                  long long ignore_index, float* m_run, float* s_run, float* tgt_logit, float* alpha_out,
                                  int write_p, int dtype, void* stream) {
  if (rows <= 0 || C <= 0 || C % 8 || (row_stride * 2) % 16 || (reinterpret_cast<uintptr_t>(lg) & 15)) return B200_ERR_BAD_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_DT_BF16)
    ce_chunk_online_kernel<__nv_bfloat16><<<(unsigned)rows, kCEThreads, 0, st>>>((__nv_bfloat16*)lg, row_stride, C, c0, targets,
                                                                                 ignore_index, m_run, s_run, tgt_logit, alpha_out, write_p);
  else if (dtype == B200_DT_FP16)
    ce_chunk_online_kernel<__half><<<(unsigned)rows, kCEThreads, 0, st>>>((__half*)lg, row_stride, C, c0, targets, ignore_index,
                                                                          m_run, s_run, tgt_logit, alpha_out, write_p);
  else return B200_ERR_UNSUPPORTED;
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_ce_chunk_grad(void* lg, long long rows, long long row_stride, int C, int c0, const long long* targets,
                                long long ignore_index, const float* m_run, const float* s_run, const float* row_scale,
                                float grad_scale, int dtype, void* stream) {
  if (rows <= 0 || C <= 0 || C % 8 || (row_stride * 2) % 16) return B200_ERR_BAD_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_DT_BF16)
    ce_chunk_grad_kernel<__nv_bfloat16><<<(unsigned)rows, kCEThreads, 0, st>>>((__nv_bfloat16*)lg, row_stride, C, c0, targets,
                                                                               ignore_index, m_run, s_run, row_scale, grad_scale);
  else if (dtype == B200_DT_FP16)
    ce_chunk_grad_kernel<__half><<<(unsigned)rows, kCEThreads, 0, st>>>((__half*)lg, row_stride, C, c0, targets, ignore_index,
                                                                        m_run, s_run, row_scale, grad_scale);
  else return B200_ERR_UNSUPPORTED;
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_ce_online_finish(const float* acc, long long acc_rs, const void* W, long long w_rs, int K,
                                   const long long* targets, long long ignore_index, const float* m_run, const float* s_run,
                                   const float* tgt_logit, const float* row_scale, float grad_scale, void* dx, long long dx_rs,
                                   float* loss_rows, long long rows, int dtype, void* stream) {
  if (rows <= 0) return B200_ERR_BAD_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_DT_BF16)
    ce_online_finish_kernel<__nv_bfloat16><<<(unsigned)rows, 256, 0, st>>>(acc, acc_rs, (const __nv_bfloat16*)W, w_rs, K, targets,
                                                                           ignore_index, m_run, s_run, tgt_logit, row_scale,
                                                                           grad_scale, (__nv_bfloat16*)dx, dx_rs, loss_rows);
  else if (dtype == B200_DT_FP16)
    ce_online_finish_kernel<__half><<<(unsigned)rows, 256, 0, st>>>(acc, acc_rs, (const __half*)W, w_rs, K, targets, ignore_index,
                                                                    m_run, s_run, tgt_logit, row_scale, grad_scale, (__half*)dx,
                                                                    dx_rs, loss_rows);
  else return B200_ERR_UNSUPPORTED;
  B200_LAUNCH_CHECK();
  return 0;
}
