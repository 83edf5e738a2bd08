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
