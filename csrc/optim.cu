// K11 / K12 / K13 / K15 — optimizer-side kernels over flat parameter arenas:
//   * AdamW (decoupled weight decay, bias correction) with fp32 master weights and an optional
//     low-precision working copy written in the same pass;
//   * 8-bit block-wise AdamW (dynamic-tree quantised m and v with one absmax per 256 values), the
//     resident equivalent of bitsandbytes' paged_adamw_8bit;
//   * sum-of-squares / non-finite scan for global-norm clipping and fp16 dynamic loss scaling.
// The clip coefficient and the overflow flag stay on the device (the update kernels read them
// through pointers), so a training step needs no host synchronisation.
//
// Replaces torch.optim.AdamW (LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:281),
// DeepSpeed FusedAdam + fp16 loss scaler (LLM_Distributed_Trainning/DeepSpeed/*/ds_config.json),
// bitsandbytes PagedAdamW8bit (Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:151) and
// clip_grad_norm_ (llm-demo/minigpt/train.py:44; gradient_clipping in Fine-Tuning/ds_zero3_config.json:24).
#include "common.cuh"
#include "../include/b200_ops.h"

namespace b200 {

struct AdamHyper {
  float lr, beta1, beta2, eps, weight_decay;
  float bias_c1, bias_c2;  // 1 - beta^t
  float grad_scale;        // static multiplier on g (1 / loss_scale, 1 / accumulation steps ...)
};

template <typename PT, typename GT, typename LT>
__global__ void __launch_bounds__(256)
adamw_kernel(PT* __restrict__ p, const GT* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
             LT* __restrict__ p_lowp, long long n, AdamHyper h, const float* __restrict__ dyn_scale,
             const float* __restrict__ found_inf, const AdamHyper* __restrict__ h_dev) {
  if (found_inf && *found_inf != 0.f) return;  // overflowed step: skip (fp16 dynamic loss scaling)
  if (h_dev) h = *h_dev;                       // CUDA-graph replay: lr / bias corrections come from device memory
  const float gs = h.grad_scale * (dyn_scale ? *dyn_scale : 1.f);
  const float step = h.lr / h.bias_c1;
  const float rs2 = rsqrtf(h.bias_c2);
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const float gi = to_f32<GT>(g[i]) * gs;
    float pi = to_f32<PT>(p[i]);
    const float mi = h.beta1 * m[i] + (1.f - h.beta1) * gi;
    const float vi = h.beta2 * v[i] + (1.f - h.beta2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    pi *= (1.f - h.lr * h.weight_decay);
    pi -= step * mi / (sqrtf(vi) * rs2 + h.eps);
    p[i] = from_f32<PT>(pi);
    if (p_lowp) p_lowp[i] = from_f32<LT>(pi);
  }
}

// ---- 8-bit block-wise AdamW ---------------------------------------------------------------
constexpr int kQBlock = 256;

__device__ __forceinline__ uint8_t quantize_nearest(const float* __restrict__ qmap, float x) {
  // qmap is sorted ascending (256 entries); binary search for the nearest entry
  int lo = 0, hi = 255;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (qmap[mid] <= x) lo = mid; else hi = mid;
  }
  return (uint8_t)((x - qmap[lo]) <= (qmap[hi] - x) ? lo : hi);
}

template <typename PT, typename GT, typename LT>
__global__ void __launch_bounds__(kQBlock)
adamw8bit_kernel(PT* __restrict__ p, const GT* __restrict__ g, uint8_t* __restrict__ s1, uint8_t* __restrict__ s2,
                 float* __restrict__ absmax1, float* __restrict__ absmax2, const float* __restrict__ qmap1,
                 const float* __restrict__ qmap2, LT* __restrict__ p_lowp, long long n, AdamHyper h,
                 const float* __restrict__ dyn_scale, const float* __restrict__ found_inf) {
  if (found_inf && *found_inf != 0.f) return;
  __shared__ float q1[256], q2[256];
  __shared__ float red[32];
  q1[threadIdx.x] = qmap1[threadIdx.x];
  q2[threadIdx.x] = qmap2[threadIdx.x];
  __syncthreads();
  const long long blk = blockIdx.x;
  const long long i = blk * kQBlock + threadIdx.x;
  const bool ok = i < n;
  const float gs = h.grad_scale * (dyn_scale ? *dyn_scale : 1.f);
  float gi = 0.f, mi = 0.f, vi = 0.f;
  if (ok) {
    gi = to_f32<GT>(g[i]) * gs;
    mi = q1[s1[i]] * absmax1[blk];
    vi = q2[s2[i]] * absmax2[blk];
    mi = h.beta1 * mi + (1.f - h.beta1) * gi;
    vi = h.beta2 * vi + (1.f - h.beta2) * gi * gi;
  }
  const float a1 = block_max(fabsf(mi), red);
  const float a2 = block_max(fabsf(vi), red);
  if (threadIdx.x == 0) {
    absmax1[blk] = a1;
    absmax2[blk] = a2;
  }
  if (ok) {
    s1[i] = quantize_nearest(q1, a1 > 0.f ? mi / a1 : 0.f);
    s2[i] = quantize_nearest(q2, a2 > 0.f ? vi / a2 : 0.f);
    float pi = to_f32<PT>(p[i]);
    pi *= (1.f - h.lr * h.weight_decay);
    pi -= (h.lr / h.bias_c1) * mi / (sqrtf(vi) * rsqrtf(h.bias_c2) + h.eps);
    p[i] = from_f32<PT>(pi);
    if (p_lowp) p_lowp[i] = from_f32<LT>(pi);
  }
}

// ---- norms / overflow ---------------------------------------------------------------------
// out[0] += sum(x^2) ; out[1] = 1 if any element is inf/nan. The caller zeroes `out` first.
template <typename T>
__global__ void __launch_bounds__(256)
sumsq_kernel(const T* __restrict__ x, long long n, float* __restrict__ out, float pre_scale) {
  __shared__ float red[32];
  float s = 0.f;
  bool bad = false;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const float v = to_f32<T>(x[i]) * pre_scale;
    if (!isfinite(v)) bad = true;
    s += v * v;
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) atomicAdd(out, s);
  if (bad) out[1] = 1.f;
}

// coef = min(1, max_norm / (sqrt(sumsq) + 1e-6)); norm_out (optional) = sqrt(sumsq)
__global__ void clip_coef_kernel(const float* __restrict__ sumsq, float max_norm, float* __restrict__ coef,
                                 float* __restrict__ norm_out) {
  const float nrm = sqrtf(sumsq[0]);
  if (norm_out) *norm_out = nrm;
  float c = 1.f;
  if (max_norm > 0.f) c = fminf(1.f, max_norm / (nrm + 1e-6f));
  *coef = c;
}

// x *= scale * (*dyn_scale)   (in place; used by clip_grad_norm_ when no fused optimizer follows)
template <typename T>
__global__ void scale_kernel(T* __restrict__ x, long long n, float scale, const float* __restrict__ dyn_scale) {
  const float s = scale * (dyn_scale ? *dyn_scale : 1.f);
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) x[i] = from_f32<T>(to_f32<T>(x[i]) * s);
}

static inline unsigned flat_grid(long long n, int tb) {
  long long g = (n + tb - 1) / tb;
  const long long cap = (long long)kNumSMs * 8;
  return (unsigned)(g > cap ? cap : (g < 1 ? 1 : g));
}

}  // namespace b200
using namespace b200;

// p: params (p_dtype fp32 or bf16); g: grads (g_dtype); m, v: fp32; p_lowp: optional bf16 copy.
static int adamw_impl(void* p, const void* g, float* m, float* v, void* p_lowp, long long n,
                      const b200_adam_hyper* hy, const void* hyper_dev, const float* dyn_scale, const float* found_inf, int p_dtype,
                      int g_dtype, void* stream) {
  if (n <= 0) return 0;
  AdamHyper h = {};
  if (hy) h = {hy->lr, hy->beta1, hy->beta2, hy->eps, hy->weight_decay, hy->bias_c1, hy->bias_c2, hy->grad_scale};
  const AdamHyper* h_dev = reinterpret_cast<const AdamHyper*>(hyper_dev);
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = flat_grid(n, 256);
  __nv_bfloat16* lp = (__nv_bfloat16*)p_lowp;
#define B200_ADAM(PT, GT) adamw_kernel<PT, GT, __nv_bfloat16><<<grid, 256, 0, st>>>((PT*)p, (const GT*)g, m, v, lp, n, h, dyn_scale, found_inf, h_dev)
  if (p_dtype == B200_DT_FP32 && g_dtype == B200_DT_FP32) B200_ADAM(float, float);
  else if (p_dtype == B200_DT_FP32 && g_dtype == B200_DT_BF16) B200_ADAM(float, __nv_bfloat16);
  else if (p_dtype == B200_DT_BF16 && g_dtype == B200_DT_BF16) B200_ADAM(__nv_bfloat16, __nv_bfloat16);
  else if (p_dtype == B200_DT_BF16 && g_dtype == B200_DT_FP32) B200_ADAM(__nv_bfloat16, float);
  else if (p_dtype == B200_DT_FP32 && g_dtype == B200_DT_FP16) B200_ADAM(float, __half);
  else return B200_ERR_UNSUPPORTED;
#undef B200_ADAM
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_adamw(void* p, const void* g, float* m, float* v, void* p_lowp, long long n,
                        const b200_adam_hyper* hy, const float* dyn_scale, const float* found_inf, int p_dtype,
                        int g_dtype, void* stream) {
  if (!hy) return B200_ERR_BAD_SHAPE;
  return adamw_impl(p, g, m, v, p_lowp, n, hy, nullptr, dyn_scale, found_inf, p_dtype, g_dtype, stream);
}

// The same update with the hyper-parameters (8 floats in b200_adam_hyper's order: lr, beta1, beta2, eps, weight_decay,
// 1 - beta1^t, 1 - beta2^t, grad_scale) read from DEVICE memory inside the kernel: a launch captured in a CUDA graph
// follows the learning-rate schedule and the bias corrections of later steps (the caller refreshes hyper_dev, e.g. with a
// captured copy from pinned host memory, before each replay).
B200_API int b200_adamw_graph(void* p, const void* g, float* m, float* v, void* p_lowp, long long n, const void* hyper_dev,
                              const float* dyn_scale, const float* found_inf, int p_dtype, int g_dtype, void* stream) {
  if (!hyper_dev) return B200_ERR_BAD_SHAPE;
  return adamw_impl(p, g, m, v, p_lowp, n, nullptr, hyper_dev, dyn_scale, found_inf, p_dtype, g_dtype, stream);
}

B200_API int b200_adamw8bit(void* p, const void* g, uint8_t* s1, uint8_t* s2, float* absmax1, float* absmax2,
                            const float* qmap1, const float* qmap2, void* p_lowp, long long n,
                            const b200_adam_hyper* hy, const float* dyn_scale, const float* found_inf, int p_dtype,
                            int g_dtype, void* stream) {
  if (n <= 0) return 0;
  AdamHyper h = {hy->lr, hy->beta1, hy->beta2, hy->eps, hy->weight_decay, hy->bias_c1, hy->bias_c2, hy->grad_scale};
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = (unsigned)((n + kQBlock - 1) / kQBlock);
  __nv_bfloat16* lp = (__nv_bfloat16*)p_lowp;
#define B200_ADAM8(PT, GT) adamw8bit_kernel<PT, GT, __nv_bfloat16><<<grid, kQBlock, 0, st>>>((PT*)p, (const GT*)g, s1, s2, absmax1, absmax2, qmap1, qmap2, lp, n, h, dyn_scale, found_inf)
  if (p_dtype == B200_DT_FP32 && g_dtype == B200_DT_FP32) B200_ADAM8(float, float);
  else if (p_dtype == B200_DT_FP32 && g_dtype == B200_DT_BF16) B200_ADAM8(float, __nv_bfloat16);
  else if (p_dtype == B200_DT_BF16 && g_dtype == B200_DT_BF16) B200_ADAM8(__nv_bfloat16, __nv_bfloat16);
  else if (p_dtype == B200_DT_BF16 && g_dtype == B200_DT_FP32) B200_ADAM8(__nv_bfloat16, float);
  else return B200_ERR_UNSUPPORTED;
#undef B200_ADAM8
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_sumsq(const void* x, long long n, float* out2, float pre_scale, int dtype, void* stream) {
  if (n <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = flat_grid(n, 256);
  if (dtype == B200_DT_FP32) sumsq_kernel<float><<<grid, 256, 0, st>>>((const float*)x, n, out2, pre_scale);
  else if (dtype == B200_DT_BF16) sumsq_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, n, out2, pre_scale);
  else if (dtype == B200_DT_FP16) sumsq_kernel<__half><<<grid, 256, 0, st>>>((const __half*)x, n, out2, pre_scale);
  else return B200_ERR_UNSUPPORTED;
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_clip_coef(const float* sumsq, float max_norm, float* coef, float* norm_out, void* stream) {
  clip_coef_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(sumsq, max_norm, coef, norm_out);
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_scale(void* x, long long n, float scale, const float* dyn_scale, int dtype, void* stream) {
  if (n <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = flat_grid(n, 256);
  if (dtype == B200_DT_FP32) scale_kernel<float><<<grid, 256, 0, st>>>((float*)x, n, scale, dyn_scale);
  else if (dtype == B200_DT_BF16) scale_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((__nv_bfloat16*)x, n, scale, dyn_scale);
  else if (dtype == B200_DT_FP16) scale_kernel<__half><<<grid, 256, 0, st>>>((__half*)x, n, scale, dyn_scale);
  else return B200_ERR_UNSUPPORTED;
  B200_LAUNCH_CHECK();
  return 0;
}
