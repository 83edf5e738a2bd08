// CPU AdamW over a flat fp32 range — the host-side optimizer of ZeRO-Offload
// ("offload_optimizer": {"device": "cpu", "pin_memory": true}). It stands in for DeepSpeedCPUAdam,
// the one native op the reference invokes on purpose
// (LLM_Distributed_Trainning/DeepSpeed/DeepSpeed-GPTLike-ZeRO-Offload/DeepSpeed-GPTLike-ZeRO-Offload.py:27,228).
// OpenMP across cores, compiler-vectorised inner loop (built with -march=x86-64-v3: AVX2 + FMA);
// optionally writes a bf16 copy of the updated parameters for the upload back to the GPU.
#include <math.h>
#include <stdint.h>

#define B200_HOST_API extern "C" __attribute__((visibility("default")))

static inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  __builtin_memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);                                            // round to nearest even
  return (uint16_t)(u >> 16);
}

B200_HOST_API void b200_cpu_adamw(float* p, const float* g, float* m, float* v, uint16_t* p_bf16, int64_t n,
                                  float lr, float beta1, float beta2, float eps, float weight_decay, int64_t step,
                                  float grad_scale) {
  const float bc1 = 1.0f - powf(beta1, (float)step);
  const float bc2 = 1.0f - powf(beta2, (float)step);
  const float step_size = lr / bc1;
  const float rs2 = 1.0f / sqrtf(bc2);
  const float decay = 1.0f - lr * weight_decay;
#pragma omp parallel for simd schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    const float gi = g[i] * grad_scale;
    const float mi = beta1 * m[i] + (1.0f - beta1) * gi;
    const float vi = beta2 * v[i] + (1.0f - beta2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] = p[i] * decay - step_size * mi / (sqrtf(vi) * rs2 + eps);
  }
  if (p_bf16) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) p_bf16[i] = f32_to_bf16(p[i]);
  }
}

// sum of squares (for clipping on the host side of an offloaded step)
B200_HOST_API double b200_cpu_sumsq(const float* g, int64_t n, float scale) {
  double s = 0.0;
#pragma omp parallel for reduction(+ : s) schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    const double x = (double)g[i] * scale;
    s += x * x;
  }
  return s;
}
