// K4 (decode) — single-query attention over a KV cache, grouped-query aware.
//
// One CTA per (sequence, query head); 4 warps split the cached positions, each keeps an online
// softmax (running max, running sum, D/32 accumulators per lane) over its share, and the partial
// results are merged through shared memory. Memory-bound by design: every K and V element of the
// head is read exactly once, in 16-byte vectors. Also appends the new K/V row to the cache.
//
// Replaces the per-token attention inside HF `generate` / vLLM decode that the reference's inference
// scripts drive (Fine-Tuning/inferences.py:50-58, Scripts/inference/01..07-*.py,
// Quantization/GPTQModel/inference_qwen3_4b_gptq.py:58-68).
#include "common.cuh"
#include "../include/b200_ops.h"

namespace b200 {

// PER consecutive elements -> fp32, as one vector load when PER * sizeof(T) is 4, 8 or 16 bytes
template <typename T, int PER>
__device__ __forceinline__ void load_row(const T* __restrict__ p, float (&out)[PER]) {
  constexpr int BYTES = PER * (int)sizeof(T);
  if constexpr (BYTES == 16 || BYTES == 8 || BYTES == 4) {
    union { uint4 v16; uint2 v8; uint32_t v4; T e[16 / sizeof(T)]; } u;
    if constexpr (BYTES == 16) u.v16 = *reinterpret_cast<const uint4*>(p);
    else if constexpr (BYTES == 8) u.v8 = *reinterpret_cast<const uint2*>(p);
    else u.v4 = *reinterpret_cast<const uint32_t*>(p);
#pragma unroll
    for (int i = 0; i < PER; ++i) out[i] = to_f32<T>(u.e[i]);
  } else {
#pragma unroll
    for (int i = 0; i < PER; ++i) out[i] = to_f32<T>(p[i]);
  }
}

// qkv: [B, (H + 2 Hkv) * D] (post-RoPE) for the new token; cache_k / cache_v: [B, Hkv, max_len, D];
// pos[b] = index the new token is written at (= number of tokens already cached); slots[b] (optional) = cache row
// of batch row b; out: [B, H * D]
template <typename T, int D>
__global__ void __launch_bounds__(128)
decode_attn_kernel(const T* __restrict__ qkv, T* __restrict__ cache_k, T* __restrict__ cache_v,
                   const int* __restrict__ pos, const int* __restrict__ slots, T* __restrict__ out, int H, int Hkv, int max_len,
                   float scale, int window) {
  constexpr int PER = D / 32;  // elements per lane
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int kvh = h / (H / Hkv);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t_new = pos[b];
  const long long row = (long long)b * (H + 2 * Hkv) * D;
  const long long slot = slots ? slots[b] : b;  // continuous batching: batch row -> cache slot
  T* kc = cache_k + (slot * Hkv + kvh) * max_len * D;
  T* vc = cache_v + (slot * Hkv + kvh) * max_len * D;
  // append the new K / V row once per kv head (the first query head of the group does it)
  if (h % (H / Hkv) == 0 && warp == 0) {
    for (int i = lane; i < D; i += 32) {
      kc[(long long)t_new * D + i] = qkv[row + (long long)(H + kvh) * D + i];
      vc[(long long)t_new * D + i] = qkv[row + (long long)(H + Hkv + kvh) * D + i];
    }
  }
  float q[PER];
#pragma unroll
  for (int i = 0; i < PER; ++i) q[i] = to_f32<T>(qkv[row + (long long)h * D + lane * PER + i]) * scale;
  __syncthreads();
  // make the appended row visible to the other CTAs of this group is unnecessary: each of them re-reads
  // the new row from qkv instead of the cache (position t_new is handled separately below)
  float m = -INFINITY, l = 0.f, acc[PER];
#pragma unroll
  for (int i = 0; i < PER; ++i) acc[i] = 0.f;
  int t0 = 0;
  if (window > 0 && t_new + 1 > window) t0 = t_new + 1 - window;
  // Each warp takes blocks of 8 consecutive positions: 16 independent vector loads in flight per lane (the loop is
  // latency-bound otherwise: one 8-byte load and a 5-step shuffle reduction per position), one softmax rescale per block.
  constexpr int UN = 8;
  const T* knew = qkv + row + (long long)(H + kvh) * D + lane * PER;
  const T* vnew = qkv + row + (long long)(H + Hkv + kvh) * D + lane * PER;
  for (int tb = t0 + warp * UN; tb <= t_new; tb += 4 * UN) {
    float kx[UN][PER], vx[UN][PER], sc[UN];
#pragma unroll
    for (int u = 0; u < UN; ++u) {
      const int t = tb + u;
      if (t <= t_new) {
        load_row<T, PER>(t == t_new ? knew : kc + (long long)t * D + lane * PER, kx[u]);
        load_row<T, PER>(t == t_new ? vnew : vc + (long long)t * D + lane * PER, vx[u]);
      } else {
#pragma unroll
        for (int i = 0; i < PER; ++i) { kx[u][i] = 0.f; vx[u][i] = 0.f; }
      }
    }
    float bm = -INFINITY;
#pragma unroll
    for (int u = 0; u < UN; ++u) {
      float d = 0.f;
#pragma unroll
      for (int i = 0; i < PER; ++i) d += q[i] * kx[u][i];
      d = warp_sum(d);
      sc[u] = (tb + u <= t_new) ? d : -INFINITY;
      bm = fmaxf(bm, sc[u]);
    }
    const float mn = fmaxf(m, bm);
    const float c = __expf(m - mn);
    l *= c;
#pragma unroll
    for (int i = 0; i < PER; ++i) acc[i] *= c;
#pragma unroll
    for (int u = 0; u < UN; ++u) {
      const float pr = __expf(sc[u] - mn);      // exp(-inf) = 0 for the padded tail
      l += pr;
#pragma unroll
      for (int i = 0; i < PER; ++i) acc[i] += pr * vx[u][i];
    }
    m = mn;
  }
  __shared__ float sm[4], sl[4], sacc[4][D];
  if (lane == 0) { sm[warp] = m; sl[warp] = l; }
#pragma unroll
  for (int i = 0; i < PER; ++i) sacc[warp][lane * PER + i] = acc[i];
  __syncthreads();
  if (warp == 0) {
    float gm = fmaxf(fmaxf(sm[0], sm[1]), fmaxf(sm[2], sm[3]));
    float gl = 0.f;
    float w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { w[k] = sm[k] == -INFINITY ? 0.f : __expf(sm[k] - gm); gl += sl[k] * w[k]; }
    const float inv = gl > 0.f ? 1.f / gl : 0.f;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      const int d = lane * PER + i;
      const float o = sacc[0][d] * w[0] + sacc[1][d] * w[1] + sacc[2][d] * w[2] + sacc[3][d] * w[3];
      out[((long long)b * H + h) * D + d] = from_f32<T>(o * inv);
    }
  }
}

}  // namespace b200
using namespace b200;

B200_API int b200_decode_attn(const void* qkv, void* cache_k, void* cache_v, const int* pos, const int* slots, void* out, int B, int H,
                              int Hkv, int D, int max_len, float scale, int window, int dtype, void* stream) {
  if (B <= 0 || H % Hkv) return B200_ERR_BAD_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = (unsigned)(B * H);
#define B200_DA(T, DD) decode_attn_kernel<T, DD><<<grid, 128, 0, st>>>((const T*)qkv, (T*)cache_k, (T*)cache_v, pos, slots, (T*)out, H, Hkv, max_len, scale, window)
  if (dtype == B200_DT_BF16) {
    if (D == 128) B200_DA(__nv_bfloat16, 128); else if (D == 64) B200_DA(__nv_bfloat16, 64);
    else if (D == 32) B200_DA(__nv_bfloat16, 32); else if (D == 96) B200_DA(__nv_bfloat16, 96); else return B200_ERR_UNSUPPORTED;
  } else if (dtype == B200_DT_FP16) {
    if (D == 128) B200_DA(__half, 128); else if (D == 64) B200_DA(__half, 64);
    else if (D == 32) B200_DA(__half, 32); else return B200_ERR_UNSUPPORTED;
  } else if (dtype == B200_DT_FP32) {
    if (D == 128) B200_DA(float, 128); else if (D == 64) B200_DA(float, 64);
    else if (D == 32) B200_DA(float, 32); else return B200_ERR_UNSUPPORTED;
  } else return B200_ERR_UNSUPPORTED;
#undef B200_DA
  B200_LAUNCH_CHECK();
  return 0;
}
