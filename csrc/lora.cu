// K3 (small side) — the skinny products of a LoRA adapter, with LoRA dropout fused in.
//
//   down : T[M, r]   = alpha * drop(X)[M, K] . A[r, K]^T          (forward; also dT = dY . B with A := B^T)
//   wgrad: G[r, K]  += alpha * P[M, r]^T . drop(X)[M, K]          (dA with P = dT; dB^T with P = T, X = dY)
//   dx   : dX[M, K] += mask/(1-p) * (dT[M, r] . A[r, K])          (only needed when dropout is on;
//                                                                  otherwise the dgrad GEMM absorbs it)
// The rank-r "up" product and the dropout-free dX term ride along in the tcgen05 GEMM as an extra
// k-block (gemm_tcgen05.cu); these CUDA-core kernels cover what is left, which is bandwidth-bound
// (each reads X once). The dropout mask is never stored: element (m, k) of X is kept iff
// philox(seed, m*K + k) >= p, so forward and backward regenerate identical masks.
//
// Replaces peft.tuners.lora.Linear's dropout -> lora_A -> lora_B chain and its autograd graph
// (LoraConfig in Fine-Tuning/qwen3-8b-lora-dist.py:118-125, Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:113-120).
#include "common.cuh"
#include "../include/b200_ops.h"

namespace b200 {

__device__ __forceinline__ uint4 lora_philox(unsigned long long seed, unsigned long long q) {
  constexpr uint32_t kA = 0xD2511F53u, kB = 0xCD9E8D57u, kW0 = 0x9E3779B9u, kW1 = 0xBB67AE85u;
  uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
  uint4 ctr = make_uint4((uint32_t)q, (uint32_t)(q >> 32), 0u, 0u);
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(kA, ctr.x), lo0 = kA * ctr.x;
    const uint32_t hi1 = __umulhi(kB, ctr.z), lo1 = kB * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += kW0;
    key.y += kW1;
  }
  return ctr;
}
// keep flags for 8 consecutive elements starting at element index e0 (e0 % 8 == 0); same stream of
// random words as b200_dropout (word j of philox(seed, idx / 4))
__device__ __forceinline__ void keep8(unsigned long long seed, unsigned long long e0, uint32_t thresh, bool (&k)[8]) {
  const uint4 a = lora_philox(seed, e0 >> 2), b = lora_philox(seed, (e0 >> 2) + 1);
  k[0] = a.x >= thresh; k[1] = a.y >= thresh; k[2] = a.z >= thresh; k[3] = a.w >= thresh;
  k[4] = b.x >= thresh; k[5] = b.y >= thresh; k[6] = b.z >= thresh; k[7] = b.w >= thresh;
}

// ---- down: one warp per row of X, A staged through shared memory in 256-column chunks ----------
template <int R>
__global__ void __launch_bounds__(256)
lora_down_kernel(const __nv_bfloat16* __restrict__ X, const __nv_bfloat16* __restrict__ A,
                 __nv_bfloat16* __restrict__ T, int M, int K, int r, long long x_rs, long long a_rs, long long t_rs,
                 float alpha, unsigned long long seed, uint32_t thresh, float inv_keep) {
  __shared__ __align__(16) __nv_bfloat16 As[R * 256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m = blockIdx.x * 8 + warp;
  float acc[R];
#pragma unroll
  for (int j = 0; j < R; ++j) acc[j] = 0.f;
  for (int k0 = 0; k0 < K; k0 += 256) {
    __syncthreads();
    // stage A[:, k0 : k0+256] (zero-padded)
    for (int i = threadIdx.x; i < R * 32; i += 256) {
      const int j = i >> 5, c = (i & 31) * 8;
      uint4 v = make_uint4(0, 0, 0, 0);
      if (j < r && k0 + c < K) v = *reinterpret_cast<const uint4*>(A + (long long)j * a_rs + k0 + c);
      *reinterpret_cast<uint4*>(As + j * 256 + c) = v;
    }
    __syncthreads();
    const int k = k0 + lane * 8;
    if (m < M && k < K) {
      const uint4 xr = *reinterpret_cast<const uint4*>(X + (long long)m * x_rs + k);
      const __nv_bfloat16* xe = reinterpret_cast<const __nv_bfloat16*>(&xr);
      float xv[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) xv[e] = __bfloat162float(xe[e]);
      if (thresh) {
        bool kp[8];
        keep8(seed, (unsigned long long)m * K + k, thresh, kp);
#pragma unroll
        for (int e = 0; e < 8; ++e) xv[e] = kp[e] ? xv[e] * inv_keep : 0.f;
      }
#pragma unroll
      for (int j = 0; j < R; ++j) {
        const uint4 ar = *reinterpret_cast<const uint4*>(As + j * 256 + lane * 8);
        const __nv_bfloat16* ae = reinterpret_cast<const __nv_bfloat16*>(&ar);
        float s = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) s += xv[e] * __bfloat162float(ae[e]);
        acc[j] += s;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const float s = warp_sum(acc[j]);
    if (lane == 0 && m < M && j < r) T[(long long)m * t_rs + j] = __float2bfloat16_rn(s * alpha);
  }
}

// ---- wgrad: grid (K / 256, row slices); thread owns one column, r accumulators; fp32 atomics ----
template <int R>
__global__ void __launch_bounds__(256)
lora_wgrad_kernel(const __nv_bfloat16* __restrict__ P, const __nv_bfloat16* __restrict__ X, float* __restrict__ G,
                  int M, int K, int r, long long p_rs, long long x_rs, long long g_rs, float alpha,
                  unsigned long long seed, uint32_t thresh, float inv_keep, int rows_per_slice) {
  __shared__ float Ps[64 * R];
  const int k = blockIdx.x * 256 + threadIdx.x;
  const int m_begin = blockIdx.y * rows_per_slice;
  const int m_end = min(M, m_begin + rows_per_slice);
  float acc[R];
#pragma unroll
  for (int j = 0; j < R; ++j) acc[j] = 0.f;
  for (int m0 = m_begin; m0 < m_end; m0 += 64) {
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * R; i += 256) {
      const int mm = i / R, j = i - mm * R;
      Ps[i] = (m0 + mm < m_end && j < r) ? __bfloat162float(P[(long long)(m0 + mm) * p_rs + j]) : 0.f;
    }
    __syncthreads();
    if (k < K) {
      const int lim = min(64, m_end - m0);
      for (int mm = 0; mm < lim; ++mm) {
        float xv = __bfloat162float(X[(long long)(m0 + mm) * x_rs + k]);
        if (thresh) {
          const unsigned long long e = (unsigned long long)(m0 + mm) * K + k;
          const uint4 b = lora_philox(seed, e >> 2);
          const uint32_t w = (e & 3) == 0 ? b.x : (e & 3) == 1 ? b.y : (e & 3) == 2 ? b.z : b.w;
          xv = w >= thresh ? xv * inv_keep : 0.f;
        }
#pragma unroll
        for (int j = 0; j < R; ++j) acc[j] += Ps[mm * R + j] * xv;
      }
    }
  }
  if (k < K) {
#pragma unroll
    for (int j = 0; j < R; ++j)
      if (j < r) atomicAdd(G + (long long)j * g_rs + k, acc[j] * alpha);
  }
}

// ---- dx: thread per 8 consecutive k ------------------------------------------------------------
template <int R>
__global__ void __launch_bounds__(256)
lora_dx_kernel(const __nv_bfloat16* __restrict__ dT, const __nv_bfloat16* __restrict__ A,
               __nv_bfloat16* __restrict__ dX, int M, int K, int r, long long dt_rs, long long a_rs, long long dx_rs,
               float alpha, unsigned long long seed, uint32_t thresh, float inv_keep) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int kchunks = K / 8;
  if (idx >= (long long)M * kchunks) return;
  const int m = (int)(idx / kchunks), k = (int)(idx - (long long)m * kchunks) * 8;
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  for (int j = 0; j < r; ++j) {
    const float t = __bfloat162float(dT[(long long)m * dt_rs + j]);
    const uint4 ar = *reinterpret_cast<const uint4*>(A + (long long)j * a_rs + k);
    const __nv_bfloat16* ae = reinterpret_cast<const __nv_bfloat16*>(&ar);
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] += t * __bfloat162float(ae[e]);
  }
  bool kp[8] = {true, true, true, true, true, true, true, true};
  if (thresh) keep8(seed, (unsigned long long)m * K + k, thresh, kp);
  uint4 old = *reinterpret_cast<const uint4*>(dX + (long long)m * dx_rs + k);
  __nv_bfloat16* oe = reinterpret_cast<__nv_bfloat16*>(&old);
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const float add = kp[e] ? acc[e] * alpha * (thresh ? inv_keep : 1.f) : 0.f;
    oe[e] = __float2bfloat16_rn(__bfloat162float(oe[e]) + add);
  }
  *reinterpret_cast<uint4*>(dX + (long long)m * dx_rs + k) = old;
}

static inline void drop_params(float p, uint32_t* thresh, float* inv_keep) {
  if (p <= 0.f) { *thresh = 0; *inv_keep = 1.f; return; }
  *thresh = p >= 1.f ? 0xFFFFFFFFu : (uint32_t)((double)p * 4294967296.0);
  *inv_keep = p >= 1.f ? 0.f : 1.f / (1.f - p);
}

}  // namespace b200
using namespace b200;

#define B200_R_DISPATCH(r, ...)                     \
  if (r <= 8) { constexpr int R = 8; __VA_ARGS__; }  \
  else if (r <= 16) { constexpr int R = 16; __VA_ARGS__; } \
  else if (r <= 32) { constexpr int R = 32; __VA_ARGS__; } \
  else if (r <= 64) { constexpr int R = 64; __VA_ARGS__; } \
  else return B200_ERR_TOO_LARGE;

B200_API int b200_lora_down(const void* X, const void* A, void* T, int M, int K, int r, long long x_rs,
                            long long a_rs, long long t_rs, float alpha, unsigned long long seed, float p,
                            void* stream) {
  if (M <= 0 || K % 8 || x_rs % 8 || a_rs % 8 || r < 1) return B200_ERR_BAD_SHAPE;
  uint32_t thresh; float inv_keep;
  drop_params(p, &thresh, &inv_keep);
  const unsigned grid = (unsigned)((M + 7) / 8);
  B200_R_DISPATCH(r, lora_down_kernel<R><<<grid, 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)X, (const __nv_bfloat16*)A, (__nv_bfloat16*)T, M, K, r, x_rs, a_rs, t_rs, alpha, seed, thresh, inv_keep));
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_lora_wgrad(const void* P, const void* X, float* G, int M, int K, int r, long long p_rs,
                             long long x_rs, long long g_rs, float alpha, unsigned long long seed, float p,
                             void* stream) {
  if (M <= 0 || r < 1) return B200_ERR_BAD_SHAPE;
  uint32_t thresh; float inv_keep;
  drop_params(p, &thresh, &inv_keep);
  const int kblocks = (K + 255) / 256;
  int slices = (2 * kNumSMs + kblocks - 1) / kblocks;
  int rows_per_slice = ((M + slices - 1) / slices + 63) / 64 * 64;
  slices = (M + rows_per_slice - 1) / rows_per_slice;
  dim3 grid(kblocks, slices);
  B200_R_DISPATCH(r, lora_wgrad_kernel<R><<<grid, 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)P, (const __nv_bfloat16*)X, G, M, K, r, p_rs, x_rs, g_rs, alpha, seed, thresh, inv_keep, rows_per_slice));
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_lora_dx(const void* dT, const void* A, void* dX, int M, int K, int r, long long dt_rs,
                          long long a_rs, long long dx_rs, float alpha, unsigned long long seed, float p,
                          void* stream) {
  if (M <= 0 || K % 8 || a_rs % 8 || dx_rs % 8 || r < 1) return B200_ERR_BAD_SHAPE;
  uint32_t thresh; float inv_keep;
  drop_params(p, &thresh, &inv_keep);
  const long long n = (long long)M * (K / 8);
  const unsigned grid = (unsigned)((n + 255) / 256);
  B200_R_DISPATCH(r, lora_dx_kernel<R><<<grid, 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dT, (const __nv_bfloat16*)A, (__nv_bfloat16*)dX, M, K, r, dt_rs, a_rs, dx_rs, alpha, seed, thresh, inv_keep));
  B200_LAUNCH_CHECK();
  return 0;
}
