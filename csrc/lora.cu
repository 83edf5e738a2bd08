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

#include <stdlib.h>

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

// ---- down: a CTA owns ROWS rows; its 8 warps split K in 256-column chunks (lane = 8 columns), every lane keeps
// ROWS x R partial sums, A is read straight from L2 (it is r x K, shared by every CTA) and reused for the ROWS rows.
// No shared-memory staging and no block barriers in the main loop: all loads of a chunk are independent.
template <int R, int ROWS>
__global__ void __launch_bounds__(256)
lora_down_kernel(const __nv_bfloat16* __restrict__ X, const __nv_bfloat16* __restrict__ A,
                 __nv_bfloat16* __restrict__ T, int M, int K, int r, long long x_rs, long long a_rs, long long t_rs,
                 float alpha, unsigned long long seed, uint32_t thresh, float inv_keep) {
  __shared__ float red[8][ROWS * R];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * ROWS;
  float acc[ROWS][R];
#pragma unroll
  for (int i = 0; i < ROWS; ++i)
#pragma unroll
    for (int j = 0; j < R; ++j) acc[i][j] = 0.f;
  for (int k = warp * 256 + lane * 8; k < K; k += 2048) {
    float xv[ROWS][8];
#pragma unroll
    for (int i = 0; i < ROWS; ++i) {
      uint4 xr = make_uint4(0, 0, 0, 0);
      if (m0 + i < M) xr = *reinterpret_cast<const uint4*>(X + (long long)(m0 + i) * x_rs + k);
      const __nv_bfloat16* xe = reinterpret_cast<const __nv_bfloat16*>(&xr);
#pragma unroll
      for (int e = 0; e < 8; ++e) xv[i][e] = __bfloat162float(xe[e]);
      if (thresh) {
        bool kp[8];
        keep8(seed, (unsigned long long)(m0 + i) * K + k, thresh, kp);
#pragma unroll
        for (int e = 0; e < 8; ++e) xv[i][e] = kp[e] ? xv[i][e] * inv_keep : 0.f;
      }
    }
#pragma unroll
    for (int j = 0; j < R; ++j) {
      if (j < r) {
        const uint4 ar = __ldg(reinterpret_cast<const uint4*>(A + (long long)j * a_rs + k));
        const __nv_bfloat16* ae = reinterpret_cast<const __nv_bfloat16*>(&ar);
        float av[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) av[e] = __bfloat162float(ae[e]);
#pragma unroll
        for (int i = 0; i < ROWS; ++i) {
          float sacc = 0.f;
#pragma unroll
          for (int e = 0; e < 8; ++e) sacc += xv[i][e] * av[e];
          acc[i][j] += sacc;
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < ROWS; ++i)
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const float v = warp_sum(acc[i][j]);
      if (lane == 0) red[warp][i * R + j] = v;
    }
  __syncthreads();
  for (int o = threadIdx.x; o < ROWS * R; o += 256) {
    const int i = o / R, j = o - i * R;
    if (m0 + i < M && j < r) {
      float v = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) v += red[w][o];
      T[(long long)(m0 + i) * t_rs + j] = __float2bfloat16_rn(v * alpha);
    }
  }
}

// ---- wgrad: grid (column blocks, row slices); a thread owns COLS consecutive columns and R x COLS accumulators,
// walks the rows of its slice with independent vector loads of X, takes P from shared memory (broadcast) and ends
// with (vector) fp32 atomics into the gradient arena. One Philox call covers the thread's 4 columns of a row.
template <int R, int COLS>
__global__ void __launch_bounds__(256)
lora_wgrad_kernel(const __nv_bfloat16* __restrict__ P, const __nv_bfloat16* __restrict__ X, float* __restrict__ G,
                  int M, int K, int r, long long p_rs, long long x_rs, long long g_rs, float alpha,
                  unsigned long long seed, uint32_t thresh, float inv_keep, int rows_per_slice, int vec_atomics) {
  __shared__ float Ps[32 * R];
  const int k = (blockIdx.x * 256 + threadIdx.x) * COLS;
  const int m_begin = blockIdx.y * rows_per_slice;
  const int m_end = min(M, m_begin + rows_per_slice);
  float acc[R][COLS];
#pragma unroll
  for (int j = 0; j < R; ++j)
#pragma unroll
    for (int c = 0; c < COLS; ++c) acc[j][c] = 0.f;
  for (int m0 = m_begin; m0 < m_end; m0 += 32) {
    __syncthreads();
    for (int i = threadIdx.x; i < 32 * R; i += 256) {
      const int mm = i / R, j = i - mm * R;
      Ps[i] = (m0 + mm < m_end && j < r) ? __bfloat162float(P[(long long)(m0 + mm) * p_rs + j]) : 0.f;
    }
    __syncthreads();
    if (k < K) {
      const int lim = min(32, m_end - m0);
#pragma unroll 4
      for (int mm = 0; mm < lim; ++mm) {
        const __nv_bfloat16* xp = X + (long long)(m0 + mm) * x_rs + k;
        float xv[COLS];
        if constexpr (COLS == 4) {
          const uint2 raw = *reinterpret_cast<const uint2*>(xp);
          const __nv_bfloat16* xe = reinterpret_cast<const __nv_bfloat16*>(&raw);
#pragma unroll
          for (int c = 0; c < 4; ++c) xv[c] = __bfloat162float(xe[c]);
        } else {
#pragma unroll
          for (int c = 0; c < COLS; ++c) xv[c] = __bfloat162float(xp[c]);
        }
        if (thresh) {
          const unsigned long long e = (unsigned long long)(m0 + mm) * K + k;
          const uint4 b = lora_philox(seed, e >> 2);
          const uint32_t w[4] = {b.x, b.y, b.z, b.w};
          if constexpr (COLS == 4) {
#pragma unroll
            for (int c = 0; c < 4; ++c) xv[c] = w[c] >= thresh ? xv[c] * inv_keep : 0.f;
          } else {
#pragma unroll
            for (int c = 0; c < COLS; ++c) {
              const uint32_t ww = w[(e + c) & 3];   // COLS in {1, 2} never crosses a 4-element Philox block when k % COLS == 0
              xv[c] = ww >= thresh ? xv[c] * inv_keep : 0.f;
            }
          }
        }
#pragma unroll
        for (int j = 0; j < R; ++j) {
          const float pv = Ps[mm * R + j];
#pragma unroll
          for (int c = 0; c < COLS; ++c) acc[j][c] += pv * xv[c];
        }
      }
    }
  }
  if (k < K) {
#pragma unroll
    for (int j = 0; j < R; ++j) {
      if (j < r) {
        float* g = G + (long long)j * g_rs + k;
        if constexpr (COLS == 4) {
          if (vec_atomics) {
            atomicAdd(reinterpret_cast<float4*>(g),
                      make_float4(acc[j][0] * alpha, acc[j][1] * alpha, acc[j][2] * alpha, acc[j][3] * alpha));
            continue;
          }
        }
#pragma unroll
        for (int c = 0; c < COLS; ++c) atomicAdd(g + c, acc[j][c] * alpha);
      }
    }
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

namespace b200 {
// ---- down on the tensor cores (r <= 16, K % 32 == 0): the adapter rows are the 16-row A operand of mma.sync.m16n8k16, eight
// tokens the B operand; a CTA owns 8 tokens, its 8 warps split K. Lane (g, t) fetches one 16-byte vector of token g and
// of adapter rows g, g + 8 at k = 32 c + 8 t: both operands use the same lane -> k mapping, so the k order inside an
// MMA is permuted consistently. Dropped elements are zeroed in place (exact) and 1 / keep is folded into the output scale.
// Same Philox stream as the SIMT kernel above and as the backward kernels.
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(256)
lora_down_mma_kernel(const uint16_t* __restrict__ X, const uint16_t* __restrict__ A, __nv_bfloat16* __restrict__ T, int M, int K,
                     int r, long long x_rs, long long a_rs, long long t_rs, float alpha_eff, unsigned long long seed,
                     uint32_t thresh) {
  __shared__ float red[8][128];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int m0 = blockIdx.x * 8, tok = m0 + g;
  const bool tok_ok = tok < M, ra_ok = g < r, rb_ok = g + 8 < r;
  const uint16_t* xp = X + (long long)(tok_ok ? tok : 0) * x_rs;
  const uint16_t* pa = A + (long long)(ra_ok ? g : 0) * a_rs;
  const uint16_t* pb = A + (long long)(rb_ok ? g + 8 : 0) * a_rs;
  float c[4] = {0.f, 0.f, 0.f, 0.f};
  const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
  for (int ch = warp; ch < K / 32; ch += 8) {
    const int k = ch * 32 + t * 8;
    uint4 xa = tok_ok ? *reinterpret_cast<const uint4*>(xp + k) : zero;
    const uint4 wa = ra_ok ? __ldg(reinterpret_cast<const uint4*>(pa + k)) : zero;
    const uint4 wb = rb_ok ? __ldg(reinterpret_cast<const uint4*>(pb + k)) : zero;
    if (thresh && tok_ok) {
      bool kp[8];
      keep8(seed, (unsigned long long)tok * K + k, thresh, kp);
      xa.x &= (kp[0] ? 0x0000FFFFu : 0u) | (kp[1] ? 0xFFFF0000u : 0u);
      xa.y &= (kp[2] ? 0x0000FFFFu : 0u) | (kp[3] ? 0xFFFF0000u : 0u);
      xa.z &= (kp[4] ? 0x0000FFFFu : 0u) | (kp[5] ? 0xFFFF0000u : 0u);
      xa.w &= (kp[6] ? 0x0000FFFFu : 0u) | (kp[7] ? 0xFFFF0000u : 0u);
    }
    mma_bf16_16816(c, wa.x, wb.x, wa.y, wb.y, xa.x, xa.y);
    mma_bf16_16816(c, wa.z, wb.z, wa.w, wb.w, xa.z, xa.w);
  }
  // c0: (row g, token 2t), c1: (row g, token 2t+1), c2: (row g+8, token 2t), c3: (row g+8, token 2t+1)
#pragma unroll
  for (int i = 0; i < 4; ++i) red[warp][lane * 4 + i] = c[i];
  __syncthreads();
  if (threadIdx.x < 128) {
    float sum = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) sum += red[w][threadIdx.x];
    const int ln = threadIdx.x >> 2, i = threadIdx.x & 3;
    const int row = (ln >> 2) + ((i & 2) ? 8 : 0), tk = m0 + 2 * (ln & 3) + (i & 1);
    if (row < r && tk < M) T[(long long)tk * t_rs + row] = __float2bfloat16_rn(sum * alpha_eff);
  }
}

}  // namespace b200

B200_API int b200_lora_down(const void* X, const void* A, void* T, int M, int K, int r, long long x_rs,
                            long long a_rs, long long t_rs, float alpha, unsigned long long seed, float p,
                            void* stream) {
  if (M <= 0 || K % 8 || x_rs % 8 || a_rs % 8 || r < 1) return B200_ERR_BAD_SHAPE;
  uint32_t thresh; float inv_keep;
  drop_params(p, &thresh, &inv_keep);
  cudaStream_t st = (cudaStream_t)stream;
  static const bool simt = getenv("B200_LORA_DOWN_SIMT") != nullptr;
  if (!simt && r <= 16 && K % 32 == 0 && ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(A)) & 15) == 0) {
    lora_down_mma_kernel<<<(unsigned)((M + 7) / 8), 256, 0, st>>>((const uint16_t*)X, (const uint16_t*)A, (__nv_bfloat16*)T, M, K, r,
                                                                x_rs, a_rs, t_rs, alpha * inv_keep, seed, thresh);
    B200_LAUNCH_CHECK();
    return 0;
  }
#define B200_DOWN(R, ROWS) lora_down_kernel<R, ROWS><<<(unsigned)((M + ROWS - 1) / ROWS), 256, 0, st>>>( \
      (const __nv_bfloat16*)X, (const __nv_bfloat16*)A, (__nv_bfloat16*)T, M, K, r, x_rs, a_rs, t_rs, alpha, seed, thresh, inv_keep)
  if (r <= 8) B200_DOWN(8, 4);
  else if (r <= 16) B200_DOWN(16, 4);
  else if (r <= 32) B200_DOWN(32, 2);
  else if (r <= 64) B200_DOWN(64, 1);
  else return B200_ERR_TOO_LARGE;
#undef B200_DOWN
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_lora_wgrad(const void* P, const void* X, float* G, int M, int K, int r, long long p_rs,
                             long long x_rs, long long g_rs, float alpha, unsigned long long seed, float p,
                             void* stream) {
  if (M <= 0 || r < 1) return B200_ERR_BAD_SHAPE;
  uint32_t thresh; float inv_keep;
  drop_params(p, &thresh, &inv_keep);
  cudaStream_t st = (cudaStream_t)stream;
  const bool vec4 = K % 4 == 0 && x_rs % 4 == 0 && ((uintptr_t)X % 8) == 0 && r <= 16;
  const bool vec2 = K % 2 == 0;
  const int cols = vec4 ? 4 : (r <= 32 && vec2 ? 2 : 1);
  const int vec_atomics = vec4 && g_rs % 4 == 0 && ((uintptr_t)G % 16) == 0;
  const int kblocks = (K / cols + 255) / 256;
  int slices = (3 * kNumSMs + kblocks - 1) / kblocks;
  int rows_per_slice = ((M + slices - 1) / slices + 31) / 32 * 32;
  slices = (M + rows_per_slice - 1) / rows_per_slice;
  dim3 grid(kblocks, slices);
#define B200_WG(R, C) lora_wgrad_kernel<R, C><<<grid, 256, 0, st>>>((const __nv_bfloat16*)P, (const __nv_bfloat16*)X, G, M, K, r, \
      p_rs, x_rs, g_rs, alpha, seed, thresh, inv_keep, rows_per_slice, vec_atomics)
  if (r <= 8) { if (cols == 4) B200_WG(8, 4); else if (cols == 2) B200_WG(8, 2); else B200_WG(8, 1); }
  else if (r <= 16) { if (cols == 4) B200_WG(16, 4); else if (cols == 2) B200_WG(16, 2); else B200_WG(16, 1); }
  else if (r <= 32) { if (cols == 2) B200_WG(32, 2); else B200_WG(32, 1); }
  else if (r <= 64) B200_WG(64, 1);
  else return B200_ERR_TOO_LARGE;
#undef B200_WG
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
