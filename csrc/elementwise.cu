// K6 / K7 / K9 / K10 — fused element-wise kernels: per-head QK-RMSNorm + RoPE, SwiGLU, GELU,
// embedding gather / scatter-add, Philox dropout, dtype casts.
//
// Replaces chains of ATen element-wise kernels: HF Qwen3Attention's q_norm/k_norm + apply_rotary_pos_emb
// and Qwen3MLP's silu(gate)*up (used by every Fine-Tuning/*.py script), nn.GELU / nn.Dropout /
// nn.Embedding in GPTLike (LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:86-165)
// and both RoPE layouts of DeepSeekLike (…/transformer_basics/DeepSeekLike_wikitext2.py:137-163,
// …/DeepSeekLike_spare_MoE_wikitext2.py:145-174).
#include "common.cuh"
#include "../include/b200_ops.h"

namespace b200 {

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (counter-based; the mask of element i depends only on (seed, offset + i / 4))
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32(uint2 key, uint4 ctr) {
  constexpr uint32_t kA = 0xD2511F53u, kB = 0xCD9E8D57u, kW0 = 0x9E3779B9u, kW1 = 0xBB67AE85u;
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
// keep-mask for 4 consecutive elements starting at element index 4*q
__device__ __forceinline__ uint4 dropout_bits(unsigned long long seed, unsigned long long q) {
  return philox4x32(make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)),
                    make_uint4((uint32_t)q, (uint32_t)(q >> 32), 0u, 0u));
}
__device__ __forceinline__ bool dropout_keep(unsigned long long seed, unsigned long long idx, uint32_t thresh) {
  uint4 b = dropout_bits(seed, idx >> 2);
  const uint32_t r = (idx & 3) == 0 ? b.x : (idx & 3) == 1 ? b.y : (idx & 3) == 2 ? b.z : b.w;
  return r >= thresh;  // P(drop) = thresh / 2^32
}

// y = keep ? x / (1-p) : 0 ; the backward is the same kernel applied to dy (same seed).
template <typename T>
__global__ void dropout_kernel(const T* __restrict__ x, T* __restrict__ y, long long n, unsigned long long seed,
                               uint32_t thresh, float inv_keep, const unsigned long long* __restrict__ seed_off) {
  const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long i0 = q * 4;
  if (i0 >= n) return;
  if (seed_off) seed += *seed_off;          // CUDA-graph replay: the per-step part of the seed lives in device memory
  const uint4 b = dropout_bits(seed, (unsigned long long)q);
  const uint32_t r[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (i0 + j < n) y[i0 + j] = from_f32<T>(r[j] >= thresh ? to_f32<T>(x[i0 + j]) * inv_keep : 0.f);
}

// ---------------------------------------------------------------------------------------------
// QK-RMSNorm + RoPE. One warp per (token, head); head_dim in {64, 96, 128} (<= 128, multiple of 32
// would be enough: each lane owns head_dim/32 elements of each half... we use the generic pairing).
//   rotate-half layout (HF Qwen3):   pair = (x[i], x[i + D/2])
//   interleaved layout (DeepSeekLike): pair = (x[2i], x[2i+1])
//   out_a = a*cos - b*sin ; out_b = b*cos + a*sin        (backward: sin -> -sin, then norm-bwd)
// src/dst: [tokens, n_heads_total, D] slices of the fused qkv projection (row stride given).
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
qknorm_rope_fwd_kernel(const T* __restrict__ src, T* __restrict__ dst, const T* __restrict__ qw,
                       const T* __restrict__ kw, const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                       const int* __restrict__ pos, int tokens, int n_q, int n_k, int n_v, int D, long long src_rs,
                       long long dst_rs, int interleaved, float eps, float* __restrict__ rstd_out) {
  pdl_launch_dependents();
  pdl_wait();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int heads = n_q + n_k + n_v;
  if (warp >= tokens * heads) return;
  const int tok = warp / heads, head = warp - tok * heads;
  const T* s = src + (long long)tok * src_rs + (long long)head * D;
  T* d = dst + (long long)tok * dst_rs + (long long)head * D;
  if (head >= n_q + n_k) {  // v heads ride along unchanged so the consumer reads one tensor
    for (int i = lane; i < D; i += 32) d[i] = s[i];
    return;
  }
  const T* w = head < n_q ? qw : kw;
  const int half = D >> 1;
  const int p = pos[tok];
  // each lane handles pairs i = lane, lane+32, ... (< half)
  float a[2], b[2];
  float ss = 0.f;
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int i = lane + t * 32;
    if (i < half) {
      const int ia = interleaved ? 2 * i : i, ib = interleaved ? 2 * i + 1 : i + half;
      a[t] = to_f32<T>(s[ia]);
      b[t] = to_f32<T>(s[ib]);
      ss += a[t] * a[t] + b[t] * b[t];
    }
  }
  float rstd = 1.f;
  if (w != nullptr) {
    ss = warp_sum(ss);
    rstd = rsqrtf(ss / D + eps);
    if (rstd_out && lane == 0) rstd_out[warp] = rstd;
  }
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int i = lane + t * 32;
    if (i < half) {
      const int ia = interleaved ? 2 * i : i, ib = interleaved ? 2 * i + 1 : i + half;
      float xa = a[t], xb = b[t];
      if (w != nullptr) {
        // HF: weight * (x * rstd).to(dtype)
        xa = to_f32<T>(from_f32<T>(xa * rstd)) * to_f32<T>(w[ia]);
        xb = to_f32<T>(from_f32<T>(xb * rstd)) * to_f32<T>(w[ib]);
        xa = to_f32<T>(from_f32<T>(xa));
        xb = to_f32<T>(from_f32<T>(xb));
      }
      const float c = cos_t[(long long)p * half + i], sn = sin_t[(long long)p * half + i];
      d[ia] = from_f32<T>(xa * c - xb * sn);
      d[ib] = from_f32<T>(xb * c + xa * sn);
    }
  }
}

// backward: g = d(out); un-rotate, then RMSNorm backward against the raw (pre-norm) input.
// dwq/dwk (fp32 [D], optional) receive the norm-weight gradients via atomics.
template <typename T>
__global__ void __launch_bounds__(256)
qknorm_rope_bwd_kernel(const T* __restrict__ g_in, const T* __restrict__ raw, T* __restrict__ g_out,
                       const T* __restrict__ qw, const T* __restrict__ kw, const float* __restrict__ cos_t,
                       const float* __restrict__ sin_t, const int* __restrict__ pos, int tokens, int n_q, int n_k,
                       int n_v, int D, long long g_rs, long long raw_rs, long long out_rs, int interleaved, float eps,
                       float* __restrict__ dwq, float* __restrict__ dwk) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int heads = n_q + n_k + n_v;
  if (warp >= tokens * heads) return;
  const int tok = warp / heads, head = warp - tok * heads;
  const T* g = g_in + (long long)tok * g_rs + (long long)head * D;
  const T* x = raw + (long long)tok * raw_rs + (long long)head * D;
  T* o = g_out + (long long)tok * out_rs + (long long)head * D;
  if (head >= n_q + n_k) {
    for (int i = lane; i < D; i += 32) o[i] = g[i];
    return;
  }
  const T* w = head < n_q ? qw : kw;
  float* dw = head < n_q ? dwq : dwk;
  const int half = D >> 1;
  const int p = pos[tok];
  float ga[2], gb[2], xa[2], xb[2];
  float ss = 0.f;
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int i = lane + t * 32;
    if (i < half) {
      const int ia = interleaved ? 2 * i : i, ib = interleaved ? 2 * i + 1 : i + half;
      const float c = cos_t[(long long)p * half + i], sn = sin_t[(long long)p * half + i];
      const float da = to_f32<T>(g[ia]), db = to_f32<T>(g[ib]);
      ga[t] = da * c + db * sn;   // transpose of the rotation
      gb[t] = db * c - da * sn;
      if (w != nullptr) {
        xa[t] = to_f32<T>(x[ia]);
        xb[t] = to_f32<T>(x[ib]);
        ss += xa[t] * xa[t] + xb[t] * xb[t];
      }
    }
  }
  if (w == nullptr) {
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      const int i = lane + t * 32;
      if (i < half) {
        const int ia = interleaved ? 2 * i : i, ib = interleaved ? 2 * i + 1 : i + half;
        o[ia] = from_f32<T>(ga[t]);
        o[ib] = from_f32<T>(gb[t]);
      }
    }
    return;
  }
  ss = warp_sum(ss);
  const float rstd = rsqrtf(ss / D + eps);
  float sgx = 0.f;
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int i = lane + t * 32;
    if (i < half) {
      const int ia = interleaved ? 2 * i : i, ib = interleaved ? 2 * i + 1 : i + half;
      const float wa = to_f32<T>(w[ia]), wb = to_f32<T>(w[ib]);
      sgx += ga[t] * wa * xa[t] * rstd + gb[t] * wb * xb[t] * rstd;
      if (dw) {
        atomicAdd(dw + ia, ga[t] * xa[t] * rstd);
        atomicAdd(dw + ib, gb[t] * xb[t] * rstd);
      }
    }
  }
  sgx = warp_sum(sgx) / D;
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int i = lane + t * 32;
    if (i < half) {
      const int ia = interleaved ? 2 * i : i, ib = interleaved ? 2 * i + 1 : i + half;
      const float wa = to_f32<T>(w[ia]), wb = to_f32<T>(w[ib]);
      o[ia] = from_f32<T>(rstd * (ga[t] * wa - xa[t] * rstd * sgx));
      o[ib] = from_f32<T>(rstd * (gb[t] * wb - xb[t] * rstd * sgx));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// SwiGLU on the fused gate|up projection: gu = [rows, 2F] (gate first), h = silu(g) * u
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void swiglu_fwd_kernel(const T* __restrict__ gu, T* __restrict__ h, long long rows, int F) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int V = 16 / sizeof(T);
  const long long idx = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * V;
  if (idx >= rows * F) return;
  const long long r = idx / F;
  const int c = (int)(idx - r * F);
  uint4 gr = *reinterpret_cast<const uint4*>(gu + r * 2 * F + c);
  uint4 ur = *reinterpret_cast<const uint4*>(gu + r * 2 * F + F + c);
  const T* g = reinterpret_cast<const T*>(&gr);
  const T* u = reinterpret_cast<const T*>(&ur);
  uint4 out;
  T* o = reinterpret_cast<T*>(&out);
#pragma unroll
  for (int j = 0; j < V; ++j) {
    const float gv = to_f32<T>(g[j]);
    const float s = gv / (1.f + __expf(-gv));
    o[j] = from_f32<T>(to_f32<T>(from_f32<T>(s)) * to_f32<T>(u[j]));
  }
  *reinterpret_cast<uint4*>(h + idx) = out;
}
template <typename T>
__global__ void swiglu_bwd_kernel(const T* __restrict__ dh, const T* __restrict__ gu, T* __restrict__ dgu,
                                  long long rows, int F) {
  constexpr int V = 16 / sizeof(T);
  const long long idx = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * V;
  if (idx >= rows * F) return;
  const long long r = idx / F;
  const int c = (int)(idx - r * F);
  uint4 gr = *reinterpret_cast<const uint4*>(gu + r * 2 * F + c);
  uint4 ur = *reinterpret_cast<const uint4*>(gu + r * 2 * F + F + c);
  uint4 dr = *reinterpret_cast<const uint4*>(dh + idx);
  const T* g = reinterpret_cast<const T*>(&gr);
  const T* u = reinterpret_cast<const T*>(&ur);
  const T* d = reinterpret_cast<const T*>(&dr);
  uint4 og, ou;
  T* pg = reinterpret_cast<T*>(&og);
  T* pu = reinterpret_cast<T*>(&ou);
#pragma unroll
  for (int j = 0; j < V; ++j) {
    const float gv = to_f32<T>(g[j]), uv = to_f32<T>(u[j]), dv = to_f32<T>(d[j]);
    const float sig = 1.f / (1.f + __expf(-gv));
    const float silu = gv * sig;
    pg[j] = from_f32<T>(dv * uv * (sig + silu * (1.f - sig)));
    pu[j] = from_f32<T>(dv * silu);
  }
  *reinterpret_cast<uint4*>(dgu + r * 2 * F + c) = og;
  *reinterpret_cast<uint4*>(dgu + r * 2 * F + F + c) = ou;
}

// ---------------------------------------------------------------------------------------------
// GELU (exact, erf) on [n] elements
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void gelu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, long long n) {
  constexpr int V = 16 / sizeof(T);
  const long long idx = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * V;
  if (idx >= n) return;
  uint4 xr = *reinterpret_cast<const uint4*>(x + idx);
  const T* xv = reinterpret_cast<const T*>(&xr);
  uint4 out;
  T* o = reinterpret_cast<T*>(&out);
#pragma unroll
  for (int j = 0; j < V; ++j) {
    const float v = to_f32<T>(xv[j]);
    o[j] = from_f32<T>(0.5f * v * (1.f + erff(v * 0.70710678118654752f)));
  }
  *reinterpret_cast<uint4*>(y + idx) = out;
}
template <typename T>
__global__ void gelu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, T* __restrict__ dx, long long n) {
  constexpr int V = 16 / sizeof(T);
  const long long idx = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * V;
  if (idx >= n) return;
  uint4 xr = *reinterpret_cast<const uint4*>(x + idx);
  uint4 dr = *reinterpret_cast<const uint4*>(dy + idx);
  const T* xv = reinterpret_cast<const T*>(&xr);
  const T* dv = reinterpret_cast<const T*>(&dr);
  uint4 out;
  T* o = reinterpret_cast<T*>(&out);
#pragma unroll
  for (int j = 0; j < V; ++j) {
    const float v = to_f32<T>(xv[j]);
    const float cdf = 0.5f * (1.f + erff(v * 0.70710678118654752f));
    const float pdf = 0.3989422804014327f * __expf(-0.5f * v * v);
    o[j] = from_f32<T>(to_f32<T>(dv[j]) * (cdf + v * pdf));
  }
  *reinterpret_cast<uint4*>(dx + idx) = out;
}

// ---------------------------------------------------------------------------------------------
// Embedding: out[t, :] = table[ids[t], :] (+ pos_table[t % period, :]); grad scatter-add (fp32 atomics)
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void embedding_fwd_kernel(const long long* __restrict__ ids, const T* __restrict__ table,
                                     const T* __restrict__ pos_table, T* __restrict__ out, long long tokens, int dim,
                                     int period, float scale) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int V = 16 / sizeof(T);
  const long long t = blockIdx.x;
  const long long id = ids[t];
  const T* src = table + id * dim;
  const T* ps = pos_table ? pos_table + (t % period) * dim : nullptr;
  for (int i = threadIdx.x * V; i < dim; i += blockDim.x * V) {
    float v[V];
    uint4 raw = *reinterpret_cast<const uint4*>(src + i);
    const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
    for (int j = 0; j < V; ++j) v[j] = to_f32<T>(e[j]) * scale;
    if (ps) {
      uint4 pr = *reinterpret_cast<const uint4*>(ps + i);
      const T* pe = reinterpret_cast<const T*>(&pr);
#pragma unroll
      for (int j = 0; j < V; ++j) v[j] += to_f32<T>(pe[j]);
    }
    uint4 o;
    T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
    for (int j = 0; j < V; ++j) oe[j] = from_f32<T>(v[j]);
    *reinterpret_cast<uint4*>(out + t * dim + i) = o;
  }
}
template <typename T>
__global__ void embedding_bwd_kernel(const long long* __restrict__ ids, const T* __restrict__ dout,
                                     float* __restrict__ dtable, long long tokens, int dim, float scale) {
  const long long t = blockIdx.x;
  const long long id = ids[t];
  for (int i = threadIdx.x; i < dim; i += blockDim.x)
    atomicAdd(dtable + id * dim + i, to_f32<T>(dout[t * dim + i]) * scale);
}

// ---------------------------------------------------------------------------------------------
// cast / scale: y = x * scale with dtype conversion (bf16 <-> fp32 master copies)
// ---------------------------------------------------------------------------------------------
template <typename TI, typename TO>
__global__ void cast_kernel(const TI* __restrict__ x, TO* __restrict__ y, long long n, float scale) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) y[i] = from_f32<TO>(to_f32<TI>(x[i]) * scale);
}

}  // namespace b200
using namespace b200;

#define B200_DISPATCH_T(dtype, ...)                                   \
  if (dtype == B200_DT_BF16) { using T = __nv_bfloat16; __VA_ARGS__; } \
  else if (dtype == B200_DT_FP16) { using T = __half; __VA_ARGS__; }   \
  else if (dtype == B200_DT_FP32) { using T = float; __VA_ARGS__; }    \
  else return B200_ERR_UNSUPPORTED;

static const unsigned long long* g_graph_seed_offset = nullptr;
const unsigned long long* b200::b200_graph_seed_offset() { return g_graph_seed_offset; }

// seed_offset_dev: device pointer to one uint64 added to the seed argument of every dropout-drawing kernel launched from
// now on (b200_dropout, the flash-attention kernels with dropout), or NULL to switch it off. See common.cuh.
B200_API int b200_set_graph_state(const void* seed_offset_dev) {
  g_graph_seed_offset = reinterpret_cast<const unsigned long long*>(seed_offset_dev);
  return 0;
}

B200_API int b200_dropout(const void* x, void* y, long long n, unsigned long long seed, float p, int dtype,
                          void* stream) {
  if (n <= 0) return 0;
  const uint32_t thresh = p >= 1.f ? 0xFFFFFFFFu : (uint32_t)((double)p * 4294967296.0);
  const float inv_keep = p >= 1.f ? 0.f : 1.f / (1.f - p);
  const long long q = (n + 3) / 4;
  const int tb = 256;
  const unsigned grid = (unsigned)((q + tb - 1) / tb);
  B200_DISPATCH_T(dtype, dropout_kernel<T><<<grid, tb, 0, (cudaStream_t)stream>>>((const T*)x, (T*)y, n, seed, thresh, inv_keep,
                                                                                   b200_graph_seed_offset()));
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_qknorm_rope_fwd(const void* src, void* dst, const void* qw, const void* kw, const float* cos_t,
                                  const float* sin_t, const int* pos, int tokens, int n_q, int n_k, int n_v, int D,
                                  long long src_rs, long long dst_rs, int interleaved, float eps, float* rstd_out,
                                  int dtype, void* stream) {
  if (D > 128 || (D & 1) || tokens <= 0) return B200_ERR_BAD_SHAPE;
  const long long warps = (long long)tokens * (n_q + n_k + n_v);
  const int tb = 256;
  const unsigned grid = (unsigned)((warps * 32 + tb - 1) / tb);
  B200_DISPATCH_T(dtype, launch_pdl(qknorm_rope_fwd_kernel<T>, grid, tb, 0, (cudaStream_t)stream, 
      (const T*)src, (T*)dst, (const T*)qw, (const T*)kw, cos_t, sin_t, pos, tokens, n_q, n_k, n_v, D, src_rs, dst_rs,
      interleaved, eps, rstd_out));
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_qknorm_rope_bwd(const void* g_in, const void* raw, void* g_out, const void* qw, const void* kw,
                                  const float* cos_t, const float* sin_t, const int* pos, int tokens, int n_q, int n_k,
                                  int n_v, int D, long long g_rs, long long raw_rs, long long out_rs, int interleaved,
                                  float eps, float* dwq, float* dwk, int dtype, void* stream) {
  if (D > 128 || (D & 1) || tokens <= 0) return B200_ERR_BAD_SHAPE;
  const long long warps = (long long)tokens * (n_q + n_k + n_v);
  const int tb = 256;
  const unsigned grid = (unsigned)((warps * 32 + tb - 1) / tb);
  B200_DISPATCH_T(dtype, qknorm_rope_bwd_kernel<T><<<grid, tb, 0, (cudaStream_t)stream>>>(
      (const T*)g_in, (const T*)raw, (T*)g_out, (const T*)qw, (const T*)kw, cos_t, sin_t, pos, tokens, n_q, n_k, n_v, D,
      g_rs, raw_rs, out_rs, interleaved, eps, dwq, dwk));
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_swiglu_fwd(const void* gu, void* h, long long rows, int F, int dtype, void* stream) {
  const int V = dtype == B200_DT_FP32 ? 4 : 8;
  if (F % V) return B200_ERR_BAD_SHAPE;
  const long long nvec = rows * F / V;
  const int tb = 256;
  const unsigned grid = (unsigned)((nvec + tb - 1) / tb);
  B200_DISPATCH_T(dtype, launch_pdl(swiglu_fwd_kernel<T>, grid, tb, 0, (cudaStream_t)stream, (const T*)gu, (T*)h, rows, F));
  B200_LAUNCH_CHECK();
  return 0;
}
B200_API int b200_swiglu_bwd(const void* dh, const void* gu, void* dgu, long long rows, int F, int dtype,
                             void* stream) {
  const int V = dtype == B200_DT_FP32 ? 4 : 8;
  if (F % V) return B200_ERR_BAD_SHAPE;
  const long long nvec = rows * F / V;
  const int tb = 256;
  const unsigned grid = (unsigned)((nvec + tb - 1) / tb);
  B200_DISPATCH_T(dtype, swiglu_bwd_kernel<T><<<grid, tb, 0, (cudaStream_t)stream>>>((const T*)dh, (const T*)gu, (T*)dgu, rows, F));
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_gelu_fwd(const void* x, void* y, long long n, int dtype, void* stream) {
  const int V = dtype == B200_DT_FP32 ? 4 : 8;
  if (n % V) return B200_ERR_BAD_SHAPE;
  const int tb = 256;
  const unsigned grid = (unsigned)((n / V + tb - 1) / tb);
  B200_DISPATCH_T(dtype, gelu_fwd_kernel<T><<<grid, tb, 0, (cudaStream_t)stream>>>((const T*)x, (T*)y, n));
  B200_LAUNCH_CHECK();
  return 0;
}
B200_API int b200_gelu_bwd(const void* dy, const void* x, void* dx, long long n, int dtype, void* stream) {
  const int V = dtype == B200_DT_FP32 ? 4 : 8;
  if (n % V) return B200_ERR_BAD_SHAPE;
  const int tb = 256;
  const unsigned grid = (unsigned)((n / V + tb - 1) / tb);
  B200_DISPATCH_T(dtype, gelu_bwd_kernel<T><<<grid, tb, 0, (cudaStream_t)stream>>>((const T*)dy, (const T*)x, (T*)dx, n));
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_embedding_fwd(const long long* ids, const void* table, const void* pos_table, void* out,
                                long long tokens, int dim, int period, float scale, int dtype, void* stream) {
  const int V = dtype == B200_DT_FP32 ? 4 : 8;
  if (dim % V || tokens <= 0) return B200_ERR_BAD_SHAPE;
  int tb = dim / V;
  tb = tb > 256 ? 256 : (tb < 32 ? 32 : (tb + 31) / 32 * 32);
  B200_DISPATCH_T(dtype, launch_pdl(embedding_fwd_kernel<T>, (unsigned)tokens, tb, 0, (cudaStream_t)stream, 
      ids, (const T*)table, (const T*)pos_table, (T*)out, tokens, dim, period > 0 ? period : 1, scale));
  B200_LAUNCH_CHECK();
  return 0;
}
B200_API int b200_embedding_bwd(const long long* ids, const void* dout, float* dtable, long long tokens, int dim,
                                float scale, int dtype, void* stream) {
  if (tokens <= 0) return 0;
  B200_DISPATCH_T(dtype, embedding_bwd_kernel<T><<<(unsigned)tokens, 256, 0, (cudaStream_t)stream>>>(
      ids, (const T*)dout, dtable, tokens, dim, scale));
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_cast(const void* x, void* y, long long n, float scale, int in_dtype, int out_dtype, void* stream) {
  if (n <= 0) return 0;
  const int tb = 256;
  long long g = (n + tb - 1) / tb;
  const unsigned grid = (unsigned)(g > 148 * 16 ? 148 * 16 : g);
  cudaStream_t st = (cudaStream_t)stream;
#define B200_CAST(TI, TO) cast_kernel<TI, TO><<<grid, tb, 0, st>>>((const TI*)x, (TO*)y, n, scale)
  if (in_dtype == B200_DT_FP32 && out_dtype == B200_DT_BF16) B200_CAST(float, __nv_bfloat16);
  else if (in_dtype == B200_DT_BF16 && out_dtype == B200_DT_FP32) B200_CAST(__nv_bfloat16, float);
  else if (in_dtype == B200_DT_FP32 && out_dtype == B200_DT_FP16) B200_CAST(float, __half);
  else if (in_dtype == B200_DT_FP16 && out_dtype == B200_DT_FP32) B200_CAST(__half, float);
  else if (in_dtype == B200_DT_FP32 && out_dtype == B200_DT_FP32) B200_CAST(float, float);
  else if (in_dtype == B200_DT_BF16 && out_dtype == B200_DT_BF16) B200_CAST(__nv_bfloat16, __nv_bfloat16);
  else return B200_ERR_UNSUPPORTED;
#undef B200_CAST
  B200_LAUNCH_CHECK();
  return 0;
}

// y(fp32) += float(x) * scale, then x = 0 when clear_x: folds a low-precision gradient arena into the fp32
// main-gradient arena in one pass (mixed-precision training; see utils/arena.py MixedArena).
namespace b200 {
template <typename T>
__global__ void fold_grad_kernel(T* __restrict__ x, float* __restrict__ y, long long n, float scale, int clear_x) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const float v = to_f32<T>(x[i]);
    if (v != 0.f) {
      y[i] += v * scale;
      if (clear_x) x[i] = from_f32<T>(0.f);
    }
  }
}
}  // namespace b200

B200_API int b200_fold_grad(void* x, float* y, long long n, float scale, int clear_x, int dtype, void* stream) {
  if (n <= 0) return 0;
  const int tb = 256;
  long long g = (n + tb - 1) / tb;
  const unsigned grid = (unsigned)(g > 148 * 16 ? 148 * 16 : g);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_DT_BF16) b200::fold_grad_kernel<__nv_bfloat16><<<grid, tb, 0, st>>>((__nv_bfloat16*)x, y, n, scale, clear_x);
  else if (dtype == B200_DT_FP16) b200::fold_grad_kernel<__half><<<grid, tb, 0, st>>>((__half*)x, y, n, scale, clear_x);
  else return B200_ERR_UNSUPPORTED;
  B200_LAUNCH_CHECK();
  return 0;
}
