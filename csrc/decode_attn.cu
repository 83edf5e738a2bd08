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

// Where position t of a sequence lives. Slot pool (tbl == nullptr): the sequence owns row `slot` of a [slots, Hkv, max_len, D]
// cache. Paged (tbl = the sequence's row of the block table): the cache is [blocks, Hkv, block, D] and logical block
// t / block maps to physical block tbl[t / block] (block = 1 << bshift) - vLLM's PagedAttention layout
// (LLM_on_Kubernetes/Inference_Platfrom/README.md:1504-1509 serves through it).
struct KvAddr {
  const int* tbl;
  long long slot_base;   // slot pool: (slot * Hkv + kvh) * max_len * D
  long long head_off;    // paged: kvh * block * D
  long long blk_stride;  // paged: Hkv * block * D
  int bshift, bmask, D;
  __device__ __forceinline__ long long row(int t) const {
    if (tbl == nullptr) return slot_base + (long long)t * D;
    return (long long)tbl[t >> bshift] * blk_stride + head_off + (long long)(t & bmask) * D;
  }
};
__device__ __forceinline__ KvAddr make_kv_addr(const int* block_table, int table_stride, int bshift, long long slot, int Hkv,
                                               int kvh, int max_len, int D) {
  KvAddr a;
  a.tbl = block_table ? block_table + slot * table_stride : nullptr;
  a.slot_base = (slot * Hkv + kvh) * (long long)max_len * D;
  a.bshift = bshift; a.bmask = (1 << bshift) - 1; a.D = D;
  a.head_off = (long long)kvh * (1 << bshift) * D;
  a.blk_stride = (long long)Hkv * (1 << bshift) * D;
  return a;
}

// qkv: [B, (H + 2 Hkv) * D] (post-RoPE) for the new token; cache_k / cache_v: [B, Hkv, max_len, D];
// pos[b] = index the new token is written at (= number of tokens already cached); slots[b] (optional) = cache row
// of batch row b; out: [B, H * D]
template <typename T, int D>
__global__ void __launch_bounds__(128)
decode_attn_kernel(const T* __restrict__ qkv, T* __restrict__ cache_k, T* __restrict__ cache_v,
                   const int* __restrict__ pos, const int* __restrict__ slots, T* __restrict__ out, int H, int Hkv, int max_len,
                   float scale, int window, const int* __restrict__ block_table, int table_stride, int bshift) {
  constexpr int PER = D / 32;  // elements per lane
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int kvh = h / (H / Hkv);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t_new = pos[b];
  const long long row = (long long)b * (H + 2 * Hkv) * D;
  const long long slot = slots ? slots[b] : b;  // continuous batching: batch row -> cache slot
  const KvAddr ka = make_kv_addr(block_table, table_stride, bshift, slot, Hkv, kvh, max_len, D);
  T* kc = cache_k;
  T* vc = cache_v;
  // append the new K / V row once per kv head (the first query head of the group does it)
  if (h % (H / Hkv) == 0 && warp == 0) {
    const long long rn = ka.row(t_new);
    for (int i = lane; i < D; i += 32) {
      kc[rn + i] = qkv[row + (long long)(H + kvh) * D + i];
      vc[rn + i] = qkv[row + (long long)(H + Hkv + kvh) * D + i];
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
        const long long rt = t == t_new ? 0 : ka.row(t);
        load_row<T, PER>(t == t_new ? knew : kc + rt + lane * PER, kx[u]);
        load_row<T, PER>(t == t_new ? vnew : vc + rt + lane * PER, vx[u]);
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

// ------------------------------------------------------------------------------------------------
// Split ("flash-decoding") variant for serving batches: one CTA per (sequence, kv head, split) serves all G query heads
// of the group, so K and V are read once per group instead of once per query head, and the cached positions are spread
// over `nsplit` CTAs. Per 32-position block a warp works in two layouts:
//   scores : lane = position. The lane streams its own K row in 16-byte vectors and accumulates the G dot products
//            against q held in shared memory (broadcast reads) - no shuffle per position, 16 independent loads in flight;
//   values : lane = D/32 output channels. The block's probabilities go through shared memory, V rows are read coalesced.
// Each CTA leaves (max, sum, unnormalised output) per head in `part`; decode_attn_combine_kernel merges the splits.
// 16-bit types, D = 128 or 64, window-less.
template <typename T, int D, int G>
__global__ void __launch_bounds__(128)
decode_attn_split_kernel(const T* __restrict__ qkv, T* __restrict__ cache_k, T* __restrict__ cache_v,
                         const int* __restrict__ pos, const int* __restrict__ slots, float* __restrict__ part, int H, int Hkv,
                         int max_len, float scale, int nsplit, const int* __restrict__ block_table, int table_stride, int bshift) {
  constexpr int PER = D / 32, CH = D / 8;   // output channels per lane; 16-byte chunks per row
  pdl_launch_dependents();
  pdl_wait();
  __shared__ __align__(16) float qs[G][D];
  __shared__ float ps[4][G][32];
  __shared__ float sm_m[4][G], sm_l[4][G];
  __shared__ float sacc[4][G][D];
  const int sp = blockIdx.x % nsplit, kvh = (blockIdx.x / nsplit) % Hkv, b = blockIdx.x / (nsplit * Hkv);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t_new = pos[b];
  const long long row = (long long)b * (H + 2 * Hkv) * D;
  const long long slot = slots ? slots[b] : b;
  const KvAddr ka = make_kv_addr(block_table, table_stride, bshift, slot, Hkv, kvh, max_len, D);
  T* kc = cache_k;
  T* vc = cache_v;
  const T* knew = qkv + row + (long long)(H + kvh) * D;
  const T* vnew = qkv + row + (long long)(H + Hkv + kvh) * D;
  if (sp == 0 && warp == 0) {   // append the new token's K / V (read back from qkv below, never from the cache)
    const long long rn = ka.row(t_new);
    for (int i = lane; i < D; i += 32) {
      kc[rn + i] = knew[i];
      vc[rn + i] = vnew[i];
    }
  }
  for (int i = threadIdx.x; i < G * D; i += 128)
    qs[i / D][i % D] = to_f32<T>(qkv[row + (long long)(kvh * G + i / D) * D + i % D]) * scale;
  __syncthreads();
  float m[G], l[G], acc[G][PER];
#pragma unroll
  for (int h = 0; h < G; ++h) {
    m[h] = -INFINITY; l[h] = 0.f;
#pragma unroll
    for (int i = 0; i < PER; ++i) acc[h][i] = 0.f;
  }
  for (int base = (sp * 4 + warp) * 32; base <= t_new; base += nsplit * 4 * 32) {
    // ---- scores: lane = position ----
    const int t = base + lane;
    const bool valid = t <= t_new;
    const T* kr = !valid ? knew : (t == t_new ? knew : kc + ka.row(t));
    float s[G];
#pragma unroll
    for (int h = 0; h < G; ++h) s[h] = 0.f;
#pragma unroll 4
    for (int c = 0; c < CH; ++c) {
      const uint4 raw = *reinterpret_cast<const uint4*>(kr + c * 8);
      const T* ke = reinterpret_cast<const T*>(&raw);
      float kf[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) kf[e] = to_f32<T>(ke[e]);
#pragma unroll
      for (int h = 0; h < G; ++h) {
        const float4 q0 = *reinterpret_cast<const float4*>(&qs[h][c * 8]);
        const float4 q1 = *reinterpret_cast<const float4*>(&qs[h][c * 8 + 4]);
        s[h] += q0.x * kf[0] + q0.y * kf[1] + q0.z * kf[2] + q0.w * kf[3] + q1.x * kf[4] + q1.y * kf[5] + q1.z * kf[6] + q1.w * kf[7];
      }
    }
#pragma unroll
    for (int h = 0; h < G; ++h) {
      const float sv = valid ? s[h] : -INFINITY;
      const float mn = fmaxf(m[h], warp_max(sv));
      const float c = __expf(m[h] - mn);
      const float p = valid ? __expf(sv - mn) : 0.f;
      ps[warp][h][lane] = p;
      l[h] = l[h] * c + warp_sum(p);
      m[h] = mn;
#pragma unroll
      for (int i = 0; i < PER; ++i) acc[h][i] *= c;
    }
    __syncwarp();
    // ---- values: lane = output channels ----
    const int nv_ = min(32, t_new - base + 1);
#pragma unroll 4
    for (int tt = 0; tt < nv_; ++tt) {
      const int tv = base + tt;
      float vf[PER];
      load_row<T, PER>((tv == t_new ? vnew : vc + ka.row(tv)) + lane * PER, vf);
#pragma unroll
      for (int h = 0; h < G; ++h) {
        const float p = ps[warp][h][tt];
#pragma unroll
        for (int i = 0; i < PER; ++i) acc[h][i] += p * vf[i];
      }
    }
    __syncwarp();
  }
  // ---- merge the 4 warps, publish this split's partial ----
#pragma unroll
  for (int h = 0; h < G; ++h) {
    if (lane == 0) { sm_m[warp][h] = m[h]; sm_l[warp][h] = l[h]; }
#pragma unroll
    for (int i = 0; i < PER; ++i) sacc[warp][h][lane * PER + i] = acc[h][i];
  }
  __syncthreads();
  for (int o = threadIdx.x; o < G * D; o += 128) {
    const int h = o / D, d = o % D;
    const float gm = fmaxf(fmaxf(sm_m[0][h], sm_m[1][h]), fmaxf(sm_m[2][h], sm_m[3][h]));
    float gl = 0.f, go = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float wgt = sm_m[w][h] == -INFINITY ? 0.f : __expf(sm_m[w][h] - gm);
      gl += sm_l[w][h] * wgt;
      go += sacc[w][h][d] * wgt;
    }
    float* dst = part + (((long long)b * H + kvh * G + h) * nsplit + sp) * (D + 2);
    dst[2 + d] = go;
    if (d == 0) { dst[0] = gm; dst[1] = gl; }
  }
}

template <typename T, int D>
__global__ void __launch_bounds__(D)
decode_attn_combine_kernel(const float* __restrict__ part, T* __restrict__ out, int nsplit) {
  pdl_launch_dependents();
  pdl_wait();
  const long long bh = blockIdx.x;   // b * H + h
  const int d = threadIdx.x;
  const float* p = part + bh * nsplit * (D + 2);
  float gm = -INFINITY;
  for (int s = 0; s < nsplit; ++s) gm = fmaxf(gm, p[s * (D + 2)]);
  float gl = 0.f, go = 0.f;
  for (int s = 0; s < nsplit; ++s) {
    const float ms = p[s * (D + 2)];
    const float w = ms == -INFINITY ? 0.f : __expf(ms - gm);
    gl += p[s * (D + 2) + 1] * w;
    go += p[s * (D + 2) + 2 + d] * w;
  }
  out[bh * D + d] = from_f32<T>(gl > 0.f ? go / gl : 0.f);
}

}  // namespace b200
using namespace b200;

// scratch: B * H * nsplit * (D + 2) floats. Returns B200_ERR_UNSUPPORTED for shapes the split kernel does not cover
// (callers use b200_decode_attn then).
static int log2_exact(int v) {
  int s = 0;
  while ((1 << s) < v) ++s;
  return (1 << s) == v ? s : -1;
}

static int decode_attn_split_impl(const void* qkv, void* cache_k, void* cache_v, const int* pos, const int* slots, void* out,
                                  float* scratch, int B, int H, int Hkv, int D, int max_len, float scale, int nsplit,
                                  int dtype, void* stream, const int* block_table, int table_stride, int bshift) {
  if (B <= 0 || H % Hkv || nsplit < 1 || nsplit > 64) return B200_ERR_BAD_SHAPE;
  if (dtype != B200_DT_BF16 && dtype != B200_DT_FP16) return B200_ERR_UNSUPPORTED;
  const int G = H / Hkv;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = (unsigned)(B * Hkv * nsplit), grid2 = (unsigned)(B * H);
#define B200_DS(T, DD, GG)                                                                                              \
  do {                                                                                                                  \
    launch_pdl(decode_attn_split_kernel<T, DD, GG>, grid, 128, 0, st, (const T*)qkv, (T*)cache_k, (T*)cache_v, pos, slots, \
               scratch, H, Hkv, max_len, scale, nsplit, block_table, table_stride, bshift);                            \
    launch_pdl(decode_attn_combine_kernel<T, DD>, grid2, DD, 0, st, (const float*)scratch, (T*)out, nsplit);            \
  } while (0)
#define B200_DS_G(T, DD)                                                         \
  if (G == 1) B200_DS(T, DD, 1); else if (G == 2) B200_DS(T, DD, 2);              \
  else if (G == 4) B200_DS(T, DD, 4); else if (G == 5) B200_DS(T, DD, 5);         \
  else if (G == 8) B200_DS(T, DD, 8); else return B200_ERR_UNSUPPORTED
  if (dtype == B200_DT_BF16) {
    if (D == 128) { B200_DS_G(__nv_bfloat16, 128); } else if (D == 64) { B200_DS_G(__nv_bfloat16, 64); } else return B200_ERR_UNSUPPORTED;
  } else {
    if (D == 128) { B200_DS_G(__half, 128); } else if (D == 64) { B200_DS_G(__half, 64); } else return B200_ERR_UNSUPPORTED;
  }
#undef B200_DS_G
#undef B200_DS
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_decode_attn_split(const void* qkv, void* cache_k, void* cache_v, const int* pos, const int* slots, void* out,
                                    float* scratch, int B, int H, int Hkv, int D, int max_len, float scale, int nsplit,
                                    int dtype, void* stream) {
  return decode_attn_split_impl(qkv, cache_k, cache_v, pos, slots, out, scratch, B, H, Hkv, D, max_len, scale, nsplit, dtype, stream,
                                nullptr, 0, 0);
}

// Paged KV cache: cache_k / cache_v [blocks, Hkv, block_size, D]; block_table [rows, table_stride] int32 (device), row
// slots[b] (or b) lists the physical blocks of sequence b in logical order; block_size a power of two.
B200_API int b200_decode_attn_split_paged(const void* qkv, void* cache_k, void* cache_v, const int* pos, const int* slots,
                                          void* out, float* scratch, int B, int H, int Hkv, int D, const int* block_table,
                                          int table_stride, int block_size, float scale, int nsplit, int dtype, void* stream) {
  const int sh = log2_exact(block_size);
  if (!block_table || sh < 0 || table_stride < 1) return B200_ERR_BAD_SHAPE;
  return decode_attn_split_impl(qkv, cache_k, cache_v, pos, slots, out, scratch, B, H, Hkv, D, block_size, scale, nsplit, dtype,
                                stream, block_table, table_stride, sh);
}

static int decode_attn_impl(const void* qkv, void* cache_k, void* cache_v, const int* pos, const int* slots, void* out, int B, int H,
                            int Hkv, int D, int max_len, float scale, int window, int dtype, void* stream,
                            const int* block_table, int table_stride, int bshift) {
  if (B <= 0 || H % Hkv) return B200_ERR_BAD_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = (unsigned)(B * H);
#define B200_DA(T, DD) launch_pdl(decode_attn_kernel<T, DD>, grid, 128, 0, st, (const T*)qkv, (T*)cache_k, (T*)cache_v, pos, slots, (T*)out, H, Hkv, max_len, scale, window, block_table, table_stride, bshift)
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

B200_API int b200_decode_attn(const void* qkv, void* cache_k, void* cache_v, const int* pos, const int* slots, void* out, int B, int H,
                              int Hkv, int D, int max_len, float scale, int window, int dtype, void* stream) {
  return decode_attn_impl(qkv, cache_k, cache_v, pos, slots, out, B, H, Hkv, D, max_len, scale, window, dtype, stream, nullptr, 0, 0);
}

B200_API int b200_decode_attn_paged(const void* qkv, void* cache_k, void* cache_v, const int* pos, const int* slots, void* out, int B,
                                    int H, int Hkv, int D, const int* block_table, int table_stride, int block_size, float scale,
                                    int window, int dtype, void* stream) {
  const int sh = log2_exact(block_size);
  if (!block_table || sh < 0 || table_stride < 1) return B200_ERR_BAD_SHAPE;
  return decode_attn_impl(qkv, cache_k, cache_v, pos, slots, out, B, H, Hkv, D, block_size, scale, window, dtype, stream, block_table,
                          table_stride, sh);
}
