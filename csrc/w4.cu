// K2 / K14 — 4-bit weight kernels shared by NF4 (QLoRA) and GPTQ/AWQ (W4A16).
//
// One storage convention serves both families: packed[N, K/2] bytes with K contiguous, element 2j
// in the high nibble and 2j+1 in the low nibble (that IS the bitsandbytes NF4 layout; GPTQ and
// compressed-tensors checkpoints are repacked into it once at load time), one fp32 scale per
// `group` consecutive K values, an optional fp32 "zero * scale" term per group, and a 16-entry
// code book:      w[n, k] = lut[q[n, k]] * scale[n, k / group] - zs[n, k / group]
//   NF4      : lut = the NormalFloat4 table, group 64, scale = absmax, zs = 0
//   GPTQ/AWQ : lut = 0..15, group 128, scale = s, zs = zero_point * s  (symmetric: zero_point = 8)
//
// Kernels: NF4 quantise (+ nested "double quant" absmax decode), dequantise to bf16/fp16 (feeds the
// tcgen05 GEMM for large M; reads shards straight from peer GPUs when given several base pointers,
// which is the ZeRO-3 parameter all-gather fused into the first consumer), and a fused small-M GEMV
// for decode (memory-bound: 4.25 bits per weight on the wire instead of 16).
//
// Replaces bitsandbytes kQuantizeBlockwise / kDequantizeBlockwise / kgemm_4bit_inference_naive
// (Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:101-106) and the Marlin / ExLlama / AWQ kernels that
// gptqmodel and vLLM dispatch to (Quantization/GPTQModel/inference_qwen3_4b_gptq.py:29,
// Quantization/LLM-Compressor/AWQ/eval_qwen3_4b_awq.py:11-21).
#include "common.cuh"
#include "../include/b200_ops.h"

namespace b200 {

__constant__ float kNF4[16] = {-1.0f, -0.6961928009986877f, -0.5250730514526367f, -0.39491748809814453f,
                               -0.28444138169288635f, -0.18477343022823334f, -0.09105003625154495f, 0.0f,
                               0.07958029955625534f, 0.16093020141124725f, 0.24611230194568634f,
                               0.33791524171829224f, 0.44070982933044434f, 0.5626170039176941f,
                               0.7229568362236023f, 1.0f};

// nearest NF4 code for x in [-1, 1] (ties resolved like a midpoint decision tree)
__device__ __forceinline__ uint32_t nf4_encode(float x) {
  uint32_t best = 0;
  float bd = fabsf(x - kNF4[0]);
#pragma unroll
  for (int i = 1; i < 16; ++i) {
    const float d = fabsf(x - kNF4[i]);
    if (d < bd) { bd = d; best = i; }
  }
  return best;
}

// One warp per 64-value block: lane l owns values 2l and 2l+1 -> one packed byte.
template <typename T>
__global__ void __launch_bounds__(256)
nf4_quantize_kernel(const T* __restrict__ w, uint8_t* __restrict__ packed, float* __restrict__ absmax,
                    long long nblocks) {
  const long long blk = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (blk >= nblocks) return;
  const float a = to_f32<T>(w[blk * 64 + 2 * lane]);
  const float b = to_f32<T>(w[blk * 64 + 2 * lane + 1]);
  const float am = warp_max(fmaxf(fabsf(a), fabsf(b)));
  const float inv = am > 0.f ? 1.f / am : 0.f;
  const uint32_t qa = nf4_encode(a * inv), qb = nf4_encode(b * inv);
  packed[blk * 32 + lane] = (uint8_t)((qa << 4) | qb);
  if (lane == 0) absmax[blk] = am;
}

// nested quantisation of absmax (bitsandbytes "double quant"): absmax = code2[q] * absmax2[i / 256] + offset
__global__ void nf4_absmax_decode_kernel(const uint8_t* __restrict__ q, const float* __restrict__ code2,
                                         const float* __restrict__ absmax2, float offset, float* __restrict__ out,
                                         long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = code2[q[i]] * absmax2[i >> 8] + offset;
}

struct W4Shards {
  const uint8_t* packed[8];
  const float* scale[8];
  const float* zs[8];
  long long rows_per_shard;  // shard s holds rows [s * rows_per_shard, (s+1) * rows_per_shard)
  int nshards;
};

// Each thread expands 16 packed bytes (32 weights). K % 32 == 0, group % 32 == 0.
template <typename T>
__global__ void __launch_bounds__(256)
w4_dequant_kernel(W4Shards sh, T* __restrict__ out, long long N, int K, int group, const float* __restrict__ lut_g,
                  long long out_rs) {
  __shared__ float lut[16];
  if (threadIdx.x < 16) lut[threadIdx.x] = lut_g ? lut_g[threadIdx.x] : kNF4[threadIdx.x];
  __syncthreads();
  const int chunks_per_row = K / 32;
  const long long cid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (cid >= N * chunks_per_row) return;
  const long long n = cid / chunks_per_row;
  const int kc = (int)(cid - n * chunks_per_row);
  const int s = sh.nshards > 1 ? (int)(n / sh.rows_per_shard) : 0;
  const long long nl = n - (long long)s * sh.rows_per_shard;
  const uint4 raw = *reinterpret_cast<const uint4*>(sh.packed[s] + (nl * K / 2) + kc * 16);
  const int g = (kc * 32) / group;
  const int groups_per_row = K / group;
  const float sc = sh.scale[s][nl * groups_per_row + g];
  const float zs = sh.zs[s] ? sh.zs[s][nl * groups_per_row + g] : 0.f;
  const uint32_t words[4] = {raw.x, raw.y, raw.z, raw.w};
  T* o = out + n * out_rs + kc * 32;
#pragma unroll
  for (int wi = 0; wi < 4; ++wi) {
    float v[8];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const uint32_t byte = (words[wi] >> (8 * b)) & 0xFFu;
      v[2 * b] = lut[byte >> 4] * sc - zs;
      v[2 * b + 1] = lut[byte & 15u] * sc - zs;
    }
    uint4 ov;
    T* oe = reinterpret_cast<T*>(&ov);
#pragma unroll
    for (int j = 0; j < 8; ++j) oe[j] = from_f32<T>(v[j]);
    *reinterpret_cast<uint4*>(o + wi * 8) = ov;
  }
}

// Same expansion with fully coalesced traffic: lane l of a warp takes the 32-bit word l of a 128-byte run of codes (8 weights)
// and writes one 16-byte vector, so a warp reads 128 contiguous bytes and writes 512 contiguous bytes per step (the kernel
// above gives every thread a 64-byte output run, i.e. 16-byte pieces at a 64-byte stride per store instruction). Four
// independent steps per thread keep loads in flight. K % 8 == 0. Measured slower than the kernel above (see b200_w4_dequant): opt-in.
template <typename T>
__global__ void __launch_bounds__(256)
w4_dequant_coalesced_kernel(W4Shards sh, T* __restrict__ out, long long N, int K, int group, const float* __restrict__ lut_g,
                            long long out_rs) {
  __shared__ float lut[16];
  if (threadIdx.x < 16) lut[threadIdx.x] = lut_g ? lut_g[threadIdx.x] : kNF4[threadIdx.x];
  __syncthreads();
  constexpr int U = 4;
  const long long nwords = N * (long long)K / 8;
  const long long w0 = ((long long)blockIdx.x * blockDim.x + (threadIdx.x & ~31)) * U + (threadIdx.x & 31);
  const int groups_per_row = K / group;
  uint32_t code[U];
  float sc[U], zs[U];
  long long oidx[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const long long w = w0 + u * 32;
    oidx[u] = -1;
    if (w < nwords) {
      const long long e = w * 8, n = e / K;
      const int k = (int)(e - n * K);
      const int s = sh.nshards > 1 ? (int)(n / sh.rows_per_shard) : 0;
      const long long nl = n - (long long)s * sh.rows_per_shard;
      code[u] = *reinterpret_cast<const uint32_t*>(sh.packed[s] + (nl * K + k) / 2);
      sc[u] = sh.scale[s][nl * groups_per_row + k / group];
      zs[u] = sh.zs[s] ? sh.zs[s][nl * groups_per_row + k / group] : 0.f;
      oidx[u] = n * out_rs + k;
    }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    if (oidx[u] < 0) continue;
    uint4 ov;
    T* oe = reinterpret_cast<T*>(&ov);
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const uint32_t byte = (code[u] >> (8 * b)) & 0xFFu;
      oe[2 * b] = from_f32<T>(lut[byte >> 4] * sc[u] - zs[u]);
      oe[2 * b + 1] = from_f32<T>(lut[byte & 15u] * sc[u] - zs[u]);
    }
    *reinterpret_cast<uint4*>(out + oidx[u]) = ov;
  }
}

// Transposed expansion: out_t[k, n] = w[n, k] (out_t row stride out_rs elements). Used by the backward pass so that
// dgrad (dY x W) reads W^T as a K-major tcgen05 operand instead of the slower MN-major path. One CTA (128 threads)
// owns a 64(n) x 64(k) tile: each thread expands one 16-byte chunk (32 weights of one row) into a padded shared
// tile t[k][n], then the tile is written out as 64 rows of 128 contiguous bytes. N % 64 == 0, K % 64 == 0.
template <typename T>
__global__ void __launch_bounds__(128)
w4_dequant_t_kernel(W4Shards sh, T* __restrict__ out, long long N, int K, int group, const float* __restrict__ lut_g,
                    long long out_rs) {
  constexpr int TS = 72;  // padded row stride (elements): 144 B keeps 16-byte alignment and staggers banks
  __shared__ float lut[16];
  __shared__ __align__(16) T tile[64 * TS];
  if (threadIdx.x < 16) lut[threadIdx.x] = lut_g ? lut_g[threadIdx.x] : kNF4[threadIdx.x];
  __syncthreads();
  const int k_tiles = K / 64;
  const long long n0 = (long long)(blockIdx.x / k_tiles) * 64;
  const int k0 = (int)(blockIdx.x % k_tiles) * 64;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  {
    const int nl_t = (warp & 1) * 32 + lane, c = warp >> 1;
    const long long n = n0 + nl_t;
    const int s = sh.nshards > 1 ? (int)(n / sh.rows_per_shard) : 0;
    const long long nl = n - (long long)s * sh.rows_per_shard;
    const int kc = k0 / 32 + c;
    const uint4 raw = *reinterpret_cast<const uint4*>(sh.packed[s] + (nl * K / 2) + kc * 16);
    const int g = (kc * 32) / group;
    const int groups_per_row = K / group;
    const float sc = sh.scale[s][nl * groups_per_row + g];
    const float zs = sh.zs[s] ? sh.zs[s][nl * groups_per_row + g] : 0.f;
    const uint32_t words[4] = {raw.x, raw.y, raw.z, raw.w};
    T* col = tile + (c * 32) * TS + nl_t;
#pragma unroll
    for (int wi = 0; wi < 4; ++wi) {
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const uint32_t byte = (words[wi] >> (8 * b)) & 0xFFu;
        col[(wi * 8 + 2 * b) * TS] = from_f32<T>(lut[byte >> 4] * sc - zs);
        col[(wi * 8 + 2 * b + 1) * TS] = from_f32<T>(lut[byte & 15u] * sc - zs);
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int it = 0; it < 4; ++it) {
    const int k = it * 16 + warp * 4 + (lane >> 3), seg = lane & 7;
    const uint4 v = *reinterpret_cast<const uint4*>(tile + k * TS + seg * 8);
    *reinterpret_cast<uint4*>(out + (long long)(k0 + k) * out_rs + n0 + seg * 8) = v;
  }
}

// Decode GEMV on the tensor cores: y[m, n] = sum_k x[m, k] * w[n, k] for M <= 8 tokens, reading the weights once at
// 4 bits. A CTA owns 16 output rows; its 8 warps take interleaved 128-wide k chunks and each issues
// mma.sync.m16n8k16 with the 16 weight rows as the A operand and the (<= 8) tokens as the 8-wide B operand, so the
// per-token cost is hidden inside the MMA instead of multiplying the ALU and shared-memory work by M. The k index
// of a tensor-core dot product may be permuted freely as long as A and B agree, which lets every lane fetch its
// operands as whole 16-byte vectors: lane (g, t) loads the 32 consecutive weights [k + 32t, k + 32t + 32) of rows g and
// g + 8 (16 packed bytes each) and the same 32 activations of token g, and MMA j of the chunk consumes elements
// 4j .. 4j+3 of those. Codes become 16-bit pairs through a byte -> (lut[hi], lut[lo]) table kept once per lane in shared
// memory (bank = lane: conflict-free for any byte values), scaled by one packed FMA. Memory-bound by construction:
// 0.5 B per weight from HBM, x from L1.
template <bool BF16>
__global__ void __launch_bounds__(256)
w4_gemv_mma_kernel(const uint16_t* __restrict__ x, const uint8_t* __restrict__ packed, const float* __restrict__ scale,
                   const float* __restrict__ zsp, const float* __restrict__ lut_g, const uint16_t* __restrict__ bias,
                   uint16_t* __restrict__ y, int M, long long N, int K, int group, long long x_rs, long long y_rs) {
  __shared__ uint32_t table[256 * 32];
  __shared__ float red[8][128];
  pdl_launch_dependents();   // the table below depends on no earlier kernel: built while the previous kernel drains
  for (int i = threadIdx.x; i < 256 * 32; i += 256) {
    const int byte = i >> 5;
    const float lo = lut_g ? lut_g[byte >> 4] : kNF4[byte >> 4];
    const float hi = lut_g ? lut_g[byte & 15] : kNF4[byte & 15];
    uint32_t v;
    if (BF16) v = pack_bf16x2(lo, hi);
    else { const __half2 h = __floats2half2_rn(lo, hi); v = *reinterpret_cast<const uint32_t*>(&h); }
    table[i] = v;
  }
  __syncthreads();
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int gpr = K / group;
  const uint32_t* tl = table + lane;
  const int nchunks = K / 128;
  const long long ntiles = (N + 15) / 16;
  // persistent over 16-row tiles (the table is built once per CTA); the next chunk's codes and scales are fetched while
  // the current chunk goes through the table and the tensor core
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long row0 = tile * 16;
    const long long rA = min(row0 + g, N - 1), rB = min(row0 + g + 8, N - 1);
    float c[4] = {0.f, 0.f, 0.f, 0.f};
    uint4 wa, wb;
    float sa, sb, za, zb;
    auto fetch = [&](int ch, uint4& a_, uint4& b_, float& sa_, float& sb_, float& za_, float& zb_) {
      const int k = ch * 128 + t * 32;
      a_ = __ldg(reinterpret_cast<const uint4*>(packed + rA * (K / 2) + k / 2));
      b_ = __ldg(reinterpret_cast<const uint4*>(packed + rB * (K / 2) + k / 2));
      sa_ = __ldg(scale + rA * gpr + k / group);
      sb_ = __ldg(scale + rB * gpr + k / group);
      za_ = zsp ? __ldg(zsp + rA * gpr + k / group) : 0.f;
      zb_ = zsp ? __ldg(zsp + rB * gpr + k / group) : 0.f;
    };
    if (warp < nchunks) fetch(warp, wa, wb, sa, sb, za, zb);
    for (int ch = warp; ch < nchunks; ch += 8) {
      const int k = ch * 128 + t * 32;
      const uint4 cwa = wa, cwb = wb;
      const float csa = sa, csb = sb, cza = za, czb = zb;
      if (ch + 8 < nchunks) fetch(ch + 8, wa, wb, sa, sb, za, zb);
      uint32_t xw[16];
      if (g < M) {
        const uint4* xp = reinterpret_cast<const uint4*>(x + (long long)g * x_rs + k);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint4 v = __ldg(xp + i);
          xw[4 * i] = v.x; xw[4 * i + 1] = v.y; xw[4 * i + 2] = v.z; xw[4 * i + 3] = v.w;
        }
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) xw[i] = 0u;
      }
      uint32_t sa2, sb2, za2, zb2;
      if (BF16) { sa2 = pack_bf16x2(csa, csa); sb2 = pack_bf16x2(csb, csb); za2 = pack_bf16x2(-cza, -cza); zb2 = pack_bf16x2(-czb, -czb); }
      else {
        __half2 h;
        h = __floats2half2_rn(csa, csa); sa2 = *reinterpret_cast<uint32_t*>(&h);
        h = __floats2half2_rn(csb, csb); sb2 = *reinterpret_cast<uint32_t*>(&h);
        h = __floats2half2_rn(-cza, -cza); za2 = *reinterpret_cast<uint32_t*>(&h);
        h = __floats2half2_rn(-czb, -czb); zb2 = *reinterpret_cast<uint32_t*>(&h);
      }
      const uint32_t wA[4] = {cwa.x, cwa.y, cwa.z, cwa.w}, wB[4] = {cwb.x, cwb.y, cwb.z, cwb.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        // bytes 2j and 2j+1 of the lane's 16 packed bytes = elements 4j .. 4j+3
        const uint32_t ba0 = (wA[j >> 1] >> ((j & 1) * 16)) & 0xFFu, ba1 = (wA[j >> 1] >> ((j & 1) * 16 + 8)) & 0xFFu;
        const uint32_t bb0 = (wB[j >> 1] >> ((j & 1) * 16)) & 0xFFu, bb1 = (wB[j >> 1] >> ((j & 1) * 16 + 8)) & 0xFFu;
        uint32_t a0 = tl[ba0 << 5], a2 = tl[ba1 << 5], a1 = tl[bb0 << 5], a3 = tl[bb1 << 5];
        if (BF16) {
          asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(a0) : "r"(a0), "r"(sa2), "r"(za2));
          asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(a2) : "r"(a2), "r"(sa2), "r"(za2));
          asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(a1) : "r"(a1), "r"(sb2), "r"(zb2));
          asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(a3) : "r"(a3), "r"(sb2), "r"(zb2));
          asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                       : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                       : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(xw[2 * j]), "r"(xw[2 * j + 1]));
        } else {
          asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(a0) : "r"(a0), "r"(sa2), "r"(za2));
          asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(a2) : "r"(a2), "r"(sa2), "r"(za2));
          asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(a1) : "r"(a1), "r"(sb2), "r"(zb2));
          asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(a3) : "r"(a3), "r"(sb2), "r"(zb2));
          asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                       : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                       : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(xw[2 * j]), "r"(xw[2 * j + 1]));
        }
      }
    }
    // c0: (row g, token 2t), c1: (row g, token 2t+1), c2: (row g+8, token 2t), c3: (row g+8, token 2t+1)
#pragma unroll
    for (int i = 0; i < 4; ++i) red[warp][lane * 4 + i] = c[i];
    __syncthreads();
    if (threadIdx.x < 128) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) s += red[w][threadIdx.x];
      const int ln = threadIdx.x >> 2, i = threadIdx.x & 3;
      const long long row = row0 + (ln >> 2) + ((i & 2) ? 8 : 0);
      const int tok = 2 * (ln & 3) + (i & 1);
      if (row < N && tok < M) {
        if (bias) s += BF16 ? __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(bias + row))
                            : __half2float(*reinterpret_cast<const __half*>(bias + row));
        uint16_t o;
        if (BF16) { const __nv_bfloat16 b = __float2bfloat16_rn(s); o = *reinterpret_cast<const uint16_t*>(&b); }
        else { const __half h = __float2half_rn(s); o = *reinterpret_cast<const uint16_t*>(&h); }
        y[(long long)tok * y_rs + row] = o;
      }
    }
    __syncthreads();   // `red` is reused by the next tile
  }
}

// y[m, n] = sum_k x[m, k] * w[n, k] for m < M <= 8: one warp per output column n.
template <typename T, int kMaxM>
__global__ void __launch_bounds__(256)
w4_gemv_kernel(const T* __restrict__ x, const uint8_t* __restrict__ packed, const float* __restrict__ scale,
               const float* __restrict__ zsp, const float* __restrict__ lut_g, const T* __restrict__ bias,
               T* __restrict__ y, int M, long long N, int K, int group, long long x_rs, long long y_rs) {
  extern __shared__ uint8_t smem_raw[];
  T* xs = reinterpret_cast<T*>(smem_raw);  // [M, K]
  __shared__ float lut[16];
  if (threadIdx.x < 16) lut[threadIdx.x] = lut_g ? lut_g[threadIdx.x] : kNF4[threadIdx.x];
  for (int i = threadIdx.x * 8; i < M * K; i += blockDim.x * 8) {
    const int m = i / K, k = i - m * K;
    *reinterpret_cast<uint4*>(xs + i) = *reinterpret_cast<const uint4*>(x + (long long)m * x_rs + k);
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long n = (long long)blockIdx.x * (blockDim.x >> 5) + warp;
  if (n >= N) return;
  const int groups_per_row = K / group;
  float acc[kMaxM];
#pragma unroll
  for (int m = 0; m < kMaxM; ++m) acc[m] = 0.f;
  const uint8_t* pr = packed + n * (K / 2);
  for (int kc = lane; kc < K / 32; kc += 32) {
    const uint4 raw = *reinterpret_cast<const uint4*>(pr + kc * 16);
    const int g = (kc * 32) / group;
    const float sc = scale[n * groups_per_row + g];
    const float zs = zsp ? zsp[n * groups_per_row + g] : 0.f;
    const uint32_t words[4] = {raw.x, raw.y, raw.z, raw.w};
    float wv[32];
#pragma unroll
    for (int wi = 0; wi < 4; ++wi)
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const uint32_t byte = (words[wi] >> (8 * b)) & 0xFFu;
        wv[wi * 8 + 2 * b] = lut[byte >> 4] * sc - zs;
        wv[wi * 8 + 2 * b + 1] = lut[byte & 15u] * sc - zs;
      }
#pragma unroll
    for (int m = 0; m < kMaxM; ++m) {
      if (m < M) {
        const T* xr = xs + m * K + kc * 32;
        float a = 0.f;
#pragma unroll
        for (int v8 = 0; v8 < 4; ++v8) {
          const uint4 xv = *reinterpret_cast<const uint4*>(xr + v8 * 8);
          const T* xe = reinterpret_cast<const T*>(&xv);
#pragma unroll
          for (int j = 0; j < 8; ++j) a += to_f32<T>(xe[j]) * wv[v8 * 8 + j];
        }
        acc[m] += a;
      }
    }
  }
#pragma unroll
  for (int m = 0; m < kMaxM; ++m) {
    if (m < M) {
      const float r = warp_sum(acc[m]);
      if (lane == 0) y[(long long)m * y_rs + n] = from_f32<T>(r + (bias ? to_f32<T>(bias[n]) : 0.f));
    }
  }
}

}  // namespace b200
using namespace b200;

B200_API int b200_nf4_quantize(const void* w, uint8_t* packed, float* absmax, long long n, int dtype, void* stream) {
  if (n <= 0 || n % 64) return B200_ERR_BAD_SHAPE;
  const long long nblocks = n / 64;
  const int tb = 256;
  const unsigned grid = (unsigned)((nblocks * 32 + tb - 1) / tb);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_DT_BF16) nf4_quantize_kernel<__nv_bfloat16><<<grid, tb, 0, st>>>((const __nv_bfloat16*)w, packed, absmax, nblocks);
  else if (dtype == B200_DT_FP32) nf4_quantize_kernel<float><<<grid, tb, 0, st>>>((const float*)w, packed, absmax, nblocks);
  else if (dtype == B200_DT_FP16) nf4_quantize_kernel<__half><<<grid, tb, 0, st>>>((const __half*)w, packed, absmax, nblocks);
  else return B200_ERR_UNSUPPORTED;
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_nf4_absmax_decode(const uint8_t* q, const float* code2, const float* absmax2, float offset,
                                    float* out, long long n, void* stream) {
  if (n <= 0) return 0;
  nf4_absmax_decode_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(q, code2, absmax2, offset, out, n);
  B200_LAUNCH_CHECK();
  return 0;
}

// packed/scale/zs: arrays of `nshards` device pointers (peer-mapped pointers allowed); zs entries may be NULL.
B200_API int b200_w4_dequant(const void* const* packed, const void* const* scale, const void* const* zs, int nshards,
                             long long rows_per_shard, void* out, long long N, int K, int group, const float* lut,
                             long long out_rs, int out_dtype, void* stream) {
  if (nshards < 1 || nshards > 8 || K % 32 || group % 32 || K % group) return B200_ERR_BAD_SHAPE;
  W4Shards sh;
  for (int i = 0; i < 8; ++i) {
    sh.packed[i] = i < nshards ? (const uint8_t*)packed[i] : nullptr;
    sh.scale[i] = i < nshards ? (const float*)scale[i] : nullptr;
    sh.zs[i] = (i < nshards && zs) ? (const float*)zs[i] : nullptr;
  }
  sh.rows_per_shard = rows_per_shard;
  sh.nshards = nshards;
  const long long chunks = N * (K / 32);
  const unsigned grid = (unsigned)((chunks + 255) / 256);
  cudaStream_t st = (cudaStream_t)stream;
  // measured inside a training step (8-layer slice, ncu launch list): 63.5 us per call for the coalesced kernel vs 47.7 us for
  // the one-chunk-per-thread kernel below (4x the load instructions, scale fetched per word) — kept as an option only
  static const bool coalesced = getenv("B200_W4_DEQUANT_COALESCED") != nullptr;
  if (coalesced && out_rs % 8 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
    const long long nwords = N * (long long)K / 8;
    const unsigned g2 = (unsigned)((nwords + 256 * 4 - 1) / (256 * 4));
    if (out_dtype == B200_DT_BF16) w4_dequant_coalesced_kernel<__nv_bfloat16><<<g2, 256, 0, st>>>(sh, (__nv_bfloat16*)out, N, K, group, lut, out_rs);
    else if (out_dtype == B200_DT_FP16) w4_dequant_coalesced_kernel<__half><<<g2, 256, 0, st>>>(sh, (__half*)out, N, K, group, lut, out_rs);
    else return B200_ERR_UNSUPPORTED;
    B200_LAUNCH_CHECK();
    return 0;
  }
  if (out_dtype == B200_DT_BF16) w4_dequant_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(sh, (__nv_bfloat16*)out, N, K, group, lut, out_rs);
  else if (out_dtype == B200_DT_FP16) w4_dequant_kernel<__half><<<grid, 256, 0, st>>>(sh, (__half*)out, N, K, group, lut, out_rs);
  else return B200_ERR_UNSUPPORTED;
  B200_LAUNCH_CHECK();
  return 0;
}

// Same as b200_w4_dequant but writes the transpose: out_t [K, N] (row stride out_rs). N % 64 == 0, K % 64 == 0.
B200_API int b200_w4_dequant_t(const void* const* packed, const void* const* scale, const void* const* zs, int nshards,
                               long long rows_per_shard, void* out, long long N, int K, int group, const float* lut,
                               long long out_rs, int out_dtype, void* stream) {
  if (nshards < 1 || nshards > 8 || K % 64 || N % 64 || group % 32 || K % group) return B200_ERR_BAD_SHAPE;
  if (nshards > 1 && rows_per_shard % 64) return B200_ERR_BAD_SHAPE;
  W4Shards sh;
  for (int i = 0; i < 8; ++i) {
    sh.packed[i] = i < nshards ? (const uint8_t*)packed[i] : nullptr;
    sh.scale[i] = i < nshards ? (const float*)scale[i] : nullptr;
    sh.zs[i] = (i < nshards && zs) ? (const float*)zs[i] : nullptr;
  }
  sh.rows_per_shard = rows_per_shard;
  sh.nshards = nshards;
  const unsigned grid = (unsigned)((N / 64) * (K / 64));
  cudaStream_t st = (cudaStream_t)stream;
  if (out_dtype == B200_DT_BF16) w4_dequant_t_kernel<__nv_bfloat16><<<grid, 128, 0, st>>>(sh, (__nv_bfloat16*)out, N, K, group, lut, out_rs);
  else if (out_dtype == B200_DT_FP16) w4_dequant_t_kernel<__half><<<grid, 128, 0, st>>>(sh, (__half*)out, N, K, group, lut, out_rs);
  else return B200_ERR_UNSUPPORTED;
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_w4_gemv(const void* x, const uint8_t* packed, const float* scale, const float* zs, const float* lut,
                          const void* bias, void* y, int M, long long N, int K, int group, long long x_rs,
                          long long y_rs, int dtype, void* stream) {
  if (M < 1 || M > 8 || K % 32 || group % 32 || K % group || x_rs % 8) return B200_ERR_BAD_SHAPE;
  if (K % 128 == 0 && (dtype == B200_DT_BF16 || dtype == B200_DT_FP16) && !getenv("B200_W4_GEMV_SIMT")) {
    const long long tiles16 = (N + 15) / 16;
    const unsigned g16 = (unsigned)(tiles16 < kNumSMs * 4 ? tiles16 : kNumSMs * 4);
    cudaStream_t s2 = (cudaStream_t)stream;
    if (dtype == B200_DT_BF16)
      return (int)launch_pdl(w4_gemv_mma_kernel<true>, g16, 256, 0, s2, (const uint16_t*)x, packed, scale, zs, lut, (const uint16_t*)bias, (uint16_t*)y, M, N, K, group, x_rs, y_rs);
    return (int)launch_pdl(w4_gemv_mma_kernel<false>, g16, 256, 0, s2, (const uint16_t*)x, packed, scale, zs, lut, (const uint16_t*)bias, (uint16_t*)y, M, N, K, group, x_rs, y_rs);
  }
  const int smem = M * K * 2;
  if (smem > 200 * 1024) return B200_ERR_TOO_LARGE;
  const int tb = 256, wpb = tb / 32;
  const unsigned grid = (unsigned)((N + wpb - 1) / wpb);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_DT_BF16) {
    static int cfg = 0;
    if (cfg < smem) { B200_CHECK(cudaFuncSetAttribute(w4_gemv_kernel<__nv_bfloat16, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); cfg = 200 * 1024; }
    w4_gemv_kernel<__nv_bfloat16, 8><<<grid, tb, smem, st>>>((const __nv_bfloat16*)x, packed, scale, zs, lut, (const __nv_bfloat16*)bias, (__nv_bfloat16*)y, M, N, K, group, x_rs, y_rs);
  } else if (dtype == B200_DT_FP16) {
    static int cfg = 0;
    if (cfg < smem) { B200_CHECK(cudaFuncSetAttribute(w4_gemv_kernel<__half, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); cfg = 200 * 1024; }
    w4_gemv_kernel<__half, 8><<<grid, tb, smem, st>>>((const __half*)x, packed, scale, zs, lut, (const __half*)bias, (__half*)y, M, N, K, group, x_rs, y_rs);
  } else {
    return B200_ERR_UNSUPPORTED;
  }
  B200_LAUNCH_CHECK();
  return 0;
}
