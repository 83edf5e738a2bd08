// Small-M products for decode / serving steps: y[m, n] = sum_k x[m, k] * w[n, k] (+ bias) for M <= 64 tokens, with the
// weight matrix read exactly once — 16-bit dense weights (`b200_gemv16`) or 4-bit codes (`b200_w4_gemv_batched`).
//
// Same scheme as w4_gemv_mma_kernel (w4.cu): a CTA owns 16 output rows, its 8 warps take interleaved 128-wide k chunks,
// the weights are the A operand of mma.sync.m16n8k16 and the tokens the 8-wide B operand; the k index is permuted so
// that every lane fetches whole 16-byte vectors (lane (g, t) owns k in [128 c + 32 t, +32) of rows g and g + 8, MMA j of
// the chunk consumes elements 4j .. 4j+3). New here: TG token groups of 8 share one pass over the weights — the A
// fragments (expanded once for 4-bit codes) are reused by TG MMAs, so up to 64 concurrent sequences cost one read of
// the model per step. That is the regime of the reference's serving table (concurrency 8..64,
// LLM_on_Kubernetes/Inference_Platfrom/README.md:1504-1509), where a 128-row tensor-core tile would be >87 % padding.
#include "common.cuh"
#include "peer.cuh"
#include "../include/b200_ops.h"

#include <stdlib.h>

namespace b200 {

__constant__ float kNF4g[16] = {-1.0f, -0.6961928009986877f, -0.5250730514526367f, -0.39491748809814453f,
                                -0.28444138169288635f, -0.18477343022823334f, -0.09105003625154495f, 0.0f,
                                0.07958029955625534f, 0.16093020141124725f, 0.24611230194568634f,
                                0.33791524171829224f, 0.44070982933044434f, 0.5626170039176941f,
                                0.7229568362236023f, 1.0f};

template <bool BF16>
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  if (BF16)
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  else
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <bool BF16>
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  if (BF16) return pack_bf16x2(a, b);
  const __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}

template <bool BF16>
__device__ __forceinline__ uint32_t fma2(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  if (BF16) asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  else asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// weights are read exactly once: bypass L1 allocation
__device__ __forceinline__ uint4 ld_nc_v4(const void* ptr) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(ptr));
  return r;
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N_>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N_) : "memory"); }

// (a & b) | c in ONE instruction: with b and c as immediates the compiler splits it into two LOP3s (the expansion of the
// codes is 22 % of all issued instructions of this kernel, profiles/r2/gemv.md)
__device__ __forceinline__ uint32_t and_or(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

struct GemvParams {
  const uint16_t* x;      // [M, K] 16-bit activations
  const void* w;          // dense: uint16 [N, w_rs]; 4-bit: packed uint8 [N, K / 2]
  const float* scale;     // 4-bit only: [N, K / group]
  const float* zs;        // 4-bit only, may be null
  const float* lut;       // 4-bit only, null = NF4
  const void* bias;       // 16-bit, may be null
  void* y;                // [M, y_rs] 16-bit or fp32
  int M, K, group, out_f32;
  int gshift;             // log2(group) when it is a power of two, else -1
  float lin_a, lin_off;   // LIN: lut[code] = lin_a * (code - lin_off), lin_off an integer
  long long N, x_rs, w_rs, y_rs;
  // ---- fused tail (fuse != 0): the product never goes to a y tensor of its own. Every CTA writes its 16-column tiles into
  // a staging buffer (fuse 1: local; fuse 2: this rank's half of a symmetric buffer, tensor-parallel row-parallel layer),
  // a ticket barrier closes the GEMV, (fuse 2) the epoch is published to the peers and theirs awaited, and CTA m then
  // finishes token m: sum of the W ranks' partials (all-reduce, K-C5) + residual -> new residual, RMSNorm with the NEXT
  // layer's weight -> its input. One launch instead of GEMV + all-reduce + residual/norm. Epoch and tickets live on the
  // device (state[0..2]), so the launch replays from a CUDA graph.
  int fuse, world, rank, channel;
  PeerPtrs stage;             // 16-bit [2][half_elems] per rank (fuse 2) / stage.p[0] = local buffer (fuse 1)
  PeerPtrs flags;
  long long half_elems;
  unsigned int* state;
  const void* res_in;         // [M, N] 16-bit, may be null (first layer)
  void* res_out;              // [M, N] 16-bit
  const void* norm_w;         // [N] 16-bit, null: no norm, x_out receives the sum
  void* x_out;                // [M, N] 16-bit
  float eps;
};

// LIN (4-bit codes that are a linear function of the code, i.e. RTN / GPTQ / AWQ integers): no lookup table. A nibble is
// OR-ed into the mantissa of 128.0 (bf16) / 1024.0 (fp16), one exact add removes that bias and the code offset, one FMA
// applies scale and zero: two FP instructions per pair of weights instead of a shared-memory lookup — the table variant is
// bound by the LSU pipe (profiles/decode_kernels_ncu.md). Nibbles are taken four pairs at a time from each 32-bit word
// ((w >> 4 i) & 0x000F000F), which pairs elements (e, e + 4); the activations get the same pairing with two PRMTs per
// word, so the k order inside an MMA is permuted consistently on both operands.
// POST (LIN only, group a multiple of 128 so that a 128-wide chunk lies inside one scale group): the codes enter the MMA
// as the exact 16-bit values (bias + code) — one logic op per pair of weights, no floating-point conversion at all — and
// the affine map is applied to the chunk's fp32 partial sum instead:  sum_k w_k x_k = A_row * (sum_k (bias + c_k) x_k -
// B_row * sum_k x_k), with sum_k x_k produced by one more MMA per k-step whose weight operand is the constant 1.0. That
// replaces 32 packed FMAs per lane and chunk by 8 MMAs and 8 scalar FMAs: the kernel is bound by instruction issue,
// not by HBM (3.5 instructions per weight before; profiles/r2/gemv.md).
template <bool BF16, bool W4, int TG, int MINB, bool LIN = false, bool POST = false>
__global__ void __launch_bounds__(256, MINB)
gemv_mma_kernel(const GemvParams p) {
  extern __shared__ uint32_t table[];   // 4-bit table variant only: byte -> two 16-bit values, one column per lane (conflict-free)
  __shared__ float red[8][128];
  pdl_launch_dependents();
  if (W4 && !LIN) {   // the table depends on no earlier kernel: it is built while the previous kernel drains
    for (int i = threadIdx.x; i < 256 * 32; i += 256) {
      const int byte = i >> 5;
      const float lo = p.lut ? p.lut[byte >> 4] : kNF4g[byte >> 4];
      const float hi = p.lut ? p.lut[byte & 15] : kNF4g[byte & 15];
      table[i] = pack2<BF16>(lo, hi);
    }
    __syncthreads();
  }
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int K = p.K, M = p.M;
  const long long N = p.N;
  // activations -> shared memory once per CTA (the host sized the allocation; 16-byte padded rows)
  const bool xsm = (long long)M * (K + 8) * 2 <= 73728;
  uint16_t* xs = reinterpret_cast<uint16_t*>(table + ((W4 && !LIN) ? 256 * 32 : 0));
  if (xsm) {
    const int vpr = K / 8;
    for (int i = threadIdx.x; i < M * vpr; i += 256) {
      const int r = i / vpr, cvec = i - r * vpr;
      *reinterpret_cast<uint4*>(xs + (long long)r * (K + 8) + cvec * 8) = __ldg(reinterpret_cast<const uint4*>(p.x + (long long)r * p.x_rs) + cvec);
    }
    __syncthreads();
  }
  uint32_t epoch = 0;
  void* ybase = p.y;
  long long y_rs = p.y_rs;
  if (p.fuse) {
    epoch = *((volatile unsigned int*)p.state) + 1u;
    const int me = p.fuse == 2 ? p.rank : 0;
    ybase = reinterpret_cast<uint16_t*>(p.stage.p[me]) + (p.fuse == 2 ? (long long)(epoch & 1u) * p.half_elems : 0ll);
    y_rs = N;
  }
  const int gpr = W4 ? K / p.group : 0;
  const uint32_t* tl = table + lane;
  const int nchunks = K / 128;
  const long long ntiles = (N + 15) / 16;
  // Work item = (16-row tile, batch of U chunks per warp). All packed weights (and scales) of an item are requested
  // before any of them is used, so an item costs ONE trip to DRAM instead of U dependent ones; the activations come from
  // shared memory (staged once per CTA when they fit), so the only long-latency operation left inside an item is that
  // one trip. (A cp.async prefetch ring in shared memory was measured 1.5x SLOWER than this: profiles/r2/gemv.md.)
  constexpr int U = MINB >= 3 ? 1 : 4;     // 3-4 CTAs per SM (<= 80 registers): one chunk at a time, occupancy hides the latency
  const int nb = (nchunks + 8 * U - 1) / (8 * U);
  const long long my_tiles = blockIdx.x < ntiles ? (ntiles - 1 - blockIdx.x) / gridDim.x + 1 : 0;
  const long long nitems = my_tiles * nb;
  float c[TG][4];
  for (long long it = 0; it < nitems; ++it) {
    const long long tile = blockIdx.x + (it / nb) * gridDim.x;
    const int bidx = (int)(it % nb);
    const int ch0 = warp + bidx * 8 * U;
    const long long row0 = tile * 16;
    const long long rA = min(row0 + g, N - 1), rB = min(row0 + g + 8, N - 1);
    if (bidx == 0) {
#pragma unroll
      for (int tg = 0; tg < TG; ++tg)
#pragma unroll
        for (int i = 0; i < 4; ++i) c[tg][i] = 0.f;
    }
    {
      uint4 wqa[U], wqb[U];        // W4: 16 packed bytes of rows g / g+8 per chunk; dense: unused
      float sca[U], scb[U], zza[U], zzb[U];
      if (W4) {
        const uint8_t* pa0 = reinterpret_cast<const uint8_t*>(p.w) + rA * (K / 2) + t * 16;
        const uint8_t* pb0 = reinterpret_cast<const uint8_t*>(p.w) + rB * (K / 2) + t * 16;
        const float* sa0 = p.scale + rA * gpr;
        const float* sb0 = p.scale + rB * gpr;
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int ch = ch0 + 8 * u;
          if (ch < nchunks) {
            wqa[u] = ld_nc_v4(pa0 + ch * 64);
            wqb[u] = ld_nc_v4(pb0 + ch * 64);
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int ch = ch0 + 8 * u;
          if (ch < nchunks) {
            const int k = ch * 128 + t * 32;
            const int gi = p.gshift >= 0 ? (k >> p.gshift) : (k / p.group);
            sca[u] = __ldg(sa0 + gi);
            scb[u] = __ldg(sb0 + gi);
            zza[u] = p.zs ? __ldg(p.zs + rA * gpr + gi) : 0.f;
            zzb[u] = p.zs ? __ldg(p.zs + rB * gpr + gi) : 0.f;
          }
        }
      }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int ch = ch0 + 8 * u;
      if (ch >= nchunks) break;
      const int k = ch * 128 + t * 32;
      uint32_t aA[16], aB[16];   // rows g / g+8: 32 weights each as 16-bit pairs
      if (W4) {
        const uint4 wa = wqa[u], wb = wqb[u];
        const float sa = sca[u], sb = scb[u];
        const float za = zza[u], zb = zzb[u];
        const uint32_t sa2 = pack2<BF16>(sa, sa), sb2 = pack2<BF16>(sb, sb), za2 = pack2<BF16>(-za, -za), zb2 = pack2<BF16>(-zb, -zb);
        const uint32_t wA[4] = {wa.x, wa.y, wa.z, wa.w}, wB[4] = {wb.x, wb.y, wb.z, wb.w};
        if (LIN && POST) {
          uint32_t kMagic = BF16 ? 0x43004300u : 0x64006400u, kMask = 0x000F000Fu;
          asm volatile("" : "+r"(kMagic), "+r"(kMask));      // keep both in registers (one LOP3 per pair of codes)
#pragma unroll
          for (int q = 0; q < 4; ++q) {   // same (e0,e4) (e1,e5) (e2,e6) (e3,e7) register order as below, values left biased
            aA[4 * q + 0] = and_or(wA[q] >> 4, kMask, kMagic);
            aA[4 * q + 1] = and_or(wA[q], kMask, kMagic);
            aA[4 * q + 2] = and_or(wA[q] >> 12, kMask, kMagic);
            aA[4 * q + 3] = and_or(wA[q] >> 8, kMask, kMagic);
            aB[4 * q + 0] = and_or(wB[q] >> 4, kMask, kMagic);
            aB[4 * q + 1] = and_or(wB[q], kMask, kMagic);
            aB[4 * q + 2] = and_or(wB[q] >> 12, kMask, kMagic);
            aB[4 * q + 3] = and_or(wB[q] >> 8, kMask, kMagic);
          }
        } else if (LIN) {
          constexpr uint32_t kMagic = BF16 ? 0x43004300u : 0x64006400u, kMask = 0x000F000Fu;
          const float bias0 = (BF16 ? 128.f : 1024.f) + p.lin_off;
          const uint32_t one2 = pack2<BF16>(1.f, 1.f), neg2 = pack2<BF16>(-bias0, -bias0);
          const uint32_t la2 = pack2<BF16>(sa * p.lin_a, sa * p.lin_a), lb2 = pack2<BF16>(sb * p.lin_a, sb * p.lin_a);
#pragma unroll
          for (int q = 0; q < 4; ++q) {   // word q = elements 8q .. 8q+7; register order (e0,e4) (e1,e5) (e2,e6) (e3,e7)
            const uint32_t a0 = (wA[q] & kMask) | kMagic, a1 = ((wA[q] >> 4) & kMask) | kMagic;
            const uint32_t a2 = ((wA[q] >> 8) & kMask) | kMagic, a3 = ((wA[q] >> 12) & kMask) | kMagic;
            const uint32_t b0 = (wB[q] & kMask) | kMagic, b1 = ((wB[q] >> 4) & kMask) | kMagic;
            const uint32_t b2 = ((wB[q] >> 8) & kMask) | kMagic, b3 = ((wB[q] >> 12) & kMask) | kMagic;
            aA[4 * q + 0] = fma2<BF16>(fma2<BF16>(a1, one2, neg2), la2, za2);
            aA[4 * q + 1] = fma2<BF16>(fma2<BF16>(a0, one2, neg2), la2, za2);
            aA[4 * q + 2] = fma2<BF16>(fma2<BF16>(a3, one2, neg2), la2, za2);
            aA[4 * q + 3] = fma2<BF16>(fma2<BF16>(a2, one2, neg2), la2, za2);
            aB[4 * q + 0] = fma2<BF16>(fma2<BF16>(b1, one2, neg2), lb2, zb2);
            aB[4 * q + 1] = fma2<BF16>(fma2<BF16>(b0, one2, neg2), lb2, zb2);
            aB[4 * q + 2] = fma2<BF16>(fma2<BF16>(b3, one2, neg2), lb2, zb2);
            aB[4 * q + 3] = fma2<BF16>(fma2<BF16>(b2, one2, neg2), lb2, zb2);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) {   // byte i of the 16 packed bytes = elements 2i, 2i+1
            const uint32_t ba = (wA[i >> 2] >> ((i & 3) * 8)) & 0xFFu, bb = (wB[i >> 2] >> ((i & 3) * 8)) & 0xFFu;
            aA[i] = fma2<BF16>(tl[ba << 5], sa2, za2);
            aB[i] = fma2<BF16>(tl[bb << 5], sb2, zb2);
          }
        }
      } else {
        const uint16_t* wd = reinterpret_cast<const uint16_t*>(p.w);
        const uint4* pa = reinterpret_cast<const uint4*>(wd + rA * p.w_rs + k);
        const uint4* pb = reinterpret_cast<const uint4*>(wd + rB * p.w_rs + k);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint4 va = ld_nc_v4(pa + i), vb = ld_nc_v4(pb + i);
          aA[4 * i] = va.x; aA[4 * i + 1] = va.y; aA[4 * i + 2] = va.z; aA[4 * i + 3] = va.w;
          aB[4 * i] = vb.x; aB[4 * i + 1] = vb.y; aB[4 * i + 2] = vb.z; aB[4 * i + 3] = vb.w;
        }
      }
#pragma unroll
      for (int tg = 0; tg < TG; ++tg) {
        const int tok = tg * 8 + g;
        uint32_t xw[16];
        if (tok < M) {
          const uint4* xp = xsm ? reinterpret_cast<const uint4*>(xs + (long long)tok * (K + 8) + k)
                                : reinterpret_cast<const uint4*>(p.x + (long long)tok * p.x_rs + k);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const uint4 v = xsm ? xp[i] : __ldg(xp + i);
            xw[4 * i] = v.x; xw[4 * i + 1] = v.y; xw[4 * i + 2] = v.z; xw[4 * i + 3] = v.w;
          }
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) xw[i] = 0u;
        }
        if (LIN) {   // same (e, e + 4) pairing as the weights
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t x0 = xw[4 * q], x1 = xw[4 * q + 1], x2 = xw[4 * q + 2], x3 = xw[4 * q + 3];
            xw[4 * q + 0] = __byte_perm(x0, x2, 0x5410);
            xw[4 * q + 1] = __byte_perm(x0, x2, 0x7632);
            xw[4 * q + 2] = __byte_perm(x1, x3, 0x5410);
            xw[4 * q + 3] = __byte_perm(x1, x3, 0x7632);
          }
        }
        if (LIN && POST) {
          constexpr uint32_t kOne2 = BF16 ? 0x3F803F80u : 0x3C003C00u;
          // two independent accumulator chains per sum (a chain of 8 dependent MMAs costs ~8 x 33 cycles of latency)
          float craw[4] = {0.f, 0.f, 0.f, 0.f}, csx[4] = {0.f, 0.f, 0.f, 0.f};
          float craw2[4] = {0.f, 0.f, 0.f, 0.f}, csx2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int j = 0; j < 8; j += 2) {
            mma16816<BF16>(craw, aA[2 * j], aB[2 * j], aA[2 * j + 1], aB[2 * j + 1], xw[2 * j], xw[2 * j + 1]);
            mma16816<BF16>(craw2, aA[2 * j + 2], aB[2 * j + 2], aA[2 * j + 3], aB[2 * j + 3], xw[2 * j + 2], xw[2 * j + 3]);
            mma16816<BF16>(csx, kOne2, kOne2, kOne2, kOne2, xw[2 * j], xw[2 * j + 1]);
            mma16816<BF16>(csx2, kOne2, kOne2, kOne2, kOne2, xw[2 * j + 2], xw[2 * j + 3]);
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) { craw[i] += craw2[i]; csx[i] += csx2[i]; }
          // w = lin_a * (code - lin_off) * scale - zs = A * ((bias + code) - B),  A = lin_a * scale,  B = bias + lin_off + zs / A
          const float bias0 = BF16 ? 128.f : 1024.f;
          const float Aa = sca[u] * p.lin_a, Ab = scb[u] * p.lin_a;
          const float Ba = bias0 + p.lin_off + __fdividef(zza[u], Aa), Bb = bias0 + p.lin_off + __fdividef(zzb[u], Ab);
          c[tg][0] += Aa * (craw[0] - Ba * csx[0]);
          c[tg][1] += Aa * (craw[1] - Ba * csx[1]);
          c[tg][2] += Ab * (craw[2] - Bb * csx[2]);
          c[tg][3] += Ab * (craw[3] - Bb * csx[3]);
        } else {
          float c2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int j = 0; j < 8; j += 2) {
            mma16816<BF16>(c[tg], aA[2 * j], aB[2 * j], aA[2 * j + 1], aB[2 * j + 1], xw[2 * j], xw[2 * j + 1]);
            mma16816<BF16>(c2, aA[2 * j + 2], aB[2 * j + 2], aA[2 * j + 3], aB[2 * j + 3], xw[2 * j + 2], xw[2 * j + 3]);
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) c[tg][i] += c2[i];
        }
      }
    }
    }
    if (bidx != nb - 1) continue;
    // c0: (row g, token 2t), c1: (row g, token 2t+1), c2: (row g+8, token 2t), c3: (row g+8, token 2t+1) of the group
#pragma unroll
    for (int tg = 0; tg < TG; ++tg) {
#pragma unroll
      for (int i = 0; i < 4; ++i) red[warp][lane * 4 + i] = c[tg][i];
      __syncthreads();
      if (threadIdx.x < 128) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) s += red[w][threadIdx.x];
        const int ln = threadIdx.x >> 2, i = threadIdx.x & 3;
        const long long row = row0 + (ln >> 2) + ((i & 2) ? 8 : 0);
        const int tok = tg * 8 + 2 * (ln & 3) + (i & 1);
        if (row < N && tok < M) {
          if (p.bias) s += BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.bias)[row])
                                : __half2float(reinterpret_cast<const __half*>(p.bias)[row]);
          if (p.out_f32 && !p.fuse) reinterpret_cast<float*>(ybase)[(long long)tok * y_rs + row] = s;
          else if (BF16) reinterpret_cast<__nv_bfloat16*>(ybase)[(long long)tok * y_rs + row] = __float2bfloat16_rn(s);
          else reinterpret_cast<__half*>(ybase)[(long long)tok * y_rs + row] = __float2half_rn(s);
        }
      }
      __syncthreads();   // `red` is reused by the next token group / tile
    }
  }
  if (!p.fuse) return;
  // ================= fused tail: (all-reduce +) residual + RMSNorm per token =================
  __syncthreads();               // all of this CTA's tiles are written ...
  if (threadIdx.x == 0) {
    if (p.fuse == 2) __threadfence_system(); else __threadfence();   // ... and made visible (to the peers too) by ONE fence
    const unsigned int old = atomicAdd(p.state + 1, 1u);
    const unsigned int target = (old / gridDim.x + 1u) * gridDim.x;
    const long long t0 = clock64();
    while ((int)(*((volatile unsigned int*)(p.state + 1)) - target) < 0) {
      if (clock64() - t0 > 20000000000ll) { printf("b200 gemv: fused-tail grid barrier timed out (block %d)\n", blockIdx.x); __trap(); }
    }
    __threadfence();
  }
  __syncthreads();
  const int W = p.fuse == 2 ? p.world : 1;
  if (p.fuse == 2) {
    if (blockIdx.x == 0) signal_all(p.flags, p.channel, W, p.rank, epoch);
    if ((int)blockIdx.x < M) wait_all(p.flags, p.channel, W, p.rank, epoch);
  }
  {
    float* redf = &red[0][0];
    const int nvec = (int)(N / 8);
    const long long hoff = p.fuse == 2 ? (long long)(epoch & 1u) * p.half_elems : 0ll;
    for (int m = blockIdx.x; m < M; m += gridDim.x) {
      // up to 4 vectors of 8 per thread (N <= 8192): the sums stay in registers between the two passes
      float v[4][8];
      float ss = 0.f;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int i = threadIdx.x + k * 256;
        if (i < nvec) {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[k][j] = 0.f;
          for (int r = 0; r < W; ++r) {
            const uint4 q = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.stage.p[p.fuse == 2 ? r : 0]) + hoff + (long long)m * N)[i];
            const uint32_t w4[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float2 f;
              if (BF16) f = unpack_bf16x2(w4[j]);
              else f = __half22float2(*reinterpret_cast<const __half2*>(&w4[j]));
              v[k][2 * j] += f.x; v[k][2 * j + 1] += f.y;
            }
          }
          if (p.res_in) {
            const uint4 q = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.res_in) + (long long)m * N)[i];
            const uint32_t w4[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float2 f;
              if (BF16) f = unpack_bf16x2(w4[j]);
              else f = __half22float2(*reinterpret_cast<const __half2*>(&w4[j]));
              v[k][2 * j] += f.x; v[k][2 * j + 1] += f.y;
            }
          }
          // the hidden state is rounded to 16 bits first (what the stand-alone norm kernel does), then normalised
          uint32_t o[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            o[j] = pack2<BF16>(v[k][2 * j], v[k][2 * j + 1]);
            float2 f;
            if (BF16) f = unpack_bf16x2(o[j]);
            else f = __half22float2(*reinterpret_cast<const __half2*>(&o[j]));
            v[k][2 * j] = f.x; v[k][2 * j + 1] = f.y;
            ss += f.x * f.x + f.y * f.y;
          }
          if (p.res_out) reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.res_out) + (long long)m * N)[i] = make_uint4(o[0], o[1], o[2], o[3]);
          if (!p.norm_w) reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.x_out) + (long long)m * N)[i] = make_uint4(o[0], o[1], o[2], o[3]);
        }
      }
      if (p.norm_w) {
        ss = block_sum(ss, redf);
        const float rstd = rsqrtf(ss / (float)N + p.eps);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int i = threadIdx.x + k * 256;
          if (i < nvec) {
            const uint4 q = reinterpret_cast<const uint4*>(p.norm_w)[i];
            const uint32_t w4[4] = {q.x, q.y, q.z, q.w};
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float2 f;
              if (BF16) f = unpack_bf16x2(w4[j]);
              else f = __half22float2(*reinterpret_cast<const __half2*>(&w4[j]));
              o[j] = pack2<BF16>(v[k][2 * j] * rstd * f.x, v[k][2 * j + 1] * rstd * f.y);
            }
            reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.x_out) + (long long)m * N)[i] = make_uint4(o[0], o[1], o[2], o[3]);
          }
        }
        __syncthreads();
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tkt = atomicAdd(p.state + 2, 1u);
    if (tkt == gridDim.x - 1) {
      p.state[2] = 0u;
      __threadfence();
      *((volatile unsigned int*)p.state) = epoch;   // the next fused launch (stream order) sees the new epoch
    }
  }
}

template <bool BF16, bool W4, int TG, int MINB, bool LIN = false, bool POST = false>
static cudaError_t launch_gemv_one(const GemvParams& p_in, cudaStream_t st) {
  GemvParams p = p_in;
  p.gshift = (p.group > 0 && (p.group & (p.group - 1)) == 0) ? __builtin_ctz(p.group) : -1;
  // persistent over 16-row tiles: one resident wave (occupancy x SMs), never more CTAs than tiles
  // [table 32 KB (4-bit, non-linear code book only)] + the activations [M, K] 16-bit when they fit in 64 KB
  const long long xbytes = (long long)p.M * (p.K + 8) * 2;      // rows padded by 16 bytes (bank spread)
  const int smem = ((W4 && !LIN) ? 256 * 32 * 4 : 0) + (xbytes <= 73728 ? (int)xbytes : 0);
  static bool attr_done = false;
  static int occ_by_kb[129] = {0};
  if (!attr_done) {
    cudaFuncSetAttribute(gemv_mma_kernel<BF16, W4, TG, MINB, LIN, POST>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
    attr_done = true;
  }
  int& occ = occ_by_kb[(smem + 1023) >> 10];
  if (!occ) {
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, gemv_mma_kernel<BF16, W4, TG, MINB, LIN, POST>, 256, smem) != cudaSuccess || occ < 1) occ = 1;
  }
  const long long tiles = (p.N + 15) / 16;
  const unsigned grid = (unsigned)(tiles < (long long)kNumSMs * occ ? tiles : (long long)kNumSMs * occ);
  return launch_pdl(gemv_mma_kernel<BF16, W4, TG, MINB, LIN, POST>, grid, 256, smem, st, p);
}

template <bool BF16>
static int launch_gemv_lin(const GemvParams& p, cudaStream_t st) {
  cudaError_t e;
  // 2 CTAs per SM (128 registers, no spills): with 4 chunks x 2 rows x 16 B requested up front per lane that is 64 KB in
  // flight per SM, above the ~44 KB that 6.5 TB/s x ~1 us of DRAM latency asks of each of the 148 SMs
  static const bool post_off = [] { const char* e = getenv("B200_GEMV_POST"); return e && atoi(e) == 0; }();
  static const int minb8 = [] { const char* e = getenv("B200_GEMV_MINB8"); return e ? atoi(e) : 2; }();
  if (p.group % 128 == 0 && !post_off) {   // scale applied to the chunk sums (see POST above)
    if (p.M <= 8) e = minb8 >= 4 ? launch_gemv_one<BF16, true, 1, 3, true, true>(p, st) : launch_gemv_one<BF16, true, 1, 2, true, true>(p, st);
    else if (p.M <= 16) e = launch_gemv_one<BF16, true, 2, 2, true, true>(p, st);
    else if (p.M <= 32) e = launch_gemv_one<BF16, true, 4, 2, true, true>(p, st);
    else e = launch_gemv_one<BF16, true, 8, 1, true, true>(p, st);
    return (int)e;
  }
  if (p.M <= 8) e = minb8 >= 4 ? launch_gemv_one<BF16, true, 1, 3, true>(p, st) : launch_gemv_one<BF16, true, 1, 2, true>(p, st);
  else if (p.M <= 16) e = launch_gemv_one<BF16, true, 2, 2, true>(p, st);
  else if (p.M <= 32) e = launch_gemv_one<BF16, true, 4, 2, true>(p, st);
  else e = launch_gemv_one<BF16, true, 8, 1, true>(p, st);
  return (int)e;
}

template <bool BF16, bool W4>
static int launch_gemv(const GemvParams& p, cudaStream_t st) {
  // dense weights: registers cap the occupancy (96 -> 2 CTAs per SM); forcing 3 trades a few spills for 50 % more loads in
  // flight, which is what a weight-streaming kernel is short of (B200_GEMV_MINB=2|3|4 to compare)
  static const int minb = [] { const char* e = getenv("B200_GEMV_MINB"); return e ? atoi(e) : 0; }();
  cudaError_t e;
#define B200_GV(TGv)                                                                     \
  (W4 || minb == 0 || minb == 2 ? launch_gemv_one<BF16, W4, TGv, W4 ? 1 : 2>(p, st)        \
   : minb == 3                  ? launch_gemv_one<BF16, W4, TGv, W4 ? 1 : 3>(p, st)        \
                                : launch_gemv_one<BF16, W4, TGv, W4 ? 1 : 4>(p, st))
  if (p.M <= 8) e = B200_GV(1);
  else if (p.M <= 16) e = B200_GV(2);
  else if (p.M <= 32) e = B200_GV(4);
  else e = B200_GV(8);
#undef B200_GV
  return (int)e;
}

}  // namespace b200
using namespace b200;

B200_API int b200_gemv16(const void* x, const void* w, const void* bias, void* y, int M, long long N, int K, long long x_rs,
                         long long w_rs, long long y_rs, int dtype, int out_f32, void* stream) {
  if (M < 1 || M > 64 || K % 128 || x_rs % 8 || w_rs % 8) return B200_ERR_BAD_SHAPE;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w)) & 15) return B200_ERR_BAD_ALIGN;
  GemvParams p = {};
  p.x = (const uint16_t*)x; p.w = w; p.bias = bias; p.y = y;
  p.M = M; p.K = K; p.N = N; p.x_rs = x_rs; p.w_rs = w_rs; p.y_rs = y_rs; p.out_f32 = out_f32; p.group = 128;
  if (dtype == B200_DT_BF16) return launch_gemv<true, false>(p, (cudaStream_t)stream);
  if (dtype == B200_DT_FP16) return launch_gemv<false, false>(p, (cudaStream_t)stream);
  return B200_ERR_UNSUPPORTED;
}

B200_API int b200_w4_gemv_batched(const void* x, const uint8_t* packed, const float* scale, const float* zs, const float* lut,
                                  const void* bias, void* y, int M, long long N, int K, int group, long long x_rs,
                                  long long y_rs, int dtype, void* stream) {
  if (M < 1 || M > 64 || K % 128 || group % 32 || K % group || x_rs % 8) return B200_ERR_BAD_SHAPE;
  GemvParams p = {};
  p.x = (const uint16_t*)x; p.w = packed; p.scale = scale; p.zs = zs; p.lut = lut; p.bias = bias; p.y = y;
  p.M = M; p.K = K; p.N = N; p.x_rs = x_rs; p.w_rs = 0; p.y_rs = y_rs; p.out_f32 = 0; p.group = group;
  if (dtype == B200_DT_BF16) return launch_gemv<true, true>(p, (cudaStream_t)stream);
  if (dtype == B200_DT_FP16) return launch_gemv<false, true>(p, (cudaStream_t)stream);
  return B200_ERR_UNSUPPORTED;
}

B200_API int b200_w4_gemv_linear(const void* x, const uint8_t* packed, const float* scale, const float* zs, const void* bias,
                                 void* y, int M, long long N, int K, int group, long long x_rs, long long y_rs, float lin_a,
                                 float lin_off, int dtype, void* stream) {
  if (M < 1 || M > 64 || K % 128 || group % 32 || K % group || x_rs % 8) return B200_ERR_BAD_SHAPE;
  if (lin_off != (float)(int)lin_off || lin_off < -64.f || lin_off > 64.f) return B200_ERR_UNSUPPORTED;
  GemvParams p = {};
  p.x = (const uint16_t*)x; p.w = packed; p.scale = scale; p.zs = zs; p.lut = nullptr; p.bias = bias; p.y = y;
  p.M = M; p.K = K; p.N = N; p.x_rs = x_rs; p.w_rs = 0; p.y_rs = y_rs; p.out_f32 = 0; p.group = group;
  p.lin_a = lin_a; p.lin_off = lin_off;
  if (dtype == B200_DT_BF16) return launch_gemv_lin<true>(p, (cudaStream_t)stream);
  if (dtype == B200_DT_FP16) return launch_gemv_lin<false>(p, (cudaStream_t)stream);
  return B200_ERR_UNSUPPORTED;
}

// Row-parallel / plain projection with the fused tail (see GemvParams::fuse): kind 0 dense 16-bit weights, 1 4-bit codes
// through the 16-entry table (NF4), 2 4-bit integer codes (GPTQ / AWQ / RTN). world == 1: residual + norm only.
B200_API int b200_gemv_fused(const b200_gemv_fused_args* a, void* stream) {
  if (!a || a->M < 1 || a->M > 64 || a->K % 128 || a->x_rs % 8 || a->N % 8 || a->N > 8192) return B200_ERR_BAD_SHAPE;
  if (a->world < 1 || a->world > kMaxRanks || !a->state || !a->stage_ptrs || !a->x_out) return B200_ERR_BAD_SHAPE;
  if (a->world > 1 && (!a->flag_ptrs || a->half_elems < (long long)a->M * a->N || a->half_elems % 8)) return B200_ERR_BAD_SHAPE;
  if (a->kind != 0 && (a->group % 32 || a->K % a->group)) return B200_ERR_BAD_SHAPE;
  GemvParams p = {};
  p.x = (const uint16_t*)a->x; p.w = a->w; p.scale = a->scale; p.zs = a->zs; p.lut = a->lut; p.bias = a->bias; p.y = nullptr;
  p.M = a->M; p.K = a->K; p.N = a->N; p.x_rs = a->x_rs; p.w_rs = a->w_rs; p.y_rs = a->N; p.out_f32 = 0;
  p.group = a->kind == 0 ? 128 : a->group;
  p.lin_a = a->lin_a; p.lin_off = a->lin_off;
  p.fuse = a->world > 1 ? 2 : 1;
  p.world = a->world; p.rank = a->rank; p.channel = a->channel;
  for (int i = 0; i < kMaxRanks; ++i) {
    p.stage.p[i] = i < a->world ? const_cast<void*>(a->stage_ptrs[i]) : nullptr;
    p.flags.p[i] = (a->flag_ptrs && i < a->world) ? const_cast<void*>(a->flag_ptrs[i]) : nullptr;
  }
  p.half_elems = a->half_elems; p.state = a->state;
  p.res_in = a->res_in; p.res_out = a->res_out; p.norm_w = a->norm_w; p.x_out = a->x_out; p.eps = a->eps;
  cudaStream_t st = (cudaStream_t)stream;
  const bool bf = a->dtype == B200_DT_BF16;
  if (!bf && a->dtype != B200_DT_FP16) return B200_ERR_UNSUPPORTED;
  if (a->kind == 0) return bf ? launch_gemv<true, false>(p, st) : launch_gemv<false, false>(p, st);
  if (a->kind == 1) return bf ? launch_gemv<true, true>(p, st) : launch_gemv<false, true>(p, st);
  if (a->kind == 2) {
    if (a->lin_off != (float)(int)a->lin_off || a->lin_off < -64.f || a->lin_off > 64.f) return B200_ERR_UNSUPPORTED;
    return bf ? launch_gemv_lin<true>(p, st) : launch_gemv_lin<false>(p, st);
  }
  return B200_ERR_UNSUPPORTED;
}
