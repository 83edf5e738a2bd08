// K4 on the 5th-generation tensor cores: causal flash-attention forward with tcgen05.mma, S, P and O in TMEM, Q / K / V
// tiles brought in by TMA straight out of the fused qkv projection (no head transposes, GQA by column offset), online
// softmax by four warps whose threads each own ONE query row (TMEM lane = row, so row max and row sum need no shuffles).
//
//   CTA = (128-query tile, head, batch); 256 threads: warps 0-3 softmax, warp 4 TMA producer, warp 5 tcgen05 issuer (the
//   whole warp runs the issue loop, elect.sync picks the lane: descriptors stay in uniform registers).
//   TMEM (256 columns, two CTAs per SM): S(even j) [0, 64), S(odd j) [64, 128), O [128, 256).
//   per 64-key block j:   S_j = Q K_j^T          (M 128 x N 64 x K 128, 8 UMMAs), issued one block AHEAD of the softmax
//                         softmax warps: tcgen05.ld S_j (64 registers), running max, P = exp2(S c - m c) -> 16 bit,
//                                        written over S_j in TMEM (tcgen05.st, two values per column)
//                         O += P V_j             (M 128 x N 128 x K 64, 4 UMMAs, A operand from TMEM, V MN-major in place)
//   O never leaves TMEM during the loop: the running max is only raised when a row's new max exceeds it by more than
//   2^8 (exponents stay representable either way), and only then is O read, scaled and written back — a rare event
//   after the first blocks of a row. The tensor pipe therefore sees QK(j+1), PV(j), QK(j+2), ... back to back while the
//   softmax warps work on block j, and the second resident CTA fills the remaining bubbles.
//   B200_FLASH_TC_PSMEM=1 selects the variant that hands P over through a 128B-swizzled shared-memory tile instead.
//
// Replaces the mma.sync kernel of flash_attn.cu for head_dim 128 forward passes (training forward, prefill); that
// kernel stays for other head sizes, dropout and the backward pass. Reference call sites: SDPA behind
// nn.MultiheadAttention (LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:89-96) and HF Qwen3 attention
// (Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:95-107).
#include "common.cuh"
#include "tma_host.h"
#include "../include/b200_ops.h"

#include <stdio.h>
#include <stdlib.h>

namespace b200 {

constexpr int kTcBM = 128, kTcBN = 64, kTcD = 128;
constexpr int kTcThreads = 256;
constexpr uint32_t kTcQBytes = 32768, kTcKVStage = 16384, kTcPBytes = 16384;
constexpr uint32_t kTcSmemQ = 0, kTcSmemK = 32768, kTcSmemV = 65536, kTcSmemP = 98304, kTcSmemBar = 114688;
constexpr uint32_t kTcSmemBytes = kTcSmemBar + 256;
constexpr uint32_t kTcTmemCols = 256, kTcTmemO = 128;

struct FlashTcParams {
  void* out;
  float* lse;         // [B, H, S] natural-log log-sum-exp of the scaled scores (what the backward kernel reads), may be null
  int B, S, H, Hkv;
  long long out_rs;   // elements between output rows
  float scale_log2;   // softmax scale * log2(e)
  int causal, is_bf16;
};

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// D[tmem] (+)= A[tmem] * B[smem desc]: the A operand (P, 16-bit, row = lane, two k values per column) comes from TMEM.
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// The same instructions issued from CONVERGED warp code (every lane of the issuing warp runs the loop, elect.sync picks the
// lane): descriptors stay in uniform registers. From inside `if (lane == 0)` each UTCHMMA costs two R2UR moves and a
// divergence loop — more cycles than a 128 x 64 x 16 product spends on the tensor pipe.
__device__ __forceinline__ void umma_f16_w(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_f16_ts_w(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_w(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar) : "memory");
}

template <bool BF16>
__device__ __forceinline__ uint32_t pack16(float a, float b) {
  if (BF16) return pack_bf16x2(a, b);
  const __half2 hh = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&hh);
}

// PT: P goes to the PV product through TMEM (over S); otherwise through the shared-memory tile at kTcSmemP.
template <bool BF16, bool PT>
__global__ void __launch_bounds__(kTcThreads, 2)
flash_fwd_tc_kernel(const __grid_constant__ CUtensorMap tm, const FlashTcParams p) {
  extern __shared__ __align__(1024) uint8_t smem_tc[];
  const uint32_t sbase = smem_u32(smem_tc);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // heavy (late) query tiles first: under the causal mask tile t has t + 1 times the work of tile 0
  const int qt = (int)gridDim.x - 1 - (int)blockIdx.x;
  const int h = blockIdx.y, b = blockIdx.z;
  const int S = p.S;
  const int q0 = qt * kTcBM;
  const int hk = h / (p.H / p.Hkv);
  const int qcol = h * kTcD, kcol = (p.H + hk) * kTcD, vcol = (p.H + p.Hkv + hk) * kTcD;
  const int row_base = b * S;
  const int kv_end = p.causal ? min(S, q0 + kTcBM) : S;
  const int nblk = (kv_end + kTcBN - 1) / kTcBN;

  const uint32_t bar = sbase + kTcSmemBar;
  const uint32_t q_full = bar, kv_full0 = bar + 8, kv_empty0 = bar + 24, s_full0 = bar + 40, s_free0 = bar + 56,
                 p_full = bar + 72, o_done = bar + 80, tmem_slot = bar + 96;
  if (threadIdx.x == 0) {
    if (sbase & 1023u) { printf("flash_tc: shared memory base is not 1024-byte aligned\n"); __trap(); }
    tma_prefetch_desc(&tm);
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(kv_full0 + 8 * s, 1); mbar_init(kv_empty0 + 8 * s, 1);
      mbar_init(s_full0 + 8 * s, 1); mbar_init(s_free0 + 8 * s, 4);
    }
    mbar_init(p_full, 4);
    mbar_init(o_done, 1);
    fence_mbar_init();
  }
  if (warp == 5) { tmem_alloc(tmem_slot, kTcTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 4) {
    if (lane == 0) {
      // ---- Q: two 64-column halves, each 128 rows (two 64-row boxes stacked = one K-major 128B-swizzled tile) ----
      mbar_arrive_expect_tx(q_full, kTcQBytes);
      for (int half = 0; half < 2; ++half)
        for (int rb = 0; rb < 2; ++rb)
          tma_load_2d(sbase + kTcSmemQ + half * 16384 + rb * 8192, &tm, q_full, qcol + half * 64, row_base + q0 + rb * 64);
      for (int j = 0; j < nblk; ++j) {
        const int s = j & 1;
        mbar_wait(kv_empty0 + 8 * s, (((uint32_t)j >> 1) & 1u) ^ 1u);
        mbar_arrive_expect_tx(kv_full0 + 8 * s, 2 * kTcKVStage);
        const int r = row_base + j * kTcBN;
        for (int half = 0; half < 2; ++half) {
          tma_load_2d(sbase + kTcSmemK + s * kTcKVStage + half * 8192, &tm, kv_full0 + 8 * s, kcol + half * 64, r);
          tma_load_2d(sbase + kTcSmemV + s * kTcKVStage + half * 8192, &tm, kv_full0 + 8 * s, vcol + half * 64, r);
        }
      }
    }
  } else if (warp == 5) {
    {   // the whole warp runs this block (see umma_f16_w)
      // descriptors: hi word = SBO 1024 | version 1 | 128B swizzle; lo word = address >> 4 (| LBO << 16 for MN-major)
      constexpr uint32_t kHi = (1024u >> 4) | (1u << 14) | (kSwizzle128B << 29);
      constexpr uint32_t kLoK = (16u >> 4) << 16, kLoMN = (8192u >> 4) << 16;
      auto desc = [](uint32_t lo) { return ((uint64_t)kHi << 32) | (uint64_t)lo; };
      const uint32_t idesc_qk = make_idesc_f16(kTcBM, kTcBN, BF16, false, false);
      const uint32_t idesc_pv = make_idesc_f16(kTcBM, kTcD, BF16, false, true);
      const uint32_t tS = tmem_base, tO = tmem_base + kTcTmemO;
      const uint32_t qlo = ((sbase + kTcSmemQ) >> 4) | kLoK, plo = ((sbase + kTcSmemP) >> 4) | kLoK;
      auto issue_qk = [&](int j) {
        const int s = j & 1;
        const uint32_t klo = ((sbase + kTcSmemK + s * kTcKVStage) >> 4) | kLoK;
#pragma unroll
        for (int half = 0; half < 2; ++half)
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)
            umma_f16_w(tS + s * kTcBN, desc(qlo + half * (16384u >> 4) + ks * 2u), desc(klo + half * (8192u >> 4) + ks * 2u),
                     idesc_qk, (half | ks) ? 1u : 0u);
        umma_commit_w(s_full0 + 8 * s);
      };
      mbar_wait_spin(q_full, 0);
      mbar_wait_spin(kv_full0, 0);
      tc_fence_after();
      issue_qk(0);
      for (int j = 0; j < nblk; ++j) {
        const int s = j & 1;
        if (j + 1 < nblk) {
          mbar_wait_spin(kv_full0 + 8 * ((j + 1) & 1), ((uint32_t)(j + 1) >> 1) & 1u);
          // S buffer (j+1)&1 last held block j-1. PT: P(j-1) sat in it and PV(j-1) was issued before this UMMA, the pipe
          // runs in order. Otherwise: the softmax warps have pulled S(j-1) into registers.
          if (!PT && j >= 1) mbar_wait_spin(s_free0 + 8 * ((j + 1) & 1), ((uint32_t)(j - 1) >> 1) & 1u);
          tc_fence_after();
          issue_qk(j + 1);
        }
        mbar_wait_spin(p_full, (uint32_t)j & 1u);               // P(j) is written (and O rescaled where a row needed it)
        tc_fence_after();
        const uint32_t vlo = ((sbase + kTcSmemV + s * kTcKVStage) >> 4) | kLoMN;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          if (PT) umma_f16_ts_w(tO, tS + s * kTcBN + ks * 8, desc(vlo + ks * (2048u >> 4)), idesc_pv, (j | ks) ? 1u : 0u);
          else umma_f16_w(tO, desc(plo + ks * 2u), desc(vlo + ks * (2048u >> 4)), idesc_pv, (j | ks) ? 1u : 0u);
        }
        umma_commit_w(o_done);
        umma_commit_w(kv_empty0 + 8 * s);
      }
    }
  } else if (warp < 4) {
    // ================= softmax: thread = query row =================
    const int q = warp;                             // TMEM lane quarter this warp may touch (warp index mod 4)
    const int r = q * 32 + lane;                    // row inside the tile
    const int grow = q0 + r;                        // position in the sequence
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    const uint32_t tS = tmem_base + lane_addr, tO = tmem_base + lane_addr + kTcTmemO;
    const float sl = p.scale_log2;
    const int lim = p.causal ? min(grow, S - 1) : S - 1;      // last visible key of this row
    float m_ref = -INFINITY, l = 0.f;
    uint32_t v[64];
    for (int j = 0; j < nblk; ++j) {
      const int kv0 = j * kTcBN, sb = j & 1;
      const bool need_mask = (p.causal && kv0 + kTcBN - 1 > q0) || (kv0 + kTcBN > S);
      mbar_wait_warp(s_full0 + 8 * sb, ((uint32_t)j >> 1) & 1u);
      tc_fence_after();
      tmem_ld_32x32(tS + sb * kTcBN, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      tmem_ld_32x32(tS + sb * kTcBN + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
      tmem_ld_wait();
      if (!PT) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_free0 + 8 * sb);
      }
      float mx = -INFINITY;
      if (need_mask) {
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (kv0 + i > lim) v[i] = 0xff800000u;    // -inf
      }
#pragma unroll
      for (int i = 0; i < 64; ++i) mx = fmaxf(mx, __uint_as_float(v[i]));
      const float m_new = fmaxf(m_ref, mx * sl);
      // raise the reference max only when it is off by more than 2^8: exp2 of at most 8 stays exact enough in 16 bits
      const bool raise = m_new > m_ref + 8.f;
      float alpha = 1.f;
      if (raise) { alpha = ex2f(m_ref - m_new); m_ref = m_new; }
      const float m_use = (m_ref == -INFINITY) ? 0.f : m_ref;
      const bool any_raise = j > 0 && __any_sync(0xffffffffu, raise);
      if (j > 0 && (!PT || any_raise)) {            // PV(j-1) complete: O is stable (and the shared P tile is free)
        mbar_wait_warp(o_done, (uint32_t)(j - 1) & 1u);
        tc_fence_after();
      }
      if (any_raise) {
        uint32_t o[32];
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          tmem_ld_32x32(tO + c * 32, o);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
          tmem_st_32x32(tO + c * 32, o);
        }
      }
      l *= alpha;
      float rs0 = 0.f, rs1 = 0.f;
      const float nm = -m_use;
      if (PT) {
        uint32_t w[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float p0 = ex2f(fmaf(__uint_as_float(v[2 * i]), sl, nm)), p1 = ex2f(fmaf(__uint_as_float(v[2 * i + 1]), sl, nm));
          rs0 += p0; rs1 += p1;
          w[i] = pack16<BF16>(p0, p1);
        }
        tmem_st_32x32(tS + sb * kTcBN, w);
        tmem_st_wait();
        tc_fence_before();
      } else {
        const uint32_t prow = sbase + kTcSmemP + (uint32_t)r * 128u;
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          uint32_t o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int i0 = g * 8 + e * 2;
            const float p0 = ex2f(fmaf(__uint_as_float(v[i0]), sl, nm)), p1 = ex2f(fmaf(__uint_as_float(v[i0 + 1]), sl, nm));
            rs0 += p0; rs1 += p1;
            o[e] = pack16<BF16>(p0, p1);
          }
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(prow + (((uint32_t)g ^ ((uint32_t)r & 7u)) << 4)),
                       "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
        }
        if (any_raise) tmem_st_wait();
        tc_fence_before();
        fence_proxy_async_smem();                        // generic-proxy stores -> visible to tcgen05.mma
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
      l += rs0 + rs1;
    }
    // ---- finish: O / l -> 16 bit through shared memory (the Q tile is dead: every QK product has completed) ----
    mbar_wait_warp(o_done, (uint32_t)(nblk - 1) & 1u);
    tc_fence_after();
    const float inv = l > 0.f ? 1.f / l : 0.f;
    const uint32_t orow = sbase + kTcSmemQ + (uint32_t)r * 256u;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      tmem_ld_32x32(tO + c * 32, o);
      tmem_ld_wait();
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint32_t w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e)
          w[e] = pack16<BF16>(__uint_as_float(o[g * 8 + e * 2]) * inv, __uint_as_float(o[g * 8 + e * 2 + 1]) * inv);
        const uint32_t chunk = (uint32_t)(c * 4 + g);
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(orow + ((chunk ^ ((uint32_t)r & 15u)) << 4)), "r"(w[0]),
                     "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
      }
    }
    if (p.lse && grow < S) p.lse[((long long)b * p.H + h) * S + grow] = m_ref * 0.6931471805599453f + __logf(l);
    __syncwarp();
    // each warp copies ITS 32 rows out (written by its own lanes): 16 lanes cover the 256 bytes of a row
    uint16_t* obase = reinterpret_cast<uint16_t*>(p.out) + (long long)h * kTcD;
#pragma unroll 4
    for (int it = 0; it < 16; ++it) {
      const int rr = q * 32 + it * 2 + (lane >> 4), chunk = lane & 15;
      uint32_t w0, w1, w2, w3;
      asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3)
                   : "r"(sbase + kTcSmemQ + (uint32_t)rr * 256u + (((uint32_t)chunk ^ ((uint32_t)rr & 15u)) << 4)));
      if (q0 + rr < S)
        *reinterpret_cast<uint4*>(obase + (long long)(row_base + q0 + rr) * p.out_rs + chunk * 8) = make_uint4(w0, w1, w2, w3);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTcTmemCols);
  }
}

}  // namespace b200
using namespace b200;

// Causal (or full) attention forward on the tcgen05 path. head_dim 128, 16-bit, no dropout. qkv [B*S, (H + 2 Hkv) * 128],
// out [B*S, H * 128], lse [B, H, S] or NULL. Returns B200_ERR_UNSUPPORTED for shapes it does not cover (the caller then
// takes b200_flash_attn_fwd).
B200_API int b200_flash_attn_fwd_tc(const void* qkv, void* out, float* lse, int B, int S, int H, int Hkv, int D, int causal,
                                    float scale, int dtype, void* stream) {
  if (D != kTcD || (dtype != B200_DT_BF16 && dtype != B200_DT_FP16)) return B200_ERR_UNSUPPORTED;
  if (B < 1 || S < 1 || H < 1 || Hkv < 1 || H % Hkv) return B200_ERR_BAD_SHAPE;
  if ((reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(out)) & 15) return B200_ERR_BAD_ALIGN;
  // autograd's backward threads have no current driver context until a runtime call binds the primary one; the tensor-map
  // encode below is a driver entry point
  { int dev = 0; B200_CHECK(cudaGetDevice(&dev)); B200_CHECK(cudaSetDevice(dev)); }
  const long long W = (long long)(H + 2 * Hkv) * kTcD;
  CUtensorMap tm;
  {
    uint64_t dims[2] = {(uint64_t)W, (uint64_t)B * (uint64_t)S}, str[1] = {(uint64_t)W * 2};
    uint32_t box[2] = {64, 64};
    if (int rc = make_tensor_map(&tm, qkv, 2, 2, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  }
  FlashTcParams p = {};
  p.out = out; p.lse = lse; p.B = B; p.S = S; p.H = H; p.Hkv = Hkv; p.out_rs = (long long)H * kTcD;
  p.scale_log2 = scale * 1.4426950408889634f; p.causal = causal; p.is_bf16 = dtype == B200_DT_BF16;
  static const bool psmem = [] { const char* e = getenv("B200_FLASH_TC_PSMEM"); return e && atoi(e) != 0; }();
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(flash_fwd_tc_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTcSmemBytes));
    B200_CHECK(cudaFuncSetAttribute(flash_fwd_tc_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTcSmemBytes));
    B200_CHECK(cudaFuncSetAttribute(flash_fwd_tc_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTcSmemBytes));
    B200_CHECK(cudaFuncSetAttribute(flash_fwd_tc_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTcSmemBytes));
    attr = true;
  }
  dim3 grid((S + kTcBM - 1) / kTcBM, H, B);
  cudaStream_t st = (cudaStream_t)stream;
  if (psmem) {
    if (p.is_bf16) flash_fwd_tc_kernel<true, false><<<grid, kTcThreads, kTcSmemBytes, st>>>(tm, p);
    else flash_fwd_tc_kernel<false, false><<<grid, kTcThreads, kTcSmemBytes, st>>>(tm, p);
  } else {
    if (p.is_bf16) flash_fwd_tc_kernel<true, true><<<grid, kTcThreads, kTcSmemBytes, st>>>(tm, p);
    else flash_fwd_tc_kernel<false, true><<<grid, kTcThreads, kTcSmemBytes, st>>>(tm, p);
  }
  B200_LAUNCH_CHECK();
  return 0;
}
