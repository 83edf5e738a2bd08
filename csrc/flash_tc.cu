// K4 on the 5th-generation tensor cores: causal flash-attention forward with tcgen05.mma, S and PV accumulators in
// TMEM, Q / K / V tiles brought in by TMA straight out of the fused qkv projection (no head transposes, GQA by
// column offset), online softmax by four warps whose threads each own ONE query row (TMEM lane = row, so row max and
// row sum need no shuffles).
//
//   CTA = (128-query tile, head, batch); 256 threads: warps 0-3 softmax (one warpgroup), warp 4 TMA producer, warp 5
//   tcgen05 issuer (setmaxnreg is a per-warpgroup instruction: the roles that give registers away and the ones that
//   take them sit in different, 4-warp-aligned groups).
//   per 64-key block j:   S = Q K_j^T           (M 128 x N 64 x K 128, 8 UMMAs)        -> TMEM columns [0, 64)
//                         softmax warps: tcgen05.ld S, m / l update, P = exp2(S c - m c) -> 16-bit, 128B-swizzled smem tile
//                         PV = P V_j             (M 128 x N 128 x K 64, 4 UMMAs, V read MN-major in place) -> TMEM [64, 192)
//                         softmax warps: O = O * alpha + tcgen05.ld(PV)                  (O lives in registers, fp32)
//   S(j+1) is issued as soon as the softmax warps have pulled S(j) into registers, so the tensor pipe works on the next
//   scores while exp2 runs; two CTAs fit on an SM (115 KB of shared memory, 256 TMEM columns each) and overlap further.
//   The register file is split with setmaxnreg: 40 per thread for the producer warpgroup, 216 for the softmax rows.
//
// Replaces the mma.sync kernel of flash_attn.cu for head_dim 128 forward passes (training forward, prefill); that
// kernel stays for other head sizes, dropout and the backward pass. Reference call sites: SDPA behind
// nn.MultiheadAttention (LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:89-96) and HF Qwen3 attention
// (Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:95-107).
#include "common.cuh"
#include "tma_host.h"
#include "../include/b200_ops.h"

#include <stdio.h>

namespace b200 {

constexpr int kTcBM = 128, kTcBN = 64, kTcD = 128;
constexpr int kTcThreads = 256;
constexpr uint32_t kTcQBytes = 32768, kTcKVStage = 16384, kTcPBytes = 16384;
constexpr uint32_t kTcSmemQ = 0, kTcSmemK = 32768, kTcSmemV = 65536, kTcSmemP = 98304, kTcSmemBar = 114688;
constexpr uint32_t kTcSmemBytes = kTcSmemBar + 256;
constexpr uint32_t kTcTmemCols = 256, kTcTmemO = 64;

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

template <bool BF16>
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
  const uint32_t q_full = bar, kv_full0 = bar + 8, kv_empty0 = bar + 24, s_full = bar + 40, s_free = bar + 48,
                 p_full = bar + 56, o_full = bar + 64, o_free = bar + 72, tmem_slot = bar + 80;
  if (threadIdx.x == 0) {
    if (sbase & 1023u) { printf("flash_tc: shared memory base is not 1024-byte aligned\n"); __trap(); }
    tma_prefetch_desc(&tm);
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(kv_full0 + 8 * s, 1); mbar_init(kv_empty0 + 8 * s, 1); }
    mbar_init(s_full, 1);
    mbar_init(s_free, 4);
    mbar_init(p_full, 4);
    mbar_init(o_full, 1);
    mbar_init(o_free, 4);
    fence_mbar_init();
  }
  if (warp == 5) { tmem_alloc(tmem_slot, kTcTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp >= 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    if (warp == 4 && lane == 0) {
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
    } else if (warp == 5 && lane == 0) {
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
            umma_f16(tS, desc(qlo + half * (16384u >> 4) + ks * 2u), desc(klo + half * (8192u >> 4) + ks * 2u), idesc_qk,
                     (half | ks) ? 1u : 0u);
        umma_commit(s_full);
      };
      mbar_wait(q_full, 0);
      mbar_wait(kv_full0, 0);
      tc_fence_after();
      issue_qk(0);
      for (int j = 0; j < nblk; ++j) {
        const int s = j & 1;
        if (j + 1 < nblk) {
          mbar_wait(kv_full0 + 8 * ((j + 1) & 1), ((uint32_t)(j + 1) >> 1) & 1u);
          mbar_wait(s_free, (uint32_t)j & 1u);            // the softmax warps hold S(j) in registers
          tc_fence_after();
          issue_qk(j + 1);
        }
        mbar_wait(p_full, (uint32_t)j & 1u);               // P(j) is in shared memory
        mbar_wait(o_free, ((uint32_t)j & 1u) ^ 1u);        // PV(j-1) has been read out of TMEM
        tc_fence_after();
        const uint32_t vlo = ((sbase + kTcSmemV + s * kTcKVStage) >> 4) | kLoMN;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
          umma_f16(tO, desc(plo + ks * 2u), desc(vlo + ks * (2048u >> 4)), idesc_pv, ks ? 1u : 0u);
        umma_commit(o_full);
        umma_commit(kv_empty0 + 8 * s);
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    // ================= softmax / accumulation: thread = query row =================
    const int q = warp;                             // TMEM lane quarter this warp may touch (warp index mod 4)
    const int r = q * 32 + lane;                    // row inside the tile
    const int grow = q0 + r;                        // position in the sequence
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    const uint32_t tS = tmem_base + lane_addr, tO = tmem_base + lane_addr + kTcTmemO;
    const float sl = p.scale_log2;
    float O[kTcD];
#pragma unroll
    for (int i = 0; i < kTcD; ++i) O[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    for (int j = 0; j < nblk; ++j) {
      const int kv0 = j * kTcBN;
      const bool need_mask = (p.causal && kv0 + kTcBN - 1 > q0) || (kv0 + kTcBN > S);
      const int lim = p.causal ? min(grow, S - 1) : S - 1;      // last visible key of this row
      mbar_wait(s_full, (uint32_t)j & 1u);
      tc_fence_after();
      // pass 1: row maximum (S is read twice out of TMEM instead of held in 64 registers next to the 128 of O)
      float mx = -INFINITY;
      uint32_t v[32];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        tmem_ld_32x32(tS + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float x = __uint_as_float(v[i]);
          if (need_mask && kv0 + c * 32 + i > lim) x = -INFINITY;
          mx = fmaxf(mx, x);
        }
      }
      const float m_new = fmaxf(m, mx * sl);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;    // a row that sees no key yet (only possible beyond S)
      const float alpha = ex2f(m - m_use);
      float rs = 0.f;
      const uint32_t prow = sbase + kTcSmemP + (uint32_t)r * 128u;
      // pass 2: probabilities -> 16-bit, straight into the swizzled A tile of the PV product
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        tmem_ld_32x32(tS + c * 32, v);
        tmem_ld_wait();
        if (c == 1) {                                   // S(j) is not needed in TMEM any more: the next QK may overwrite it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(s_free);
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int i0 = g * 8 + e * 2;
            float x0 = __uint_as_float(v[i0]) * sl - m_use, x1 = __uint_as_float(v[i0 + 1]) * sl - m_use;
            if (need_mask) {
              if (kv0 + c * 32 + i0 > lim) x0 = -INFINITY;
              if (kv0 + c * 32 + i0 + 1 > lim) x1 = -INFINITY;
            }
            const float p0 = ex2f(x0), p1 = ex2f(x1);
            if (BF16) o[e] = pack_bf16x2(p0, p1);
            else { const __half2 hh = __floats2half2_rn(p0, p1); o[e] = *reinterpret_cast<const uint32_t*>(&hh); }
            // the row sum uses the rounded probabilities, i.e. exactly what the tensor core multiplies
            float2 f;
            if (BF16) f = unpack_bf16x2(o[e]);
            else f = __half22float2(*reinterpret_cast<const __half2*>(&o[e]));
            rs += f.x + f.y;
          }
          const uint32_t chunk = (uint32_t)(c * 4 + g);
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(prow + ((chunk ^ ((uint32_t)r & 7u)) << 4)), "r"(o[0]),
                       "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
        }
      }
      fence_proxy_async_smem();                          // generic-proxy stores -> visible to tcgen05.mma
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
      l = l * alpha + rs;
      m = m_use;
      // O = O * alpha + PV(j)
      mbar_wait(o_full, (uint32_t)j & 1u);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        tmem_ld_32x32(tO + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) O[c * 32 + i] = O[c * 32 + i] * alpha + __uint_as_float(v[i]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(o_free);
    }
    // ---- finish: O / l -> 16 bit, lse ----
    if (grow < S) {
      const float inv = l > 0.f ? 1.f / l : 0.f;
      uint16_t* orow = reinterpret_cast<uint16_t*>(p.out) + (long long)(row_base + grow) * p.out_rs + (long long)h * kTcD;
#pragma unroll
      for (int c = 0; c < kTcD / 8; ++c) {
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float a0 = O[c * 8 + e * 2] * inv, a1 = O[c * 8 + e * 2 + 1] * inv;
          if (BF16) o[e] = pack_bf16x2(a0, a1);
          else { const __half2 hh = __floats2half2_rn(a0, a1); o[e] = *reinterpret_cast<const uint32_t*>(&hh); }
        }
        reinterpret_cast<uint4*>(orow)[c] = make_uint4(o[0], o[1], o[2], o[3]);
      }
      if (p.lse) p.lse[((long long)b * p.H + h) * S + grow] = m * 0.6931471805599453f + __logf(l);
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
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(flash_fwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTcSmemBytes));
    B200_CHECK(cudaFuncSetAttribute(flash_fwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTcSmemBytes));
    attr = true;
  }
  dim3 grid((S + kTcBM - 1) / kTcBM, H, B);
  cudaStream_t st = (cudaStream_t)stream;
  if (p.is_bf16) flash_fwd_tc_kernel<true><<<grid, kTcThreads, kTcSmemBytes, st>>>(tm, p);
  else flash_fwd_tc_kernel<false><<<grid, kTcThreads, kTcSmemBytes, st>>>(tm, p);
  B200_LAUNCH_CHECK();
  return 0;
}
