// K4 backward on the 5th-generation tensor cores: dQ, dK, dV of causal flash attention with every product on tcgen05.mma,
// accumulators in TMEM, operands by TMA straight out of the fused qkv / dout matrices, no atomics and no fp32 gradient
// staging buffer.
//
// One kernel body, two modes (launched one after the other):
//   mode KV  CTA = (64-key block j, kv head, batch): K_j, V_j stay in shared memory; the (query head of the group, 128-row
//            query tile) pairs that see the block stream through a 2-stage TMA ring. TMEM holds dK_j^T and dV_j^T
//            ([head_dim lanes] x [64 key columns], accumulated over ALL those pairs, GQA group included).
//   mode Q   CTA = (128-row query tile i, query head, batch): Q_i, dO_i stay; the key blocks j <= i stream. TMEM holds dQ_i.
// Per (query tile, key block) pair, in both modes:
//     S  = Q K^T      (M 128 x N 64 x K 128)   -> TMEM, double-buffered            \  issued one pair AHEAD of the
//     dP = dO V^T     (M 128 x N 64 x K 128)   -> TMEM, double-buffered            /  element-wise warps
//     warps 0-7, thread = (query row, half of the keys):  P = exp2(S c - lse),  dS = P (dP - delta) * scale  -> two 16-bit
//                                                          128B-swizzled tiles in shared memory
//     mode KV:  dV^T += dO^T P   (M 128 (d) x N 64 x K 128 (rows), both operands MN-major, read in place)
//               dK^T += Q^T dS
//     mode Q :  dQ   += dS K     (M 128 x N 128 x K 64)
// S and dP are recomputed in mode Q (7 products instead of 5) in exchange for: no red.global traffic (the atomics of a
// one-pass kernel are ~64 KB per pair — 2 GB at 4k tokens x 32 heads), no zero-filled fp32 [tokens, 3 x heads x 128]
// buffer, bit-reproducible gradients.
//
// 320 threads: warps 0-7 element-wise (warp w and w + 4 share a TMEM lane quarter and split the 64 key columns), warp 8
// TMA producer, warp 9 tcgen05 issuer; 1 CTA per SM (384 TMEM columns, up to
// 193 KB of shared memory). head_dim 128, 16-bit, no dropout — anything else goes to the mma.sync kernel of
// flash_attn.cu. Reference call sites: the SDPA backward behind nn.MultiheadAttention
// (LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:89-96) and HF Qwen3 attention under the Trainer
// (Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:95-107, Fine-Tuning/qwen3-8b-lora-dist.py:118).
#include "common.cuh"
#include "tma_host.h"
#include "../include/b200_ops.h"

#include <stdio.h>
#include <stdlib.h>

namespace b200 {

constexpr int kBwBM = 128, kBwBN = 64, kBwD = 128;
constexpr int kBwThreads = 320;       // warps 0-7 element-wise (two per TMEM lane quarter), warp 8 TMA, warp 9 tcgen05 issuer
constexpr uint32_t kBwQTile = 32768;     // 128 rows x 128 columns x 2 B (two 64-column halves of 16 KB)
constexpr uint32_t kBwKTile = 16384;     // 64 rows x 128 columns x 2 B (two halves of 8 KB)
constexpr uint32_t kBwPTile = 16384;     // 128 rows x 64 keys x 2 B
constexpr uint32_t kBwTmemCols = 512;
constexpr uint32_t kBwTS = 0, kBwTdP = 64, kBwTBuf = 128, kBwTAcc = 256;   // S(b) at b*128, dP(b) at 64 + b*128, accumulators from 256

struct FlashBwdTcParams {
  void* dqkv;             // [B*S, W] 16-bit, same layout as qkv
  const float* lse;       // [B, H, S] natural log
  const float* delta;     // [B, H, S] rowsum(dO * O)
  int B, S, H, Hkv;
  long long W;            // row stride of qkv / dqkv in elements
  float scale, scale_log2;
  int causal;
};

// delta[b, h, s] = sum_d dO[b*S + s, h*128 + d] * O[...]: one warp per (row, head)
template <bool BF16>
__global__ void __launch_bounds__(256)
flash_delta_tc_kernel(const uint16_t* __restrict__ o, const uint16_t* __restrict__ dout, float* __restrict__ delta, int B, int S, int H) {
  const long long w = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (w >= (long long)B * S * H) return;
  const long long row = w / H;
  const int h = (int)(w % H);
  const uint2 a = *reinterpret_cast<const uint2*>(o + (row * H + h) * kBwD + lane * 4);
  const uint2 b = *reinterpret_cast<const uint2*>(dout + (row * H + h) * kBwD + lane * 4);
  float s;
  if (BF16) {
    const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), b0 = unpack_bf16x2(b.x), b1 = unpack_bf16x2(b.y);
    s = a0.x * b0.x + a0.y * b0.y + a1.x * b1.x + a1.y * b1.y;
  } else {
    const float2 a0 = __half22float2(*reinterpret_cast<const __half2*>(&a.x)), a1 = __half22float2(*reinterpret_cast<const __half2*>(&a.y));
    const float2 b0 = __half22float2(*reinterpret_cast<const __half2*>(&b.x)), b1 = __half22float2(*reinterpret_cast<const __half2*>(&b.y));
    s = a0.x * b0.x + a0.y * b0.y + a1.x * b1.x + a1.y * b1.y;
  }
  s = warp_sum(s);
  if (lane == 0) delta[((row / S) * H + h) * S + (row % S)] = s;
}

__device__ __forceinline__ float bw_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Whole-warp wait, one lane spinning on the non-blocking test_wait (try_wait may park the thread for a while: these
// hand-offs are on the critical path of every pair)
__device__ __forceinline__ void bw_wait_warp(uint32_t bar, uint32_t parity) {
  if ((threadIdx.x & 31) == 0) mbar_wait_spin(bar, parity);
  __syncwarp();
}

// tcgen05.mma / commit issued from CONVERGED warp code (every lane runs the loop, elect.sync picks the issuing lane): the
// descriptors stay in uniform registers. Issued from inside `if (lane == 0)` each UTCHMMA costs two R2UR moves and a
// divergence loop, ~100 cycles per instruction — more than these 128 x 64 x 16 products take on the tensor pipe.
__device__ __forceinline__ void umma_f16_w(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_w(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar) : "memory");
}

template <bool BF16>
__device__ __forceinline__ uint32_t bw_pack16(float a, float b) {
  if (BF16) return pack_bf16x2(a, b);
  const __half2 hh = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&hh);
}

// MODE_Q false: dK / dV of one 64-key block (K, V resident, query tiles stream). MODE_Q true: dQ of one 128-row query tile
// (Q, dO resident, key blocks stream). tmq: tensor map over qkv [B*S, W]; tmd: over dout [B*S, H*128]; both box {64, 64}.
template <bool BF16, bool MODE_Q>
__global__ void __launch_bounds__(kBwThreads, 1)
flash_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmd, const FlashBwdTcParams p) {
  extern __shared__ __align__(1024) uint8_t smem_bw[];
  const uint32_t sbase = smem_u32(smem_bw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = p.S, H = p.H, Hkv = p.Hkv, G = H / Hkv;
  const int b = blockIdx.z;
  const int row_base = b * S;
  const int nq = (S + kBwBM - 1) / kBwBM;

  // ---- what this CTA owns and what streams through it ----
  int kv0 = 0, hk = 0, q0 = 0, hq = 0, i_min = 0, nqt = 0, n_iter = 0;
  if (!MODE_Q) {
    kv0 = blockIdx.x * kBwBN; hk = blockIdx.y;
    i_min = p.causal ? kv0 / kBwBM : 0;
    nqt = nq - i_min;
    n_iter = G * nqt;
  } else {
    q0 = ((int)gridDim.x - 1 - (int)blockIdx.x) * kBwBM; hq = blockIdx.y; hk = hq / G;
    const int kv_end = p.causal ? min(S, q0 + kBwBM) : S;
    n_iter = (kv_end + kBwBN - 1) / kBwBN;
  }
  // shared memory map (bytes): resident tile(s) | 2 streaming stages | P | dS | barriers
  constexpr uint32_t kRes = MODE_Q ? 2 * kBwQTile : 2 * kBwKTile;
  constexpr uint32_t kStage = MODE_Q ? 2 * kBwKTile : 2 * kBwQTile;
  constexpr int NST = MODE_Q ? 3 : 2;      // streaming stages: 3 x 32 KB next to the resident Q / dO, 2 x 64 KB next to K / V
  constexpr uint32_t kOffP = 163840, kOffDS = 180224, kOffBar = 196608;
  const uint32_t sRes = sbase, sStage0 = sbase + kRes, sP = sbase + kOffP, sDS = sbase + kOffDS;
  const uint32_t bar = sbase + kOffBar;
  const uint32_t res_full = bar, st_full0 = bar + 8, st_empty0 = bar + 40, s_full0 = bar + 72, s_free0 = bar + 88,
                 pds_full = bar + 104, pds_free = bar + 112, acc_done = bar + 120, tmem_slot = bar + 128;
  if (threadIdx.x == 0) {
    if (sbase & 1023u) { printf("flash_bwd_tc: shared memory base is not 1024-byte aligned\n"); __trap(); }
    tma_prefetch_desc(&tmq);
    tma_prefetch_desc(&tmd);
    mbar_init(res_full, 1);
    for (int s = 0; s < NST; ++s) { mbar_init(st_full0 + 8 * s, 1); mbar_init(st_empty0 + 8 * s, 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(s_full0 + 8 * s, 1); mbar_init(s_free0 + 8 * s, 8); }
    mbar_init(pds_full, 8);
    mbar_init(pds_free, 1);
    mbar_init(acc_done, 1);
    fence_mbar_init();
  }
  if (warp == 9) { tmem_alloc(tmem_slot, kBwTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  // query tile / head / key block of streaming iteration n
  auto pair_of = [&](int n, int& qh, int& qrow0, int& kvrow0) {
    if (!MODE_Q) { const int g = n / nqt; qh = hk * G + g; qrow0 = (i_min + n - g * nqt) * kBwBM; kvrow0 = kv0; }
    else { qh = hq; qrow0 = q0; kvrow0 = n * kBwBN; }
  };

  if (warp == 8) {
    if (lane == 0) {
      // 128-row x 128-column tile (Q or dO): two 64-column halves, each two 64-row boxes; 64-row tile (K or V): two boxes
      auto load_q = [&](uint32_t dst, const CUtensorMap* tm, uint32_t fb, int col, int row) {
        for (int half = 0; half < 2; ++half)
          for (int rb = 0; rb < 2; ++rb) tma_load_2d(dst + half * 16384 + rb * 8192, tm, fb, col + half * 64, row + rb * 64);
      };
      auto load_k = [&](uint32_t dst, uint32_t fb, int col, int row) {
        for (int half = 0; half < 2; ++half) tma_load_2d(dst + half * 8192, &tmq, fb, col + half * 64, row);
      };
      mbar_arrive_expect_tx(res_full, kRes);
      if (!MODE_Q) {
        load_k(sRes, res_full, (H + hk) * kBwD, row_base + kv0);
        load_k(sRes + kBwKTile, res_full, (H + Hkv + hk) * kBwD, row_base + kv0);
      } else {
        load_q(sRes, &tmq, res_full, hq * kBwD, row_base + q0);
        load_q(sRes + kBwQTile, &tmd, res_full, hq * kBwD, row_base + q0);
      }
      for (int n = 0; n < n_iter; ++n) {
        const int s = n % NST;
        int qh, qr, kr;
        pair_of(n, qh, qr, kr);
        mbar_wait(st_empty0 + 8 * s, (((uint32_t)(n / NST)) & 1u) ^ 1u);
        mbar_arrive_expect_tx(st_full0 + 8 * s, kStage);
        const uint32_t dst = sStage0 + s * kStage;
        if (!MODE_Q) {
          load_q(dst, &tmq, st_full0 + 8 * s, qh * kBwD, row_base + qr);
          load_q(dst + kBwQTile, &tmd, st_full0 + 8 * s, qh * kBwD, row_base + qr);
        } else {
          load_k(dst, st_full0 + 8 * s, (H + hk) * kBwD, row_base + kr);
          load_k(dst + kBwKTile, st_full0 + 8 * s, (H + Hkv + hk) * kBwD, row_base + kr);
        }
      }
    }
  } else if (warp == 9) {
    {   // the whole warp runs this block (see umma_f16_w)
      constexpr uint32_t kHi = (1024u >> 4) | (1u << 14) | (kSwizzle128B << 29);
      constexpr uint32_t kLoK = (16u >> 4) << 16;
      constexpr uint32_t kLoMNq = (16384u >> 4) << 16;   // MN-major 128-row tile: 64-column chunks 16 KB apart
      constexpr uint32_t kLoMNk = (8192u >> 4) << 16;    // MN-major 64-row tile: chunks 8 KB apart
      constexpr uint32_t kStepMN = 2048u >> 4;           // 16 rows of 128 B
      auto desc = [](uint32_t lo) { return ((uint64_t)kHi << 32) | (uint64_t)lo; };
      const uint32_t idesc_s = make_idesc_f16(kBwBM, kBwBN, BF16, false, false);
      const uint32_t idesc_kv = make_idesc_f16(kBwD, kBwBN, BF16, true, true);      // dV^T / dK^T: both MN-major
      const uint32_t idesc_q = make_idesc_f16(kBwBM, kBwD, BF16, false, true);      // dQ: dS K-major, K MN-major
      // operand addresses of iteration n
      auto q_of = [&](int n) { return MODE_Q ? sRes : sStage0 + (uint32_t)(n % NST) * kStage; };
      auto do_of = [&](int n) { return q_of(n) + kBwQTile; };
      auto k_of = [&](int n) { return MODE_Q ? sStage0 + (uint32_t)(n % NST) * kStage : sRes; };
      auto v_of = [&](int n) { return k_of(n) + kBwKTile; };
      auto issue_sdp = [&](int n) {
        const uint32_t buf = (uint32_t)(n & 1) * kBwTBuf;
        const uint32_t ql = (q_of(n) >> 4) | kLoK, kl = (k_of(n) >> 4) | kLoK, dl = (do_of(n) >> 4) | kLoK, vl = (v_of(n) >> 4) | kLoK;
#pragma unroll
        for (int half = 0; half < 2; ++half)
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)
            umma_f16_w(tmem_base + kBwTS + buf, desc(ql + half * (16384u >> 4) + ks * 2u), desc(kl + half * (8192u >> 4) + ks * 2u),
                     idesc_s, (half | ks) ? 1u : 0u);
#pragma unroll
        for (int half = 0; half < 2; ++half)
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)
            umma_f16_w(tmem_base + kBwTdP + buf, desc(dl + half * (16384u >> 4) + ks * 2u), desc(vl + half * (8192u >> 4) + ks * 2u),
                     idesc_s, (half | ks) ? 1u : 0u);
        umma_commit_w(s_full0 + 8 * (n & 1));
      };
      mbar_wait_spin(res_full, 0);
      mbar_wait_spin(st_full0, 0);
      tc_fence_after();
      issue_sdp(0);
      for (int n = 0; n < n_iter; ++n) {
        if (n + 1 < n_iter) {
          mbar_wait_spin(st_full0 + 8 * ((n + 1) % NST), ((uint32_t)((n + 1) / NST)) & 1u);
          if (n >= 1) mbar_wait_spin(s_free0 + 8 * ((n + 1) & 1), ((uint32_t)(n - 1) >> 1) & 1u);   // S / dP of pair n-1 are in registers
          tc_fence_after();
          issue_sdp(n + 1);
        }
        mbar_wait_spin(pds_full, (uint32_t)n & 1u);           // P and dS of pair n are in shared memory
        tc_fence_after();
        const uint32_t pl = sP >> 4, dsl = sDS >> 4;
        if (!MODE_Q) {
          const uint32_t dol = (do_of(n) >> 4) | kLoMNq, ql = (q_of(n) >> 4) | kLoMNq;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma_f16_w(tmem_base + kBwTAcc + 64, desc(dol + ks * kStepMN), desc((pl | kLoMNq) + ks * kStepMN), idesc_kv, (n | ks) ? 1u : 0u);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma_f16_w(tmem_base + kBwTAcc, desc(ql + ks * kStepMN), desc((dsl | kLoMNq) + ks * kStepMN), idesc_kv, (n | ks) ? 1u : 0u);
        } else {
          const uint32_t kl = (k_of(n) >> 4) | kLoMNk;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)
            umma_f16_w(tmem_base + kBwTAcc, desc((dsl | kLoK) + ks * 2u), desc(kl + ks * kStepMN), idesc_q, (n | ks) ? 1u : 0u);
        }
        umma_commit_w(pds_free);
        umma_commit_w(st_empty0 + 8 * (n % NST));
      }
      umma_commit_w(acc_done);
    }
  } else if (warp < 8) {
    // ================= element-wise: thread = (query row, half of the key columns) =================
    const int wq = warp & 3, c = warp >> 2;        // TMEM lane quarter, 32-column half of S / dP
    const int r = wq * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
    const float sl = p.scale_log2, sc = p.scale;
    // per-row softmax statistics of the pair: loaded one pair ahead (in mode Q they never change)
    auto row_stats = [&](int n, float& lse2, float& dl) {
      int qh, qr, kr;
      pair_of(n, qh, qr, kr);
      lse2 = INFINITY; dl = 0.f;
      if (qr + r < S) {
        const long long o = ((long long)b * H + qh) * S + qr + r;
        lse2 = p.lse[o] * 1.4426950408889634f;
        dl = p.delta[o];
      }
    };
    float lse2_next, dl_next;
    row_stats(0, lse2_next, dl_next);
    for (int n = 0; n < n_iter; ++n) {
      int qh, qr, kr;
      pair_of(n, qh, qr, kr);
      const int grow = qr + r;
      const float lse2 = lse2_next, dl = dl_next;
      if (!MODE_Q && n + 1 < n_iter) row_stats(n + 1, lse2_next, dl_next);
      const bool need_mask = (p.causal && kr + kBwBN - 1 > qr) || (kr + kBwBN > S);
      const int lim = p.causal ? min(grow, S - 1) : S - 1;
      const uint32_t buf = (uint32_t)(n & 1) * kBwTBuf;
      bw_wait_warp(s_full0 + 8 * (n & 1), ((uint32_t)n >> 1) & 1u);
      tc_fence_after();
      const uint32_t prow = sP + (uint32_t)r * 128u, dsrow = sDS + (uint32_t)r * 128u;
      {
        uint32_t sv[32], dv[32];
        tmem_ld_32x32(tmem_base + lane_addr + kBwTS + buf + c * 32, sv);
        tmem_ld_32x32(tmem_base + lane_addr + kBwTdP + buf + c * 32, dv);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_free0 + 8 * (n & 1));
        uint32_t pw[16], dw[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float p0 = bw_ex2(fmaf(__uint_as_float(sv[2 * i]), sl, -lse2)), p1 = bw_ex2(fmaf(__uint_as_float(sv[2 * i + 1]), sl, -lse2));
          if (need_mask) {
            if (kr + c * 32 + 2 * i > lim) p0 = 0.f;
            if (kr + c * 32 + 2 * i + 1 > lim) p1 = 0.f;
          }
          pw[i] = bw_pack16<BF16>(p0, p1);
          dw[i] = bw_pack16<BF16>(p0 * (__uint_as_float(dv[2 * i]) - dl) * sc, p1 * (__uint_as_float(dv[2 * i + 1]) - dl) * sc);
        }
        if (n > 0) bw_wait_warp(pds_free, (uint32_t)(n - 1) & 1u);    // the products of pair n-1 have read P / dS
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const uint32_t off = (((uint32_t)(c * 4 + g)) ^ ((uint32_t)r & 7u)) << 4;
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(prow + off), "r"(pw[4 * g]), "r"(pw[4 * g + 1]),
                       "r"(pw[4 * g + 2]), "r"(pw[4 * g + 3]) : "memory");
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(dsrow + off), "r"(dw[4 * g]), "r"(dw[4 * g + 1]),
                       "r"(dw[4 * g + 2]), "r"(dw[4 * g + 3]) : "memory");
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(pds_full);
    }
    // ---- write the accumulators out ----
    bw_wait_warp(acc_done, 0);
    tc_fence_after();
    uint16_t* dq = reinterpret_cast<uint16_t*>(p.dqkv);
    if (!MODE_Q) {
      // TMEM lane = head-dim index d (this thread), column = key: dK^T at kBwTAcc, dV^T at kBwTAcc + 64
      {
        const int which = c;                       // warps 0-3 write dK, warps 4-7 dV
        const long long col = (long long)(which == 0 ? (H + hk) : (H + Hkv + hk)) * kBwD + r;
#pragma unroll 1
        for (int cc = 0; cc < 2; ++cc) {
          uint32_t v[32];
          tmem_ld_32x32(tmem_base + lane_addr + kBwTAcc + which * 64 + cc * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int key = kv0 + cc * 32 + i;
            if (key < S) {
              const float x = __uint_as_float(v[i]);
              uint16_t o;
              if (BF16) { const __nv_bfloat16 t = __float2bfloat16_rn(x); o = *reinterpret_cast<const uint16_t*>(&t); }
              else { const __half t = __float2half_rn(x); o = *reinterpret_cast<const uint16_t*>(&t); }
              dq[(long long)(row_base + key) * p.W + col] = o;
            }
          }
        }
      }
    } else {
      // dQ rows through shared memory (the resident Q tile is dead) for 16-byte coalesced stores
      const uint32_t orow = sRes + (uint32_t)r * 256u;
#pragma unroll 1
      for (int cc = 2 * c; cc < 2 * c + 2; ++cc) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + lane_addr + kBwTAcc + cc * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t w[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) w[e] = bw_pack16<BF16>(__uint_as_float(v[g * 8 + e * 2]), __uint_as_float(v[g * 8 + e * 2 + 1]));
          const uint32_t chunk = (uint32_t)(cc * 4 + g);
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(orow + ((chunk ^ ((uint32_t)r & 15u)) << 4)), "r"(w[0]),
                       "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
        }
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");      // both column halves of every row are in shared memory
#pragma unroll 4
      for (int it = 0; it < 8; ++it) {
        const int rr = warp * 16 + it * 2 + (lane >> 4), chunk = lane & 15;
        uint32_t w0, w1, w2, w3;
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3)
                     : "r"(sRes + (uint32_t)rr * 256u + (((uint32_t)chunk ^ ((uint32_t)rr & 15u)) << 4)));
        if (q0 + rr < S)
          *reinterpret_cast<uint4*>(dq + (long long)(row_base + q0 + rr) * p.W + (long long)hq * kBwD + chunk * 8) = make_uint4(w0, w1, w2, w3);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kBwTmemCols);
  }
}

}  // namespace b200
using namespace b200;

// dqkv [B*S, (H + 2 Hkv) * 128] = gradient of the fused projection for dout [B*S, H * 128] (16-bit). out / lse are the
// forward's outputs (lse natural log, [B, H, S]); delta [B, H, S] is a caller-provided fp32 work buffer. head_dim 128, no
// dropout; B200_ERR_UNSUPPORTED otherwise (callers then use b200_flash_attn_bwd).
B200_API int b200_flash_attn_bwd_tc(const void* qkv, const void* out, const void* dout, const float* lse, float* delta, void* dqkv,
                                    int B, int S, int H, int Hkv, int D, int causal, float scale, int dtype, void* stream) {
  if (D != kBwD || (dtype != B200_DT_BF16 && dtype != B200_DT_FP16)) return B200_ERR_UNSUPPORTED;
  if (B < 1 || S < 1 || H < 1 || Hkv < 1 || H % Hkv) return B200_ERR_BAD_SHAPE;
  if ((reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(dout) |
       reinterpret_cast<uintptr_t>(dqkv)) & 15) return B200_ERR_BAD_ALIGN;
  // autograd's backward threads have no current driver context until a runtime call binds the primary one; the tensor-map
  // encode below is a driver entry point
  { int dev = 0; B200_CHECK(cudaGetDevice(&dev)); B200_CHECK(cudaSetDevice(dev)); }
  const long long W = (long long)(H + 2 * Hkv) * kBwD;
  CUtensorMap tmq, tmd;
  {
    uint64_t dims[2] = {(uint64_t)W, (uint64_t)B * (uint64_t)S}, str[1] = {(uint64_t)W * 2};
    uint32_t box[2] = {64, 64};
    if (int rc = make_tensor_map(&tmq, qkv, 2, 2, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
    uint64_t dims2[2] = {(uint64_t)H * kBwD, (uint64_t)B * (uint64_t)S}, str2[1] = {(uint64_t)H * kBwD * 2};
    if (int rc = make_tensor_map(&tmd, dout, 2, 2, dims2, str2, box, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  }
  FlashBwdTcParams p = {};
  p.dqkv = dqkv; p.lse = lse; p.delta = delta; p.B = B; p.S = S; p.H = H; p.Hkv = Hkv; p.W = W;
  p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f; p.causal = causal;
  const bool bf16 = dtype == B200_DT_BF16;
  constexpr int kSmem = 196608 + 256;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(flash_bwd_tc_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    B200_CHECK(cudaFuncSetAttribute(flash_bwd_tc_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    B200_CHECK(cudaFuncSetAttribute(flash_bwd_tc_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    B200_CHECK(cudaFuncSetAttribute(flash_bwd_tc_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    attr = true;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const long long warps = (long long)B * S * H;
  const unsigned dgrid = (unsigned)((warps + 7) / 8);
  if (bf16) flash_delta_tc_kernel<true><<<dgrid, 256, 0, st>>>((const uint16_t*)out, (const uint16_t*)dout, delta, B, S, H);
  else flash_delta_tc_kernel<false><<<dgrid, 256, 0, st>>>((const uint16_t*)out, (const uint16_t*)dout, delta, B, S, H);
  B200_LAUNCH_CHECK();
  dim3 gkv((S + kBwBN - 1) / kBwBN, Hkv, B), gq((S + kBwBM - 1) / kBwBM, H, B);
  if (bf16) {
    flash_bwd_tc_kernel<true, false><<<gkv, kBwThreads, kSmem, st>>>(tmq, tmd, p);
    flash_bwd_tc_kernel<true, true><<<gq, kBwThreads, kSmem, st>>>(tmq, tmd, p);
  } else {
    flash_bwd_tc_kernel<false, false><<<gkv, kBwThreads, kSmem, st>>>(tmq, tmd, p);
    flash_bwd_tc_kernel<false, true><<<gq, kBwThreads, kSmem, st>>>(tmq, tmd, p);
  }
  B200_LAUNCH_CHECK();
  return 0;
}
