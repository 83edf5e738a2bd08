// K1 / K3 — persistent, warp-specialised tcgen05 GEMM for sm_100a.
//
//   D[b] = alpha * A[b] x B[b / b_div]^T  (+ A2 x B2^T)  (+ bias)  + beta * D[b]
//
// One CTA per SM loops over 128 x BN output tiles. Warp 0 feeds a ring of shared-memory stages
// with TMA (128-byte swizzle), warp 1 issues tcgen05.mma into one of two TMEM accumulators,
// warps 2-5 drain the other accumulator with tcgen05.ld and write D, so the epilogue of tile i
// overlaps the main loop of tile i+1. BN (32..256) and the stage count are run-time values: the
// host picks the BN that wastes the fewest SM-waves for the problem at hand.
//
// Operands are read in place in either storage order ("K-major" or "MN-major" UMMA descriptors),
// so forward (X W^T), dgrad (dY W) and wgrad (dY^T X) all run through this one kernel with no
// transpose copies, for bf16/fp16 (kind::f16) and fp32 (kind::tf32) inputs. The optional A2/B2
// operands append a rank-K2 update as one extra k-block: that is the LoRA term of peft's
// lora.Linear (Fine-Tuning/qwen3-8b-lora-dist.py:118-125), computed on the tensor cores inside
// the same accumulator instead of as three extra cuBLAS calls.
//
// Replaces: cuBLAS/cuBLASLt behind every nn.Linear in the reference, e.g.
// LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:86-165.
#include "common.cuh"
#include "tma_host.h"
#include "../include/b200_ops.h"

#include <stdio.h>

namespace b200 {

constexpr int kBM = 128;             // tile rows = TMEM lanes
constexpr int kStageABytes = 16384;  // 128 rows x 128 B
constexpr int kGemmThreads = 192;    // producer, mma, 4 epilogue warps
constexpr int kAccStride = 256;      // TMEM columns between the two accumulators

struct GemmKernelParams {
  void* D;
  const void* bias;
  int M, N, K, K2;
  int batch, nb0, b_div, nb0_b;
  long long d_rs, d_bs0, d_bs1;
  float alpha, beta;
  int out_dtype, bias_dtype;
  uint32_t idesc_main, idesc_lora;
  int num_m, num_n, num_k;  // k-blocks of the main operands
  int bn, stages, a_mn, b_mn;
  int vec_ok;
};

template <int ES>
__device__ __forceinline__ void umma_issue(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  if constexpr (ES == 2) umma_f16(d, a, b, idesc, acc);
  else umma_tf32(d, a, b, idesc, acc);
}

template <typename OutT>
__device__ __forceinline__ void epilogue_store(const GemmKernelParams& p, OutT* __restrict__ drow, int col0,
                                               const uint32_t (&r)[32], bool row_ok) {
  // drow points at D[row, 0]; this thread owns columns col0 .. col0+31 of its row.
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;
  if (p.bias != nullptr) {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      int c = col0 + j;
      if (c < p.N) {
        float b = p.bias_dtype == B200_DT_FP32 ? reinterpret_cast<const float*>(p.bias)[c]
                  : p.bias_dtype == B200_DT_BF16
                      ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.bias)[c])
                      : __half2float(reinterpret_cast<const __half*>(p.bias)[c]);
        v[j] += b;
      }
    }
  }
  if (!row_ok) return;
  const bool full = (col0 + 32 <= p.N) && p.vec_ok;
  if (full) {
    constexpr int kPerVec = 16 / sizeof(OutT);
    constexpr int kVecs = 32 / kPerVec;
    uint4* dst = reinterpret_cast<uint4*>(drow + col0);
    if (p.beta != 0.f) {
#pragma unroll
      for (int i = 0; i < kVecs; ++i) {
        uint4 old = dst[i];
        const OutT* o = reinterpret_cast<const OutT*>(&old);
#pragma unroll
        for (int j = 0; j < kPerVec; ++j) v[i * kPerVec + j] += p.beta * to_f32<OutT>(o[j]);
      }
    }
#pragma unroll
    for (int i = 0; i < kVecs; ++i) {
      uint4 out;
      OutT* o = reinterpret_cast<OutT*>(&out);
#pragma unroll
      for (int j = 0; j < kPerVec; ++j) o[j] = from_f32<OutT>(v[i * kPerVec + j]);
      dst[i] = out;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      int c = col0 + j;
      if (c < p.N) {
        float x = v[j];
        if (p.beta != 0.f) x += p.beta * to_f32<OutT>(drow[c]);
        drow[c] = from_f32<OutT>(x);
      }
    }
  }
}

template <int ES>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                    const GemmKernelParams p) {
  constexpr int EPR = 128 / ES;  // elements per 128-byte swizzle row = k-block depth
  constexpr int UK = 32 / ES;    // elements per UMMA k-step

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t stage_b_bytes = (uint32_t)p.bn * 128u;
  const uint32_t sA0 = smem_base;
  const uint32_t sB0 = smem_base + (uint32_t)p.stages * kStageABytes;
  const uint32_t bar0 = sB0 + (uint32_t)p.stages * stage_b_bytes;
  // barrier block: full[stages] | empty[stages] | tmem_full[2] | tmem_empty[2] | tmem_ptr
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (p.stages + s); };
  auto tfull_bar = [&](int a) { return bar0 + 8u * (2 * p.stages + a); };
  auto tempty_bar = [&](int a) { return bar0 + 8u * (2 * p.stages + 2 + a); };
  const uint32_t tmem_slot = bar0 + 8u * (2 * p.stages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.K2 > 0) {
      tma_prefetch_desc(&tmA2);
      tma_prefetch_desc(&tmB2);
    }
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 4);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const int tiles_per_batch = p.num_m * p.num_n;
  const int total_tiles = tiles_per_batch * p.batch;
  const int num_k_total = p.num_k + (p.K2 > 0 ? 1 : 0);

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t tx_bytes = kStageABytes + stage_b_bytes;
      const uint32_t a_box_bytes = EPR * 128;  // MN-major A: box of EPR k-rows x 128 B
      const uint32_t b_box_bytes = EPR * 128;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int bt = tile / tiles_per_batch;
        const int rem = tile - bt * tiles_per_batch;
        const int n_blk = rem / p.num_m;
        const int m_blk = rem - n_blk * p.num_m;
        const int m0 = m_blk * kBM, n0 = n_blk * p.bn;
        const int ab0 = bt % p.nb0, ab1 = bt / p.nb0;
        const int btb = bt / p.b_div;
        const int bb0 = btb % p.nb0_b, bb1 = btb / p.nb0_b;
        for (int kb = 0; kb < num_k_total; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t fb = full_bar(stage);
          mbar_arrive_expect_tx(fb, tx_bytes);
          const uint32_t sa = sA0 + (uint32_t)stage * kStageABytes;
          const uint32_t sb = sB0 + (uint32_t)stage * stage_b_bytes;
          if (kb < p.num_k) {
            const int k0 = kb * EPR;
            if (!p.a_mn) {
              tma_load_4d(sa, &tmA, fb, k0, m0, ab0, ab1);
            } else {
              for (int j = 0; j < kBM / EPR; ++j)
                tma_load_4d(sa + j * a_box_bytes, &tmA, fb, m0 + j * EPR, k0, ab0, ab1);
            }
            if (!p.b_mn) {
              tma_load_4d(sb, &tmB, fb, k0, n0, bb0, bb1);
            } else {
              for (int j = 0; j < p.bn / EPR; ++j)
                tma_load_4d(sb + j * b_box_bytes, &tmB, fb, n0 + j * EPR, k0, bb0, bb1);
            }
          } else {
            tma_load_4d(sa, &tmA2, fb, 0, m0, 0, 0);
            tma_load_4d(sb, &tmB2, fb, 0, n0, 0, 0);
          }
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)acc * kAccStride;
        for (int kb = 0; kb < num_k_total; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = sA0 + (uint32_t)stage * kStageABytes;
          const uint32_t sb = sB0 + (uint32_t)stage * stage_b_bytes;
          const bool lora = kb >= p.num_k;
          const int ksteps = lora ? (p.K2 + UK - 1) / UK : EPR / UK;
          const bool amn = p.a_mn && !lora, bmn = p.b_mn && !lora;
          const uint32_t idesc = lora ? p.idesc_lora : p.idesc_main;
#pragma unroll 1
          for (int k = 0; k < ksteps; ++k) {
            // K-major: rows of 128 B, 8-row groups 1024 B apart, k-step = +32 B inside the swizzle row.
            // MN-major: k rows of 128 B (8-row groups 1024 B apart), 128-B column chunks one TMA box
            // (EPR rows x 128 B) apart, k-step = UK rows further down.
            const uint64_t adesc = amn ? make_smem_desc(sa + k * (UK * 128), EPR * 128, 1024, kSwizzle128B)
                                       : make_smem_desc(sa + k * 32, 16, 1024, kSwizzle128B);
            const uint64_t bdesc = bmn ? make_smem_desc(sb + k * (UK * 128), EPR * 128, 1024, kSwizzle128B)
                                       : make_smem_desc(sb + k * 32, 16, 1024, kSwizzle128B);
            umma_issue<ES>(d_tmem, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(empty_bar(stage));
          if (kb == num_k_total - 1) umma_commit(tfull_bar(acc));
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  } else {
    // ================= epilogue (4 warps, one TMEM lane quarter each) =================
    const int q = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int bt = tile / tiles_per_batch;
      const int rem = tile - bt * tiles_per_batch;
      const int n_blk = rem / p.num_m;
      const int m_blk = rem - n_blk * p.num_m;
      const int m0 = m_blk * kBM, n0 = n_blk * p.bn;
      const int ab0 = bt % p.nb0, ab1 = bt / p.nb0;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < p.M;
      const long long doff = (long long)ab0 * p.d_bs0 + (long long)ab1 * p.d_bs1 + (long long)row * p.d_rs;
      const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)acc * kAccStride;
      const int nchunks = p.bn / 32;
      for (int c = 0; c < nchunks; ++c) {
        const int col0 = n0 + c * 32;
        if (col0 >= p.N) break;
        uint32_t r[32];
        tmem_ld_32x32(t0 + c * 32, r);
        tmem_ld_wait();
        if (p.out_dtype == B200_DT_BF16)
          epilogue_store<__nv_bfloat16>(p, reinterpret_cast<__nv_bfloat16*>(p.D) + doff, col0, r, row_ok);
        else if (p.out_dtype == B200_DT_FP32)
          epilogue_store<float>(p, reinterpret_cast<float*>(p.D) + doff, col0, r, row_ok);
        else
          epilogue_store<__half>(p, reinterpret_cast<__half*>(p.D) + doff, col0, r, row_ok);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(acc));
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1u;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int g_num_sms = 0;
static int g_max_smem = 0;

static int device_props() {
  if (g_num_sms == 0) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&g_max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  }
  return 0;
}

// Tile-width choice: fewest SM-waves times the per-k-step cost of a 128 x bn tile
// (tensor pipe: bn/2 cycles; shared-memory operand reads: 32 + bn/4 cycles).
static int pick_bn(int M, int N, int batch, int step, int num_sms) {
  if (N <= 32 && step <= 32) return 32;
  const int num_m = ceil_div(M, kBM);
  int best_bn = 256;
  double best = 1e30;
  for (int bn = 256; bn >= 64; bn -= step) {
    if (bn % step) continue;
    const long long tiles = (long long)batch * num_m * ceil_div(N, bn);
    const long long waves = (tiles + num_sms - 1) / num_sms;
    const double per = (bn / 2.0 > 32 + bn / 4.0) ? bn / 2.0 : 32 + bn / 4.0;
    const double cost = (double)waves * (per + 6.0);  // + fixed per-k-block issue/barrier overhead
    if (cost < best - 1e-9) { best = cost; best_bn = bn; }
  }
  return best_bn;
}

}  // namespace b200

using namespace b200;

B200_API int b200_gemm(const b200_gemm_args* a, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (a->M <= 0 || a->N <= 0 || a->K <= 0 || a->batch <= 0) return B200_ERR_BAD_SHAPE;
  const int ES = a->in_dtype == B200_DT_FP32 ? 4 : 2;
  const int EPR = 128 / ES;
  const int nb0 = a->nb0 > 0 ? a->nb0 : 1;
  const int b_div = a->b_div > 0 ? a->b_div : 1;
  if (nb0 % b_div != 0 && a->batch > 1 && nb0 > 1) return B200_ERR_BAD_SHAPE;
  const bool lora = a->K2 > 0 && a->A2 && a->B2;
  if (lora && a->K2 > EPR) return B200_ERR_UNSUPPORTED;
  // 32-bit MN-major operands need the 128B/32B-atom swizzle mode; callers transpose instead.
  if (ES == 4 && (a->a_mn || a->b_mn)) return B200_ERR_UNSUPPORTED;
  auto al16 = [&](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  auto st16 = [&](long long s) { return (s * ES) % 16 == 0; };
  if (!al16(a->A) || !al16(a->B)) return B200_ERR_BAD_ALIGN;
  if (!st16(a->a_rs) || !st16(a->b_rs) || !st16(a->a_bs0) || !st16(a->a_bs1) || !st16(a->b_bs0) || !st16(a->b_bs1))
    return B200_ERR_BAD_ALIGN;
  if (lora && (!al16(a->A2) || !al16(a->B2) || !st16(a->a2_rs) || !st16(a->b2_rs))) return B200_ERR_BAD_ALIGN;
  int rc = device_props();
  if (rc) return rc;

  GemmKernelParams p;
  p.D = a->D;
  p.bias = a->bias;
  p.M = a->M; p.N = a->N; p.K = a->K; p.K2 = lora ? a->K2 : 0;
  p.batch = a->batch; p.nb0 = nb0; p.b_div = b_div;
  p.nb0_b = nb0 / b_div > 0 ? nb0 / b_div : 1;
  p.d_rs = a->d_rs; p.d_bs0 = a->d_bs0; p.d_bs1 = a->d_bs1;
  p.alpha = a->alpha; p.beta = a->beta;
  p.out_dtype = a->out_dtype; p.bias_dtype = a->bias_dtype;
  p.a_mn = a->a_mn ? 1 : 0; p.b_mn = a->b_mn ? 1 : 0;
  const int step = p.b_mn ? EPR : 32;  // MN-major B is loaded in 128-byte column chunks
  p.bn = pick_bn(a->M, a->N, a->batch, step, g_num_sms);
  p.num_m = ceil_div(a->M, kBM);
  p.num_n = ceil_div(a->N, p.bn);
  p.num_k = ceil_div(a->K, EPR);
  const int stage_bytes = kStageABytes + p.bn * 128;
  int stages = (g_max_smem - 2048) / stage_bytes;
  if (stages > 8) stages = 8;
  if (stages < 2) return B200_ERR_UNSUPPORTED;
  p.stages = stages;
  const bool is_bf16 = a->in_dtype == B200_DT_BF16;
  if (ES == 2) {
    p.idesc_main = make_idesc_f16(kBM, p.bn, is_bf16, p.a_mn, p.b_mn);
    p.idesc_lora = make_idesc_f16(kBM, p.bn, is_bf16, false, false);
  } else {
    // kind::tf32: a_format = b_format = 2
    p.idesc_main = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16) |
                   ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)(kBM >> 4) << 24);
    p.idesc_lora = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)(kBM >> 4) << 24);
  }
  const int out_es = a->out_dtype == B200_DT_FP32 ? 4 : 2;
  const int per_vec = 16 / out_es;
  p.vec_ok = ((reinterpret_cast<uintptr_t>(a->D) & 15) == 0) && (a->d_rs % per_vec == 0) &&
             (a->d_bs0 % per_vec == 0) && (a->d_bs1 % per_vec == 0);

  // ---- tensor maps (dims innermost first; 4-D: contiguous, rows, batch0, batch1) ----
  CUtensorMap tmA, tmB, tmA2, tmB2;
  const int nb1 = ceil_div(a->batch, nb0);
  const int batch_b = ceil_div(a->batch, b_div);
  const int nb1_b = ceil_div(batch_b, p.nb0_b);
  {
    uint64_t dims[4], str[3];
    uint32_t box[4];
    if (!p.a_mn) { dims[0] = a->K; dims[1] = a->M; box[0] = EPR; box[1] = kBM; }
    else         { dims[0] = a->M; dims[1] = a->K; box[0] = EPR; box[1] = EPR; }
    dims[2] = nb0; dims[3] = nb1; box[2] = 1; box[3] = 1;
    str[0] = (uint64_t)a->a_rs * ES; str[1] = (uint64_t)a->a_bs0 * ES; str[2] = (uint64_t)a->a_bs1 * ES;
    rc = make_tensor_map(&tmA, a->A, ES, 4, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  {
    uint64_t dims[4], str[3];
    uint32_t box[4];
    if (!p.b_mn) { dims[0] = a->K; dims[1] = a->N; box[0] = EPR; box[1] = p.bn; }
    else         { dims[0] = a->N; dims[1] = a->K; box[0] = EPR; box[1] = EPR; }
    dims[2] = p.nb0_b; dims[3] = nb1_b; box[2] = 1; box[3] = 1;
    str[0] = (uint64_t)a->b_rs * ES; str[1] = (uint64_t)a->b_bs0 * ES; str[2] = (uint64_t)a->b_bs1 * ES;
    rc = make_tensor_map(&tmB, a->B, ES, 4, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  if (lora) {
    uint64_t dims[4] = {(uint64_t)a->K2, (uint64_t)a->M, 1, 1}, str[3] = {(uint64_t)a->a2_rs * ES, 0, 0};
    uint32_t box[4] = {(uint32_t)EPR, (uint32_t)kBM, 1, 1};
    rc = make_tensor_map(&tmA2, a->A2, ES, 4, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    uint64_t dimsb[4] = {(uint64_t)a->K2, (uint64_t)a->N, 1, 1}, strb[3] = {(uint64_t)a->b2_rs * ES, 0, 0};
    uint32_t boxb[4] = {(uint32_t)EPR, (uint32_t)p.bn, 1, 1};
    rc = make_tensor_map(&tmB2, a->B2, ES, 4, dimsb, strb, boxb, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  } else {
    tmA2 = tmA;
    tmB2 = tmB;
  }

  const long long total_tiles = (long long)p.num_m * p.num_n * p.batch;
  const int grid = (int)(total_tiles < g_num_sms ? total_tiles : g_num_sms);
  const int smem = stages * stage_bytes + 8 * (2 * stages + 5) + 1024;
  if (ES == 2) {
    static int configured = 0;
    if (configured < smem) {
      B200_CHECK(cudaFuncSetAttribute(gemm_tcgen05_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_max_smem));
      configured = g_max_smem;
    }
    gemm_tcgen05_kernel<2><<<grid, kGemmThreads, smem, stream>>>(tmA, tmB, tmA2, tmB2, p);
  } else {
    static int configured = 0;
    if (configured < smem) {
      B200_CHECK(cudaFuncSetAttribute(gemm_tcgen05_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_max_smem));
      configured = g_max_smem;
    }
    gemm_tcgen05_kernel<4><<<grid, kGemmThreads, smem, stream>>>(tmA, tmB, tmA2, tmB2, p);
  }
  B200_LAUNCH_CHECK();
  return 0;
}
