// K1 / K3 — persistent, warp-specialised tcgen05 GEMM for sm_100a.
//
//   D[b] = alpha * A[b] x B[b / b_div]^T  (+ A2 x B2^T)  (+ bias)  + beta * D[b]
//
// Work unit = one CTA (128 x BN tile) or one CTA *pair* (cta_group::2, 256 x BN tile: each CTA
// stages its own 128 rows of A and half of B, the leader issues one 256-row UMMA for both SMs).
// The pair halves the B traffic per flop, which matters because at 128 x 256 tiles the kernel is
// bound by L2->SM bandwidth (~12.5 TB/s), not by the tensor pipe: 85 flop/B caps a single-CTA
// kernel near 1.0 PFLOP/s, 128 flop/B lifts the cap to ~1.6 (measured; see profiles/).
// Per unit: warp 0 feeds a ring of shared-memory stages with TMA (128-byte swizzle), warp 1 issues
// tcgen05.mma into one of two TMEM accumulators, warps 2-5 drain the other accumulator with
// tcgen05.ld and write D, so the epilogue of tile i overlaps the main loop of tile i+1. BN
// (16..256), the CTA-group size and the stage count are run-time choices: the host picks the
// combination with the lowest modelled time (SM-waves x max(tensor-pipe, L2) per k-block).
//
// Operands are read in place in either storage order ("K-major" or "MN-major" UMMA descriptors),
// so forward (X W^T), dgrad (dY W) and wgrad (dY^T X) all run through this one kernel with no
// transpose copies, for bf16/fp16 (kind::f16) and fp32 (kind::tf32, K-major only) inputs. The
// optional A2/B2 operands append a rank-K2 update as one extra k-block: that is the LoRA term of
// peft's lora.Linear (Fine-Tuning/qwen3-8b-lora-dist.py:118-125), computed on the tensor cores
// inside the same accumulator instead of as three extra cuBLAS calls.
//
// Replaces: cuBLAS/cuBLASLt behind every nn.Linear in the reference, e.g.
// LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:86-165.
#include "common.cuh"
#include "tma_host.h"
#include "../include/b200_ops.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

namespace b200 {

constexpr int kBM = 128;             // rows per CTA = TMEM lanes
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
  const float* beta_rows;   // optional per-row beta (overrides `beta`)
  int out_dtype, bias_dtype;
  uint32_t idesc_main, idesc_lora;
  int num_m, num_n, num_k;  // num_m counts unit tiles (128 * CTAS rows); num_k = k-blocks of the main operands
  int bn, stages, a_mn, b_mn;
  int vec_ok;
  // stream-K: every unit gets an equal contiguous range of (tile, k-block) work items; a tile that is
  // split between units is summed through `sk_ws` (one fp32 128 x 256 slot per CTA) by the unit
  // that owns the tile's first k-block, after the others published their partial (sk_flags = epoch).
  int stream_k, sk_group;
  float* sk_ws;
  unsigned int* sk_flags;
  unsigned int sk_epoch;      // host-provided epoch (used when sk_state is null)
  unsigned int* sk_state;     // device epoch: [0] = epoch of the last stream-K launch, [1] = ticket counter (graph-replayable)
  // K16 (grouped GEMM): b_map[batch] = which B matrix (and bias row) batch entry `batch` multiplies with, -1 = skip the
  // entry. Read on the device, so the grouping (tokens sorted by expert) never visits the host.
  const int* b_map;
  long long bias_bs;          // elements between the bias rows of consecutive B matrices (0 = one shared bias)
  // K-C5: all-reduce of D over the tensor-parallel ranks inside this launch (row-parallel projection). D is every rank's
  // copy of one symmetric bf16 buffer and ar_mc its multicast address. A finished tile is announced to its owner rank
  // (tile % world) with an epoch flag per (tile, CTA of the pair, source rank); once a CTA has no GEMM tile left, its
  // epilogue warps take the owned tiles whose partials have arrived from every rank and sum them IN PLACE inside NVSwitch
  // (multimem.ld_reduce of the W copies, multimem.st of the sum back to all of them). The last CTA of a rank then tells
  // the peers that this rank's share is reduced and waits for theirs, so the launch ends with D complete everywhere.
  int ar_world, ar_rank, ar_channel;
  unsigned int ar_epoch;
  char* ar_mc;
  unsigned int* ar_tile_flags[8];   // per rank: uint32 [tile][2][8]
  unsigned int* ar_sync[8];         // per rank: uint32 [channel][8] epoch flags
  unsigned int* ar_ticket;          // local ticket counter (returns to 0)
  char* ar_d[2];                    // two ranks: both copies of D by peer pointer (ar_mc null)
};

// Fused 4-bit B operand (W4 = true instantiation): the weight matrix W [rows, kw] is stored packed (two codes per
// byte, kw contiguous, even element in the high nibble) with one fp32 scale (and optional zero*scale) per `group`
// elements, possibly split by rows over up to 8 buffers (ZeRO-3 shards in peer memory, read through NVLink by TMA).
// w = lut[code] * scale - zs. b_mn = 0: B = W (forward, GEMM N = rows, K = kw); b_mn = 1: B = W^T read in place
// (dgrad, GEMM N = kw, K = rows).
// Timing experiments on the fused 4-bit path (skip table look-ups, fences, barrier legs, ...) are compiled out of the
// shipped kernel: W4DBG(mask) is the constant 0 unless the library is built with -DB200_W4_EXPERIMENT.
#ifdef B200_W4_EXPERIMENT
#define W4DBG(mask) (w4->debug & (mask))
#else
#define W4DBG(mask) 0
#endif

struct W4Params {
  CUtensorMap tmW[8];
  const float* scale[8];
  const float* zs[8];
  const float* lut;  // 16 entries, nullptr = NF4
  int nshards, rows_per_shard, group, kw;
  int raw_stages, is_bf16;
  int lut_bytes;
  int debug;  // B200_W4_DEBUG bit mask; only read when the library is built with -DB200_W4_EXPERIMENT (timing experiments)
};

constexpr int kW4Threads = 480;      // + 8 dequant warps + 1 packed-tile TMA warp
constexpr int kW4DqWarps = 8;
constexpr int kRawStageBytes = 4096; // 128 rows x 32 B (K-major) or 64 rows x 64 B (MN-major) of packed codes
constexpr int kLutBytes = 256 * 32 * 4;

__constant__ float kNF4Table[16] = {-1.0f, -0.6961928009986877f, -0.5250730514526367f, -0.39491748809814453f,
                                    -0.28444138169288635f, -0.18477343022823334f, -0.09105003625154495f, 0.0f,
                                    0.07958029955625534f, 0.16093020141124725f, 0.24611230194568634f,
                                    0.33791524171829224f, 0.44070982933044434f, 0.5626170039176941f,
                                    0.7229568362236023f, 1.0f};

// One accumulator's worth of work: k-blocks [kb0, kb1) of `tile`.
struct Segment {
  int tile, kb0, kb1;
};

// Iterates the segments of one unit, identically in the producer, MMA and epilogue roles.
struct SegmentIter {
  long long item, end;  // stream-K: current / last work item of this unit
  int tile, stride, total_tiles, nkt, grp_size, member;
  bool stream_k;
  __device__ SegmentIter(const GemmKernelParams& p, int unit, int nunits, int total_tiles_, int nkt_) {
    stream_k = p.stream_k != 0;
    total_tiles = total_tiles_;
    nkt = nkt_;
    tile = unit;
    stride = nunits;
    // stream-K work space = (panel, k-block), a panel being the sk_group M-tiles that share one B panel. The units
    // form groups of sk_group members: a group walks one contiguous (panel, k) range and member j computes M-tile j
    // of it, so the members read the same B tiles at the same time (one DRAM fetch, sk_group - 1 L2 hits). With
    // sk_group = 1 this is plain per-tile stream-K.
    grp_size = p.stream_k ? p.sk_group : 1;
    member = unit % grp_size;
    const int ngroups = nunits / grp_size, grp = unit / grp_size;
    const long long total = (long long)(total_tiles_ / grp_size) * nkt_;
    if (grp < ngroups) {
      item = total * grp / ngroups;
      end = total * (grp + 1) / ngroups;
    } else {
      item = end = 0;   // spare units beyond the last whole group
    }
  }
  __device__ bool next(Segment& s) {
    if (!stream_k) {
      if (tile >= total_tiles) return false;
      s.tile = tile; s.kb0 = 0; s.kb1 = nkt;
      tile += stride;
      return true;
    }
    if (item >= end) return false;
    const int panel = (int)(item / nkt);
    s.tile = panel * grp_size + member;
    s.kb0 = (int)(item - (long long)panel * nkt);
    const long long left = end - item;
    s.kb1 = (long long)(nkt - s.kb0) <= left ? nkt : s.kb0 + (int)left;
    item += s.kb1 - s.kb0;
    return true;
  }
};

template <int ES, int CTAS>
__device__ __forceinline__ void umma_issue(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  if constexpr (CTAS == 1) {
    if constexpr (ES == 2) umma_f16(d, a, b, idesc, acc);
    else umma_tf32(d, a, b, idesc, acc);
  } else {
    if constexpr (ES == 2) umma_f16_2sm(d, a, b, idesc, acc);
    else umma_tf32_2sm(d, a, b, idesc, acc);
  }
}

template <typename OutT>
__device__ __forceinline__ void epilogue_store(const GemmKernelParams& p, OutT* __restrict__ drow, int col0,
                                               int col_limit, const uint32_t (&r)[32], bool row_ok, float beta,
                                               long long bias_off) {
  // drow points at D[row, 0]; this thread owns columns col0 .. col0+31 (< col_limit) of its row.
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;
  if (p.bias != nullptr) {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const long long c = bias_off + col0 + j;
      if (col0 + j < col_limit) {
        float b = p.bias_dtype == B200_DT_FP32 ? reinterpret_cast<const float*>(p.bias)[c]
                  : p.bias_dtype == B200_DT_BF16
                      ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.bias)[c])
                      : __half2float(reinterpret_cast<const __half*>(p.bias)[c]);
        v[j] += b;
      }
    }
  }
  if (!row_ok) return;
  const bool full = (col0 + 32 <= col_limit) && p.vec_ok;
  if (full) {
    constexpr int kPerVec = 16 / sizeof(OutT);
    constexpr int kVecs = 32 / kPerVec;
    uint4* dst = reinterpret_cast<uint4*>(drow + col0);
    if (beta != 0.f) {
#pragma unroll
      for (int i = 0; i < kVecs; ++i) {
        uint4 old = dst[i];
        const OutT* o = reinterpret_cast<const OutT*>(&old);
#pragma unroll
        for (int j = 0; j < kPerVec; ++j) v[i * kPerVec + j] += beta * to_f32<OutT>(o[j]);
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
      if (c < col_limit) {
        float x = v[j];
        if (beta != 0.f) x += beta * to_f32<OutT>(drow[c]);
        drow[c] = from_f32<OutT>(x);
      }
    }
  }
}

// ---- K-C5: the all-reduce half of the fused row-parallel GEMM (called by the four epilogue warps, 128 threads) ----
// A reduction item is a slab of kArSlabRows rows of a tile this rank owns (tile % world == rank); item i of the rank goes
// to CTA i % gridDim.x. An item can be summed once every rank has announced the tile (both CTAs of a pair). With
// `block` false the call returns at the first item that is not ready yet (it is made between two GEMM tiles, while the
// tensor pipe works on the next one). The sum of the W copies of a 16-byte vector is taken inside NVSwitch
// (multimem.ld_reduce, fp32 accumulation) and written back to every copy (multimem.st) when ar_mc is set and W > 2; with
// two ranks plain peer loads / stores move the same bytes faster.
constexpr int kArSlabRows = 32;

template <int CTAS>
__device__ __noinline__ int ar_reduce(const GemmKernelParams& p, int next, int n_items, int budget, bool block,
                                         int total_tiles, uint32_t poll_slot) {
  const int et = (int)threadIdx.x - 64;     // 0..127
  const int W = p.ar_world;
  constexpr int kSlabs = (kBM * CTAS) / kArSlabRows;
  while (next < n_items && budget > 0) {
    const int tile = (next / kSlabs) * W + p.ar_rank, slab = next % kSlabs;
    // ---- have all ranks (and both CTAs of their pairs) stored this tile? ----
    if (et < 32) {
      bool ok = true;
      if (et < CTAS * W) {
        const unsigned int* f = p.ar_tile_flags[p.ar_rank] + ((long long)(tile * 2 + et / W) * 8 + et % W);
        unsigned int seen;
        const long long tw = clock64();
        do {
          asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(f) : "memory");
          ok = (int)(seen - p.ar_epoch) >= 0;
          if (block && !ok && clock64() - tw > 20000000000ll) {
            printf("b200 gemm+allreduce: rank %d timed out waiting for tile %d (CTA %d of the pair) of rank %d: flag %u, epoch %u\n",
                   p.ar_rank, tile, et / W, et % W, seen, p.ar_epoch);
            __trap();
          }
        } while (block && !ok);
      }
      ok = __all_sync(0xffffffffu, ok);
      if (et == 0) asm volatile("st.shared.u32 [%0], %1;" ::"r"(poll_slot), "r"(ok ? 1u : 0u) : "memory");
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    uint32_t ready;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(ready) : "r"(poll_slot) : "memory");
    asm volatile("bar.sync 1, 128;" ::: "memory");      // everyone has read the verdict before the slot is rewritten
    if (!ready) break;
    const int n_blk = tile / p.num_m, m_blk = tile - n_blk * p.num_m;
    const int m0 = m_blk * (kBM * CTAS) + slab * kArSlabRows, n0 = n_blk * p.bn;
    const int rows = min(kArSlabRows, p.M - m0);
    const int vpr = (min(p.N, n0 + p.bn) - n0) / 8;            // 16-byte vectors per row (bf16)
    const int nvec = rows > 0 ? rows * vpr : 0;
    const long long off0 = ((long long)m0 * p.d_rs + n0) * 2;
    constexpr int U = 8;
    if (p.ar_mc != nullptr) {
      char* base = p.ar_mc + off0;
      for (int v0 = et; v0 < nvec; v0 += 128 * U) {
        uint4 x[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int v = v0 + u * 128;
          if (v < nvec) {
            const char* a = base + (long long)(v / vpr) * p.d_rs * 2 + (v % vpr) * 16;
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                         : "=r"(x[u].x), "=r"(x[u].y), "=r"(x[u].z), "=r"(x[u].w) : "l"(a) : "memory");
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int v = v0 + u * 128;
          if (v < nvec) {
            char* a = base + (long long)(v / vpr) * p.d_rs * 2 + (v % vpr) * 16;
            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a), "r"(x[u].x), "r"(x[u].y),
                         "r"(x[u].z), "r"(x[u].w) : "memory");
          }
        }
      }
    } else {
      for (int v0 = et; v0 < nvec; v0 += 128 * U) {
        uint4 x[U][2];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int v = v0 + u * 128;
          if (v < nvec) {
            const long long o = off0 + (long long)(v / vpr) * p.d_rs * 2 + (v % vpr) * 16;
            x[u][0] = *reinterpret_cast<const uint4*>(p.ar_d[0] + o);
            x[u][1] = *reinterpret_cast<const uint4*>(p.ar_d[1] + o);
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int v = v0 + u * 128;
          if (v < nvec) {
            const long long o = off0 + (long long)(v / vpr) * p.d_rs * 2 + (v % vpr) * 16;
            uint4 y;
            const uint32_t* a = reinterpret_cast<const uint32_t*>(&x[u][0]);
            const uint32_t* b = reinterpret_cast<const uint32_t*>(&x[u][1]);
            uint32_t* c = reinterpret_cast<uint32_t*>(&y);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 fa = unpack_bf16x2(a[e]), fb = unpack_bf16x2(b[e]);
              c[e] = pack_bf16x2(fa.x + fb.x, fa.y + fb.y);
            }
            *reinterpret_cast<uint4*>(p.ar_d[0] + o) = y;
            *reinterpret_cast<uint4*>(p.ar_d[1] + o) = y;
          }
        }
      }
    }
    next += gridDim.x;
    --budget;
  }
  return next;
}

template <int ES, int CTAS, bool W4>
__device__ __forceinline__ void gemm_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmA2,
                                          const CUtensorMap& tmB2, const GemmKernelParams& p, const W4Params* w4) {
  constexpr int EPR = 128 / ES;  // elements per 128-byte swizzle row = k-block depth
  constexpr int UK = 32 / ES;    // elements per UMMA k-step

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int bn_cta = p.bn / CTAS;  // B rows (output columns) staged by this CTA
  const uint32_t stage_b_bytes = (uint32_t)bn_cta * 128u;
  const uint32_t sA0 = smem_base;
  const uint32_t sB0 = smem_base + (uint32_t)p.stages * kStageABytes;
  const int raw_stages = W4 ? w4->raw_stages : 0;
  const uint32_t raw0 = sB0 + (uint32_t)p.stages * stage_b_bytes;          // packed-code ring (W4 only)
  const uint32_t lut0 = raw0 + (uint32_t)raw_stages * kRawStageBytes;      // byte -> 2 x 16-bit table, one copy per lane
  const uint32_t bar0 = lut0 + (W4 ? (uint32_t)w4->lut_bytes : 0u);
  // barrier block: full[stages] | empty[stages] | tmem_full[2] | tmem_empty[2] | tmem_ptr | raw_full[] | raw_empty[]
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (p.stages + s); };
  auto tfull_bar = [&](int a) { return bar0 + 8u * (2 * p.stages + a); };
  auto tempty_bar = [&](int a) { return bar0 + 8u * (2 * p.stages + 2 + a); };
  const uint32_t tmem_slot = bar0 + 8u * (2 * p.stages + 4);
  auto rfull_bar = [&](int s) { return bar0 + 8u * (2 * p.stages + 5 + s); };
  auto rempty_bar = [&](int s) { return bar0 + 8u * (2 * p.stages + 5 + raw_stages + s); };

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = CTAS == 2 ? cluster_ctarank() : 0u;
  const bool leader = rank == 0;
  const int unit = blockIdx.x / CTAS, nunits = gridDim.x / CTAS;

  const uint32_t sk_epoch_slot = tmem_slot + 4u;   // second half of the 8-byte TMEM-pointer slot (no static shared memory: the
                                                   // dynamic allocation already takes the whole 227 KB)
  if (warp == 0 && lane == 0) {
    if (p.stream_k) {
      unsigned int e = p.sk_epoch;
      if (p.sk_state) {
        // every CTA of the launch reads the old epoch, then takes a ticket; the last ticket publishes the new epoch for the
        // next launch on this stream. Nothing here comes from the host, so a captured launch can be replayed.
        e = *((volatile unsigned int*)p.sk_state) + 1u;
        __threadfence();
        if (atomicAdd(p.sk_state + 1, 1u) == gridDim.x - 1) {
          p.sk_state[1] = 0u;
          __threadfence();
          *((volatile unsigned int*)p.sk_state) = e;
        }
      }
      asm volatile("st.shared.u32 [%0], %1;" ::"r"(sk_epoch_slot), "r"(e) : "memory");
    }
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.K2 > 0) {
      tma_prefetch_desc(&tmA2);
      tma_prefetch_desc(&tmB2);
    }
    for (int s = 0; s < p.stages; ++s) {
      // one arrival per CTA's producer (+ the TMA transaction bytes) (+ W4: the dequant warps of every CTA)
      mbar_init(full_bar(s), (W4 && !W4DBG(128)) ? CTAS * 2 : CTAS);   // W4: + one arrival per CTA from its dequant warps
      mbar_init(empty_bar(s), 1);     // tcgen05.commit (multicast to both CTAs of a pair)
    }
    for (int s = 0; s < raw_stages; ++s) {
      mbar_init(rfull_bar(s), 1);
      mbar_init(rempty_bar(s), W4DBG(256) ? 1 : kW4DqWarps);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 4 * CTAS);  // epilogue warps of every CTA in the unit
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    if constexpr (CTAS == 1) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    else { tmem_alloc_2sm(tmem_slot, 512); tmem_relinquish_2sm(); }
  }
  if (W4 && w4->lut_bytes) {
    // entry (byte, lane) = {16-bit lut[byte >> 4] (even element, low half), 16-bit lut[byte & 15]}: lane l only ever
    // reads column l, so the 32 lanes of a warp hit 32 different banks whatever their bytes are
    for (int i = threadIdx.x; i < 256 * 32; i += blockDim.x) {
      const int byte = i >> 5;
      const float lo = w4->lut ? w4->lut[byte >> 4] : kNF4Table[byte >> 4];
      const float hi = w4->lut ? w4->lut[byte & 15] : kNF4Table[byte & 15];
      uint32_t v;
      if (w4->is_bf16) v = pack_bf16x2(lo, hi);
      else { const __half2 h = __floats2half2_rn(lo, hi); v = *reinterpret_cast<const uint32_t*>(&h); }
      asm volatile("st.shared.u32 [%0], %1;" ::"r"(lut0 + 4u * i), "r"(v) : "memory");
    }
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (CTAS == 2) cluster_sync_all();  // the peer's barriers exist before anyone signals them
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const int tiles_per_batch = p.num_m * p.num_n;
  const int total_tiles = tiles_per_batch * p.batch;
  const int num_k_total = p.num_k + (p.K2 > 0 ? 1 : 0);

  if (warp == 0) {
    // ================= TMA producer (every CTA loads its own rows of A and its share of B) =========
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t tx_full = (uint32_t)CTAS * (kStageABytes + stage_b_bytes);
      const uint32_t tx_a_only = (uint32_t)CTAS * kStageABytes;
      const uint32_t box_bytes = EPR * 128;  // MN-major operands: one box = EPR k-rows x 128 B
      SegmentIter it(p, unit, nunits, total_tiles, num_k_total);
      Segment sg;
      while (it.next(sg)) {
        const int tile = sg.tile;
        const int bt = tile / tiles_per_batch;
        const int rem = tile - bt * tiles_per_batch;
        const int n_blk = rem / p.num_m;
        const int m_blk = rem - n_blk * p.num_m;
        const int m0 = m_blk * (kBM * CTAS) + (int)rank * kBM;
        const int n0 = n_blk * p.bn + (int)rank * bn_cta;
        const int ab0 = bt % p.nb0, ab1 = bt / p.nb0;
        const int btb = p.b_map ? p.b_map[bt] : bt / p.b_div;
        if (btb < 0) continue;                       // grouped GEMM: unused row tile
        const int bb0 = btb % p.nb0_b, bb1 = btb / p.nb0_b;
        for (int kb = sg.kb0; kb < sg.kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t fb_local = full_bar(stage);
          const uint32_t fb = leader ? fb_local : mapa_shared(fb_local, 0);  // the leader's barrier
          const bool dbg_b = W4 && W4DBG(64);   // timing experiment: also stream a (dummy) B tile through TMA
          if (leader) mbar_arrive_expect_tx(fb_local, (W4 && kb < p.num_k && !dbg_b) ? tx_a_only : tx_full);
          else mbar_arrive_cluster(fb);
          const uint32_t sa = sA0 + (uint32_t)stage * kStageABytes;
          const uint32_t sb = sB0 + (uint32_t)stage * stage_b_bytes;
          auto load = [&](uint32_t dst, const CUtensorMap* tm, int c0, int c1, int c2, int c3) {
            if constexpr (CTAS == 1) tma_load_4d(dst, tm, fb, c0, c1, c2, c3);
            else tma_load_4d_2sm(dst, tm, fb, c0, c1, c2, c3);
          };
          if (kb < p.num_k) {
            const int k0 = kb * EPR;
            if (!p.a_mn) {
              load(sa, &tmA, k0, m0, ab0, ab1);
            } else {
              for (int j = 0; j < kBM / EPR; ++j) load(sa + j * box_bytes, &tmA, m0 + j * EPR, k0, ab0, ab1);
            }
            if constexpr (W4) { if (dbg_b) load(sb, &tmB, k0, (n0 % 512), bb0, bb1); }
            if constexpr (!W4) {
              if (!p.b_mn) {
                load(sb, &tmB, k0, n0, bb0, bb1);
              } else {
                for (int j = 0; j < bn_cta / EPR; ++j) load(sb + j * box_bytes, &tmB, n0 + j * EPR, k0, bb0, bb1);
              }
            }
          } else {
            load(sa, &tmA2, 0, m0, 0, 0);
            load(sb, &tmB2, 0, n0, 0, 0);
          }
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (one thread of the leader CTA) =================
    if (lane == 0 && leader) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      // Shared-memory descriptors, split so that the per-UMMA work is ONE 32-bit add per operand.
      // (Building each 64-bit descriptor from scratch costs ~45 dependent uniform-datapath ops per
      // k-block, ~540 cycles, which made this thread — not the tensor pipe — the bottleneck.)
      //   hi word: stride-byte-offset 1024 (8-row groups) | version 1 | 128-byte swizzle
      //   lo word: (address >> 4) | (leading-byte-offset >> 4) << 16
      // K-major: rows of 128 B, k-step = +32 B inside the swizzle row, LBO unused (16).
      // MN-major: k rows of 128 B, 128-B column chunks one TMA box (EPR rows x 128 B) apart (LBO),
      //           k-step = UK rows further down.
      constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (kSwizzle128B << 29);
      constexpr uint32_t kLoK = (16u >> 4) << 16, kLoMN = ((uint32_t)(EPR * 128) >> 4) << 16;
      constexpr uint32_t kStepK = 32u >> 4, kStepMN = (uint32_t)(UK * 128) >> 4;
      const uint32_t a_lo_main = (sA0 >> 4) | (p.a_mn ? kLoMN : kLoK);
      const uint32_t b_lo_main = (sB0 >> 4) | (p.b_mn ? kLoMN : kLoK);
      const uint32_t a_step_main = p.a_mn ? kStepMN : kStepK, b_step_main = p.b_mn ? kStepMN : kStepK;
      const uint32_t a_lo_lora = (sA0 >> 4) | kLoK, b_lo_lora = (sB0 >> 4) | kLoK;
      const uint32_t a_stage16 = kStageABytes >> 4, b_stage16 = stage_b_bytes >> 4;
      const int lora_ksteps = (p.K2 + UK - 1) / UK;
      auto desc = [](uint32_t lo) { return ((uint64_t)kDescHi << 32) | (uint64_t)lo; };
      SegmentIter it(p, unit, nunits, total_tiles, num_k_total);
      Segment sg;
      while (it.next(sg)) {
        if (p.b_map && p.b_map[sg.tile / tiles_per_batch] < 0) continue;   // grouped GEMM: unused row tile
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)acc * kAccStride;
        for (int kb = sg.kb0; kb < sg.kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          if (kb < p.num_k) {
            const uint32_t a_lo = a_lo_main + (uint32_t)stage * a_stage16;
            const uint32_t b_lo = b_lo_main + (uint32_t)stage * b_stage16;
            umma_issue<ES, CTAS>(d_tmem, desc(a_lo), desc(b_lo), p.idesc_main, kb != sg.kb0 ? 1u : 0u);
#pragma unroll
            for (int k = 1; k < EPR / UK; ++k)
              umma_issue<ES, CTAS>(d_tmem, desc(a_lo + k * a_step_main), desc(b_lo + k * b_step_main), p.idesc_main, 1u);
          } else {
            const uint32_t a_lo = a_lo_lora + (uint32_t)stage * a_stage16;
            const uint32_t b_lo = b_lo_lora + (uint32_t)stage * b_stage16;
            for (int k = 0; k < lora_ksteps; ++k)
              umma_issue<ES, CTAS>(d_tmem, desc(a_lo + k * kStepK), desc(b_lo + k * kStepK), p.idesc_lora,
                                   (kb != sg.kb0 || k != 0) ? 1u : 0u);
          }
          if constexpr (CTAS == 1) {
            umma_commit(empty_bar(stage));
            if (kb == sg.kb1 - 1) umma_commit(tfull_bar(acc));
          } else {
            umma_commit_2sm(empty_bar(stage), 3);
            if (kb == sg.kb1 - 1) umma_commit_2sm(tfull_bar(acc), 3);
          }
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  } else if (W4 && warp >= 6 && (W4DBG(128) || (W4DBG(256) && warp > 6 && warp < 6 + kW4DqWarps))) {
    // timing experiment: the extra warps stay out of the way
  } else if (W4 && warp == 6 + kW4DqWarps) {
    // ================= packed-weight TMA producer: runs `raw_stages` k-blocks ahead of the dequant warps ========
    if constexpr (W4) if (lane == 0) {
      int rs = 0;
      uint32_t rphase = 0;
      SegmentIter it(p, unit, nunits, total_tiles, num_k_total);
      Segment sg;
      while (it.next(sg)) {
        const int rem = sg.tile % tiles_per_batch;
        const int n_blk = rem / p.num_m;
        const int n0 = n_blk * p.bn + (int)rank * bn_cta;
        for (int kb = sg.kb0; kb < sg.kb1 && kb < p.num_k; ++kb) {
          mbar_wait(rempty_bar(rs), rphase ^ 1u);
          if (W4DBG(32)) {   // timing experiment: no packed-tile traffic
            mbar_arrive(rfull_bar(rs));
            if (++rs == raw_stages) { rs = 0; rphase ^= 1u; }
            continue;
          }
          mbar_arrive_expect_tx(rfull_bar(rs), kRawStageBytes);
          const uint32_t dst = raw0 + (uint32_t)rs * kRawStageBytes;
          if (!p.b_mn) {   // rows n0.. of W, bytes k0/2..
            const int sh = n0 / w4->rows_per_shard;
            tma_load_2d(dst, &w4->tmW[sh], rfull_bar(rs), kb * (EPR / 2), n0 - sh * w4->rows_per_shard);
          } else {         // rows k0.. of W, bytes n0/2..
            const int k0 = kb * EPR;
            const int sh = k0 / w4->rows_per_shard;
            tma_load_2d(dst, &w4->tmW[sh], rfull_bar(rs), n0 / 2, k0 - sh * w4->rows_per_shard);
          }
          if (++rs == raw_stages) { rs = 0; rphase ^= 1u; }
        }
      }
    }
  } else if (W4 && warp >= 6) {
    // ================= dequant warps: packed codes -> swizzled 16-bit B stage =================
    if constexpr (W4) {
      const int t = threadIdx.x - 192;   // 0..255: two threads per 128-byte row of the B stage
      int stage = 0, rs = 0;
      uint32_t phase = 0, rphase = 0;
      const int gpr = w4->kw / w4->group;     // scale groups per stored row of W
      const uint32_t lut_lane = lut0 + 4u * lane;
      const bool bf = w4->is_bf16 != 0;
      SegmentIter it(p, unit, nunits, total_tiles, num_k_total);
      Segment sg;
      while (it.next(sg)) {
        const int rem = sg.tile % tiles_per_batch;
        const int n_blk = rem / p.num_m;
        const int n0 = n_blk * p.bn + (int)rank * bn_cta;
        // K-major: B row n0 + t/2, k half t&1 of the block. MN-major: k row t/4, column half (t>>1)&1, 32-column part t&1.
        const int r_loc = p.b_mn ? (t >> 2) : (t >> 1);
        const int half = p.b_mn ? ((t >> 1) & 1) : 0;
        const int sub = t & 1;                 // which 16 packed bytes (4 of the row's 8 sixteen-byte chunks)
        const uint32_t raw_off = p.b_mn ? (uint32_t)(r_loc * 64 + half * 32 + sub * 16) : (uint32_t)(r_loc * 32 + sub * 16);
        const uint32_t dst_off = p.b_mn ? (uint32_t)(half * EPR * 128 + r_loc * 128) : (uint32_t)(r_loc * 128);
        const uint32_t key = (uint32_t)(r_loc & 7);
        auto scale_at = [&](int kb, float& sc, float& zs) {
          long long idx;
          int sh;
          if (!p.b_mn) {
            const int n = n0 + r_loc;
            sh = n / w4->rows_per_shard;
            idx = (long long)(n - sh * w4->rows_per_shard) * gpr + (kb * EPR) / w4->group;
          } else {
            const int o = kb * EPR + r_loc;
            sh = o / w4->rows_per_shard;
            idx = (long long)(o - sh * w4->rows_per_shard) * gpr + (n0 + half * 64) / w4->group;
          }
          sc = __ldg(w4->scale[sh] + idx);
          zs = w4->zs[sh] ? __ldg(w4->zs[sh] + idx) : 0.f;
        };
        // scales are fetched 3 k-blocks ahead (L2 latency is longer than one k-block of tensor work)
        float scq[3] = {0.f, 0.f, 0.f}, zsq[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < 3; ++j)
          if (sg.kb0 + j < sg.kb1 && sg.kb0 + j < p.num_k) scale_at(sg.kb0 + j, scq[j], zsq[j]);
        for (int kb = sg.kb0; kb < sg.kb1; ++kb) {
          const uint32_t fb_local = full_bar(stage);
          if (kb < p.num_k) {
            const float sc = scq[0], zs = zsq[0];
            scq[0] = scq[1]; scq[1] = scq[2]; zsq[0] = zsq[1]; zsq[1] = zsq[2];
            if (kb + 3 < sg.kb1 && kb + 3 < p.num_k) scale_at(kb + 3, scq[2], zsq[2]);
            uint32_t sc2, nz2;
            if (w4->is_bf16) { sc2 = pack_bf16x2(sc, sc); nz2 = pack_bf16x2(-zs, -zs); }
            else {
              const __half2 a = __floats2half2_rn(sc, sc), b = __floats2half2_rn(-zs, -zs);
              sc2 = *reinterpret_cast<const uint32_t*>(&a); nz2 = *reinterpret_cast<const uint32_t*>(&b);
            }
            if (W4DBG(8)) {   // timing experiment: no expansion at all, only the barrier traffic
              mbar_wait(rfull_bar(rs), rphase);
              __syncwarp();
              if (lane == 0) mbar_arrive(rempty_bar(rs));
              if (++rs == raw_stages) { rs = 0; rphase ^= 1u; }
              mbar_wait(empty_bar(stage), phase ^ 1u);
              goto signal;
            }
            mbar_wait(rfull_bar(rs), rphase);
            uint32_t w[4];
            const uint32_t src = raw0 + (uint32_t)rs * kRawStageBytes + raw_off;
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(src) : "memory");
            // table look-ups first (16 independent loads in flight), arithmetic and stores after
            uint32_t v[16];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
#pragma unroll
              for (int b = 0; b < 4; ++b) {
                const uint32_t off = b == 0 ? (w[c] << 7) & 0x7F80u : (w[c] >> (8 * b - 7)) & 0x7F80u;
                if (W4DBG(1)) v[c * 4 + b] = off;
                else asm("ld.shared.u32 %0, [%1];" : "=r"(v[c * 4 + b]) : "r"(lut_lane + off));
              }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(rempty_bar(rs));
            if (++rs == raw_stages) { rs = 0; rphase ^= 1u; }
            mbar_wait(empty_bar(stage), phase ^ 1u);
            const uint32_t dst = sB0 + (uint32_t)stage * stage_b_bytes + dst_off;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint32_t o[4];
              if (bf) {
#pragma unroll
                for (int b = 0; b < 4; ++b) asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(o[b]) : "r"(v[c * 4 + b]), "r"(sc2), "r"(nz2));
              } else {
#pragma unroll
                for (int b = 0; b < 4; ++b) asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(o[b]) : "r"(v[c * 4 + b]), "r"(sc2), "r"(nz2));
              }
              if (!W4DBG(16))
              asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (((uint32_t)(sub * 4 + c) ^ key) << 4)), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
              else asm volatile("" ::"r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]));
            }
            if (!W4DBG(2)) fence_proxy_async_smem();   // generic-proxy stores to this CTA's smem -> visible to tcgen05.mma
          } else {
            mbar_wait(empty_bar(stage), phase ^ 1u);             // LoRA k-block: B2 comes from TMA, nothing to expand
          }
        signal:
          if (!W4DBG(256)) asm volatile("bar.sync 2, 256;" ::: "memory");   // all dequant warps of this CTA wrote (and fenced) their part
          if (t == 0 && !W4DBG(4)) {
            // same signalling as a TMA-less CUTLASS producer: proxy fence by every writer, then a plain arrive
            if (leader) mbar_arrive(fb_local);
            else mbar_arrive_cluster(mapa_shared(fb_local, 0));
          } else if (t == 0) {
            if (leader) asm volatile("mbarrier.arrive.release.cluster.shared::cta.b64 _, [%0];" ::"r"(fb_local) : "memory");
            else asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(mapa_shared(fb_local, 0)) : "memory");
          }
          if (++stage == p.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp >= 2 && warp < 6) {
    // ================= epilogue (4 warps per CTA, one TMEM lane quarter each) =================
    const int q = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    SegmentIter it(p, unit, nunits, total_tiles, num_k_total);
    Segment sg;
    const int sk_g = p.stream_k ? p.sk_group : 1;
    unsigned int sk_epoch_v = 0;
    if (p.stream_k) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(sk_epoch_v) : "r"(sk_epoch_slot) : "memory");
    const int sk_groups = nunits / sk_g, sk_grp = unit / sk_g;
    const long long sk_total = (long long)(total_tiles / sk_g) * num_k_total;
    // fused all-reduce: this CTA's reduction items are ar_next, ar_next + gridDim.x, ... < ar_items (see ar_reduce)
    const int ar_slabs = (kBM * CTAS) / kArSlabRows;
    const int ar_items = p.ar_world > 1 ? ((total_tiles - p.ar_rank + p.ar_world - 1) / p.ar_world) * ar_slabs : 0;
    int ar_next = blockIdx.x;
    const uint32_t ar_poll_slot = sk_epoch_slot;     // stream-K is off in this mode: the slot is free
    while (it.next(sg)) {
      const int tile = sg.tile;
      const int bt = tile / tiles_per_batch;
      const int rem = tile - bt * tiles_per_batch;
      const int n_blk = rem / p.num_m;
      const int m_blk = rem - n_blk * p.num_m;
      const int m0 = m_blk * (kBM * CTAS) + (int)rank * kBM, n0 = n_blk * p.bn;
      const int ab0 = bt % p.nb0, ab1 = bt / p.nb0;
      const int bsel = p.b_map ? p.b_map[bt] : 0;
      if (bsel < 0) continue;                        // grouped GEMM: unused row tile
      const long long bias_off = (long long)bsel * p.bias_bs;
      mbar_wait_sleep(tfull_bar(acc), acc_phase, 128);
      tc_fence_after();
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < p.M;
      const long long doff = (long long)ab0 * p.d_bs0 + (long long)ab1 * p.d_bs1 + (long long)row * p.d_rs;
      const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)acc * kAccStride;
      const int col_limit = min(p.N, n0 + p.bn);
      const int nchunks = (p.bn + 31) / 32;
      const float beta = (p.beta_rows != nullptr && row_ok) ? p.beta_rows[row] : p.beta;
      if (sg.kb0 > 0) {
        // ---- stream-K contributor: publish this CTA's fp32 partial, the tile's owner finishes it ----
        float* slot = p.sk_ws + ((long long)blockIdx.x * kBM + (q * 32 + lane)) * kAccStride;
        for (int c = 0; c < nchunks; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(t0 + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            reinterpret_cast<uint4*>(slot + c * 32)[j] = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
        }
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (threadIdx.x == 64) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p.sk_flags + blockIdx.x), "r"(sk_epoch_v) : "memory");
      } else {
        // number of following units that hold later k-blocks of this tile (0 unless stream-K split it)
        int ncontrib = 0;
        if (sg.kb1 < num_k_total) {
          const long long tile_end = (long long)(tile / sk_g + 1) * num_k_total;
          for (int v = sk_grp + 1; v < sk_groups && sk_total * v / sk_groups < tile_end; ++v) ++ncontrib;
          for (int v = 1; v <= ncontrib; ++v) {
            const unsigned int* f = p.sk_flags + (unit + v * sk_g) * CTAS + rank;   // same member of the next group
            unsigned int seen;
            const long long tw = clock64();
            do {
              asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(f) : "memory");
              if (clock64() - tw > 20000000000ll) {
                printf("b200 gemm: stream-K wait timed out (block %d waits for block %d)\n", blockIdx.x, (unit + v * sk_g) * CTAS + (int)rank);
                __trap();
              }
            } while (seen != sk_epoch_v);
          }
        }
      for (int c = 0; c < nchunks; ++c) {
        const int col0 = n0 + c * 32;
        if (col0 >= col_limit) break;
        uint32_t r[32];
        tmem_ld_32x32(t0 + c * 32, r);
        tmem_ld_wait();
        for (int v = 1; v <= ncontrib; ++v) {
          const float* slot = p.sk_ws + ((long long)((unit + v * sk_g) * CTAS + rank) * kBM + (q * 32 + lane)) * kAccStride + c * 32;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 w = reinterpret_cast<const float4*>(slot)[j];
            r[4 * j] = __float_as_uint(__uint_as_float(r[4 * j]) + w.x);
            r[4 * j + 1] = __float_as_uint(__uint_as_float(r[4 * j + 1]) + w.y);
            r[4 * j + 2] = __float_as_uint(__uint_as_float(r[4 * j + 2]) + w.z);
            r[4 * j + 3] = __float_as_uint(__uint_as_float(r[4 * j + 3]) + w.w);
          }
        }
        if (p.out_dtype == B200_DT_BF16)
          epilogue_store<__nv_bfloat16>(p, reinterpret_cast<__nv_bfloat16*>(p.D) + doff, col0, col_limit, r, row_ok, beta, bias_off);
        else if (p.out_dtype == B200_DT_FP32)
          epilogue_store<float>(p, reinterpret_cast<float*>(p.D) + doff, col0, col_limit, r, row_ok, beta, bias_off);
        else
          epilogue_store<__half>(p, reinterpret_cast<__half*>(p.D) + doff, col0, col_limit, r, row_ok, beta, bias_off);
      }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (leader) mbar_arrive(tempty_bar(acc));
        else mbar_arrive_cluster(mapa_shared(tempty_bar(acc), 0));
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1u;
      if (!W4 && p.ar_world > 1) {
        // this CTA's rows of the tile are in D: one system-scope fence + release by one thread covers the stores the
        // other epilogue threads made before the barrier
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (threadIdx.x == 64) {
          __threadfence_system();
          unsigned int* f = p.ar_tile_flags[tile % p.ar_world] + ((long long)(tile * 2 + (int)rank) * 8 + p.ar_rank);
          asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f), "r"(p.ar_epoch) : "memory");
        }
        // while the tensor pipe works on this CTA's next tile: sum up to two owned row slabs whose partials have arrived
        ar_next = ar_reduce<CTAS>(p, ar_next, ar_items, 2, false, total_tiles, ar_poll_slot);
      }
    }
    if (!W4 && p.ar_world > 1) {
      // ---- the rest of this CTA's share of the reduction, then the cross-rank completion barrier ----
      ar_reduce<CTAS>(p, ar_next, ar_items, 1 << 30, true, total_tiles, ar_poll_slot);
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (threadIdx.x == 64) {
        const int W = p.ar_world;
        __threadfence_system();
        if (atomicAdd(p.ar_ticket, 1u) == gridDim.x - 1) {
          *((volatile unsigned int*)p.ar_ticket) = 0u;
          __threadfence_system();
          for (int r = 0; r < W; ++r)
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.ar_sync[r] + p.ar_channel * 8 + p.ar_rank), "r"(p.ar_epoch) : "memory");
          for (int r = 0; r < W; ++r) {
            const unsigned int* f = p.ar_sync[p.ar_rank] + p.ar_channel * 8 + r;
            unsigned int seen;
            const long long tw = clock64();
            do {
              asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(f) : "memory");
              if (clock64() - tw > 20000000000ll) {
                printf("b200 gemm+allreduce: rank %d timed out waiting for rank %d to finish its share (flag %u, epoch %u)\n",
                       p.ar_rank, r, seen, p.ar_epoch);
                __trap();
              }
            } while ((int)(seen - p.ar_epoch) < 0);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if constexpr (CTAS == 2) cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    if constexpr (CTAS == 1) tmem_dealloc(tmem_base, 512);
    else tmem_dealloc_2sm(tmem_base, 512);
  }
}

template <int ES, int CTAS>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                    const __grid_constant__ GemmKernelParams p) {
  gemm_body<ES, CTAS, false>(tmA, tmB, tmA2, tmB2, p, nullptr);
}

// Same pipeline with the B operand expanded from 4-bit codes inside the kernel (NF4 / GPTQ / AWQ weights, optionally
// sharded over peer GPUs): no 16-bit copy of the weight ever exists in global memory.
__global__ void __launch_bounds__(kW4Threads, 1)
gemm_w4_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2,
                       const __grid_constant__ CUtensorMap tmB2, const __grid_constant__ W4Params w4,
                       const __grid_constant__ GemmKernelParams p) {
  gemm_body<2, 2, true>(tmA, tmA, tmA2, tmB2, p, &w4);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int g_num_sms = 0;
static int g_max_smem = 0;
static int g_force_ctas = 0;  // B200_GEMM_CTAS=1|2 pins the CTA-group size (debugging / A-B timing)

static int device_props() {
  if (g_num_sms == 0) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&g_max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    const char* f = getenv("B200_GEMM_CTAS");
    if (f) g_force_ctas = atoi(f);
  }
  return 0;
}

// Modelled cycles per k-block (four k-steps) of one unit tile (128*ctas x bn): the slower of the
// tensor pipe (2*bn) and this SM's shared-memory port, which carries the TMA fill and the UMMA
// operand reads of the same bytes ((16 KB + 128 B * bn / ctas) * 2 at 128 B/cycle), plus issue and
// barrier overhead. A CTA pair halves the B bytes per SM, which is what makes 256-wide tiles
// tensor-bound.
static double kblock_cost(int ctas, int bn) {
  const double t_mma = 2.0 * bn;
  const double t_smem = 256.0 + 2.0 * bn / ctas;
  return (t_mma > t_smem ? t_mma : t_smem) + 40.0;
}

struct GemmConfig {
  int ctas, bn, stream_k;
  double cost;
};

static GemmConfig pick_config(int M, int N, int batch, int nkt, int b_mn, int EPR, int num_sms, bool allow_sk) {
  GemmConfig best = {1, 256, 0, 1e30};
  static const bool no_sk = getenv("B200_GEMM_NO_STREAMK") != nullptr;     // tuning switches are read once per process
  for (int ctas = 2; ctas >= 1; --ctas) {
    if (g_force_ctas && ctas != g_force_ctas) continue;
    if (ctas == 2 && (M <= kBM || num_sms < 2)) continue;
    const int step = b_mn ? EPR * ctas : 16 * ctas;  // MN-major B is staged in 128-byte column chunks per CTA
    const int units = num_sms / ctas;
    for (int bn = 256; bn >= 16 * ctas; bn -= 16) {
      if (bn % step) continue;
      if (bn > 32 && bn > 2 * ((N + 15) / 16 * 16)) continue;  // do not pick tiles far wider than N
      const long long tiles = (long long)batch * ceil_div(M, kBM * ctas) * ceil_div(N, bn);
      const long long waves = (tiles + units - 1) / units;
      const double per = kblock_cost(ctas, bn);
      const double classic = (double)waves * nkt * per;
      if (classic < best.cost * 0.999) best = {ctas, bn, 0, classic};
      if (allow_sk && !no_sk && tiles % units != 0 && tiles * nkt >= 4ll * units) {
        // equal k-block ranges per unit; ~10 k-blocks' worth of partial-tile write/read/wait per unit
        const double sk = ((double)tiles * nkt / units + 10.0) * per;
        // measured: worth it only when classic scheduling wastes a large part of a wave (few tiles);
        // with many waves the uneven k-ranges cost more L2 locality than the tail they recover
        if (waves <= 4 && sk < classic * 0.85 && sk < best.cost * 0.97) best = {ctas, bn, 1, sk};
      }
    }
  }
  return best;
}

template <int ES, int CTAS>
static int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmA2, const CUtensorMap& tmB2,
                  const GemmKernelParams& p, int grid, int smem, cudaStream_t stream) {
  static bool configured = false;
  if (!configured) {
    B200_CHECK(cudaFuncSetAttribute(gemm_tcgen05_kernel<ES, CTAS>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_max_smem));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CTAS;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CHECK(cudaLaunchKernelEx(&cfg, gemm_tcgen05_kernel<ES, CTAS>, tmA, tmB, tmA2, tmB2, p));
  return 0;
}

}  // namespace b200

using namespace b200;

// A full-size workspace (b200_gemm_workspace_bytes()) carries the stream-K epoch on the device, in its last 16 bytes:
// every stream-K launch on that workspace (b200_gemm and b200_gemm_w4 alike) then takes its epoch from there and the launch
// can be captured in a CUDA graph. A smaller workspace falls back to the host-side counters.
static unsigned int* sk_state_of(const b200_gemm_args* a) {
  const long long full = 160ll * kBM * kAccStride * 4 + 160ll * 4 + 16;
  if (a->workspace_bytes < full) return nullptr;
  return reinterpret_cast<unsigned int*>(reinterpret_cast<char*>(a->workspace) + full - 16);
}

B200_API long long b200_gemm_workspace_bytes(void) {
  // one fp32 128 x 256 partial-tile slot per CTA + one flag per CTA, sized for the largest part
  return 160ll * kBM * kAccStride * 4 + 160ll * 4 + 16;   // + the device epoch state of stream-K
}

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

  GemmKernelParams p = {};
  p.D = a->D;
  p.bias = a->bias;
  p.M = a->M; p.N = a->N; p.K = a->K; p.K2 = lora ? a->K2 : 0;
  p.batch = a->batch; p.nb0 = nb0; p.b_div = b_div;
  p.nb0_b = nb0 / b_div > 0 ? nb0 / b_div : 1;
  p.b_map = a->b_map; p.bias_bs = a->b_map ? a->bias_bs : 0;
  if (a->b_map) {
    if (a->b_count < 1 || a->ar) return B200_ERR_BAD_SHAPE;
    p.nb0_b = a->b_count;
  }
  p.d_rs = a->d_rs; p.d_bs0 = a->d_bs0; p.d_bs1 = a->d_bs1;
  p.alpha = a->alpha; p.beta = a->beta;
  p.beta_rows = (a->batch == 1) ? a->beta_rows : nullptr;
  p.out_dtype = a->out_dtype; p.bias_dtype = a->bias_dtype;
  p.a_mn = a->a_mn ? 1 : 0; p.b_mn = a->b_mn ? 1 : 0;
  p.num_k = ceil_div(a->K, EPR);
  const int nkt = p.num_k + (lora ? 1 : 0);
  const long long sk_need = (long long)g_num_sms * kBM * kAccStride * 4 + (long long)g_num_sms * 4;
  const b200_gemm_allreduce* ar = a->ar;
  if (ar) {
    if (ar->world < 2 || ar->world > 8 || ar->rank < 0 || ar->rank >= ar->world || !ar->d_mc || !ar->ticket) return B200_ERR_BAD_SHAPE;
    if (a->batch != 1 || a->out_dtype != B200_DT_BF16 || a->beta != 0.f || a->beta_rows) return B200_ERR_UNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(a->D) | reinterpret_cast<uintptr_t>(ar->d_mc)) & 15 || a->d_rs % 8 || a->N % 8) return B200_ERR_BAD_ALIGN;
  }
  const bool allow_sk = !ar && !a->b_map && a->workspace != nullptr && a->workspace_bytes >= sk_need;
  const GemmConfig cfg = pick_config(a->M, a->N, a->batch, nkt, p.b_mn, EPR, g_num_sms, allow_sk);
  const int ctas = cfg.ctas;
  p.bn = cfg.bn;
  const int bn_cta = p.bn / ctas;
  p.num_m = ceil_div(a->M, kBM * ctas);
  p.num_n = ceil_div(a->N, p.bn);
  p.stream_k = cfg.stream_k;
  p.sk_group = 1;
  p.sk_ws = nullptr; p.sk_flags = nullptr; p.sk_epoch = 0; p.sk_state = nullptr;
  if (cfg.stream_k) {
    static const bool sk_flat = getenv("B200_GEMM_SK_FLAT") != nullptr;
    // lock-step groups when the M-tiles of a panel fit a group and at most ~5 % of the units are left without one
    const int units_sk = g_num_sms / ctas;
    if (p.num_m > 1 && p.num_m <= units_sk && (units_sk % p.num_m) * 20 <= units_sk && !sk_flat)
      p.sk_group = p.num_m;
    static unsigned int epoch = 0;
    p.sk_ws = reinterpret_cast<float*>(a->workspace);
    p.sk_flags = reinterpret_cast<unsigned int*>(reinterpret_cast<char*>(a->workspace) +
                                                 (long long)g_num_sms * kBM * kAccStride * 4);
    p.sk_epoch = ++epoch;
    p.sk_state = sk_state_of(a);
  }
  const int stage_bytes = kStageABytes + bn_cta * 128;
  int stages = (g_max_smem - 2048) / stage_bytes;
  // 6 stages sustain the pipe (5..7 measure the same); the 32 KB left free let bandwidth-bound kernels of other streams
  // (the next weight's dequant, the ZeRO-3 shard pull) co-reside with the persistent GEMM CTAs
  if (stages > 6) stages = 6;
  static const int env_stages = [] { const char* e = getenv("B200_GEMM_STAGES"); return e ? atoi(e) : 0; }();
  if (env_stages >= 2 && env_stages < stages) stages = env_stages;
  if (stages < 2) return B200_ERR_UNSUPPORTED;
  p.stages = stages;
  const int umma_m = kBM * ctas;
  const uint32_t fmt = ES == 4 ? 2u : (a->in_dtype == B200_DT_BF16 ? 1u : 0u);  // tf32 / bf16 / f16
  const uint32_t idesc_base = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(p.bn >> 3) << 17) |
                              ((uint32_t)(umma_m >> 4) << 24);
  p.idesc_main = idesc_base | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16);
  p.idesc_lora = idesc_base;
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
    if (!p.b_mn) { dims[0] = a->K; dims[1] = a->N; box[0] = EPR; box[1] = bn_cta; }
    else         { dims[0] = a->N; dims[1] = a->K; box[0] = EPR; box[1] = EPR; }
    dims[2] = p.nb0_b; dims[3] = a->b_map ? 1 : nb1_b; box[2] = 1; box[3] = 1;
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
    uint32_t boxb[4] = {(uint32_t)EPR, (uint32_t)bn_cta, 1, 1};
    rc = make_tensor_map(&tmB2, a->B2, ES, 4, dimsb, strb, boxb, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  } else {
    tmA2 = tmA;
    tmB2 = tmB;
  }

  const long long total_tiles = (long long)p.num_m * p.num_n * p.batch;
  p.ar_world = 0; p.ar_rank = 0; p.ar_channel = 0; p.ar_epoch = 0; p.ar_mc = nullptr; p.ar_ticket = nullptr;
  p.ar_d[0] = p.ar_d[1] = nullptr;
  for (int r = 0; r < 8; ++r) { p.ar_tile_flags[r] = nullptr; p.ar_sync[r] = nullptr; }
  if (ar) {
    if (total_tiles > ar->max_tiles) return B200_ERR_TOO_LARGE;
    p.ar_world = ar->world; p.ar_rank = ar->rank; p.ar_channel = ar->channel; p.ar_epoch = ar->epoch;
    p.ar_mc = reinterpret_cast<char*>(ar->d_mc); p.ar_ticket = ar->ticket;
    // two ranks: the switch has nothing to save (each byte crosses one link either way) and plain peer loads are faster
    static const int force_peer = [] { const char* e = getenv("B200_GEMM_AR_PEER"); return e ? atoi(e) : -1; }();
    if (ar->d_ptrs && ar->world == 2 && force_peer != 0) {
      p.ar_mc = nullptr;
      p.ar_d[0] = reinterpret_cast<char*>(const_cast<void*>(ar->d_ptrs[0]));
      p.ar_d[1] = reinterpret_cast<char*>(const_cast<void*>(ar->d_ptrs[1]));
    }
    for (int r = 0; r < ar->world; ++r) {
      p.ar_tile_flags[r] = reinterpret_cast<unsigned int*>(const_cast<void*>(ar->tile_flag_ptrs[r]));
      p.ar_sync[r] = reinterpret_cast<unsigned int*>(const_cast<void*>(ar->sync_flag_ptrs[r]));
    }
  }
  const int units = g_num_sms / ctas;
  const int grid = (cfg.stream_k ? units : (int)(total_tiles < units ? total_tiles : units)) * ctas;
  const int smem = stages * stage_bytes + 8 * (2 * stages + 5) + 1024;
  if (ES == 2) rc = ctas == 2 ? launch<2, 2>(tmA, tmB, tmA2, tmB2, p, grid, smem, stream)
                              : launch<2, 1>(tmA, tmB, tmA2, tmB2, p, grid, smem, stream);
  else rc = ctas == 2 ? launch<4, 2>(tmA, tmB, tmA2, tmB2, p, grid, smem, stream)
                      : launch<4, 1>(tmA, tmB, tmA2, tmB2, p, grid, smem, stream);
  if (rc) return rc;
  B200_LAUNCH_CHECK();
  return 0;
}

B200_API int b200_gemm_w4(const b200_gemm_args* a, const b200_w4_desc* w, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  constexpr int ES = 2, EPR = 64, ctas = 2, bn = 256, bn_cta = 128;
  if (a->M <= 0 || a->N <= 0 || a->K <= 0) return B200_ERR_BAD_SHAPE;
  if (a->in_dtype != B200_DT_BF16 && a->in_dtype != B200_DT_FP16) return B200_ERR_UNSUPPORTED;
  if (a->batch > 1 || a->a_mn || a->M <= kBM || a->N % bn || a->K % EPR) return B200_ERR_UNSUPPORTED;
  if (w->nshards < 1 || w->nshards > 8 || w->group % 64 || w->kw % 64 || w->kw % w->group) return B200_ERR_UNSUPPORTED;
  const long long rps = w->nshards > 1 ? w->rows_per_shard : w->rows;
  if (rps * w->nshards < w->rows || rps % (a->b_mn ? 64 : 128)) return B200_ERR_UNSUPPORTED;
  if (!a->b_mn ? (a->N != w->rows || a->K != w->kw) : (a->N != w->kw || a->K != w->rows)) return B200_ERR_BAD_SHAPE;
  const bool lora = a->K2 > 0 && a->A2 && a->B2;
  if (lora && a->K2 > EPR) return B200_ERR_UNSUPPORTED;
  auto al16 = [&](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!al16(a->A) || (a->a_rs * ES) % 16) return B200_ERR_BAD_ALIGN;
  if (lora && (!al16(a->A2) || !al16(a->B2) || (a->a2_rs * ES) % 16 || (a->b2_rs * ES) % 16)) return B200_ERR_BAD_ALIGN;
  for (int i = 0; i < w->nshards; ++i)
    if (!al16(w->packed[i])) return B200_ERR_BAD_ALIGN;
  int rc = device_props();
  if (rc) return rc;

  GemmKernelParams p = {};
  p.D = a->D;
  p.bias = a->bias;
  p.M = a->M; p.N = a->N; p.K = a->K; p.K2 = lora ? a->K2 : 0;
  p.batch = 1; p.nb0 = 1; p.b_div = 1; p.nb0_b = 1;
  p.d_rs = a->d_rs; p.d_bs0 = 0; p.d_bs1 = 0;
  p.alpha = a->alpha; p.beta = a->beta;
  p.beta_rows = (a->batch == 1) ? a->beta_rows : nullptr;
  p.out_dtype = a->out_dtype; p.bias_dtype = a->bias_dtype;
  p.a_mn = 0; p.b_mn = a->b_mn ? 1 : 0;
  p.num_k = a->K / EPR;
  const int nkt = p.num_k + (lora ? 1 : 0);
  p.bn = bn;
  p.num_m = ceil_div(a->M, kBM * ctas);
  p.num_n = a->N / bn;
  const int units = g_num_sms / ctas;
  const long long tiles = (long long)p.num_m * p.num_n;
  const long long sk_need = (long long)g_num_sms * kBM * kAccStride * 4 + (long long)g_num_sms * 4;
  static const bool no_sk_w4 = getenv("B200_GEMM_NO_STREAMK") != nullptr;
  const bool allow_sk = a->workspace != nullptr && a->workspace_bytes >= sk_need && !no_sk_w4;
  const long long waves = (tiles + units - 1) / units;
  const double classic = (double)waves * nkt, sk = (double)tiles * nkt / units + 10.0;
  p.stream_k = (allow_sk && tiles % units != 0 && tiles * nkt >= 4ll * units && waves <= 4 && sk < classic * 0.85) ? 1 : 0;
  p.sk_group = 1;
  p.sk_ws = nullptr; p.sk_flags = nullptr; p.sk_epoch = 0; p.sk_state = nullptr;
  if (p.stream_k) {
    static unsigned int epoch = 0x40000000u;   // own epoch space: flags are shared with b200_gemm's workspace
    p.sk_ws = reinterpret_cast<float*>(a->workspace);
    p.sk_flags = reinterpret_cast<unsigned int*>(reinterpret_cast<char*>(a->workspace) +
                                                 (long long)g_num_sms * kBM * kAccStride * 4);
    p.sk_epoch = ++epoch;
    p.sk_state = sk_state_of(a);
  }
  const int stage_bytes = kStageABytes + bn_cta * 128;
  int raw_stages = 8, lut_bytes = kLutBytes;
  static const int env_raw = [] { const char* e = getenv("B200_W4_RAW"); return e ? atoi(e) : -1; }();
  static const bool env_nolut = getenv("B200_W4_NOLUT") != nullptr;
  if (env_raw >= 0) raw_stages = env_raw;
  if (env_nolut) lut_bytes = 0;
  int stages = (g_max_smem - 1024 - lut_bytes - raw_stages * kRawStageBytes - 8 * (2 * 8 + 5 + 2 * raw_stages)) / stage_bytes;
  if (stages > 8) stages = 8;
  if (stages < 3) return B200_ERR_UNSUPPORTED;
  p.stages = stages;
  const uint32_t fmt = a->in_dtype == B200_DT_BF16 ? 1u : 0u;
  const uint32_t idesc_base = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(bn >> 3) << 17) |
                              ((uint32_t)((kBM * ctas) >> 4) << 24);
  p.idesc_main = idesc_base | ((uint32_t)p.b_mn << 16);
  p.idesc_lora = idesc_base;
  const int out_es = a->out_dtype == B200_DT_FP32 ? 4 : 2;
  const int per_vec = 16 / out_es;
  p.vec_ok = ((reinterpret_cast<uintptr_t>(a->D) & 15) == 0) && (a->d_rs % per_vec == 0);

  CUtensorMap tmA, tmA2, tmB2;
  {
    uint64_t dims[4] = {(uint64_t)a->K, (uint64_t)a->M, 1, 1}, str[3] = {(uint64_t)a->a_rs * ES, 0, 0};
    uint32_t box[4] = {(uint32_t)EPR, (uint32_t)kBM, 1, 1};
    rc = make_tensor_map(&tmA, a->A, ES, 4, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  if (lora) {
    uint64_t dims[4] = {(uint64_t)a->K2, (uint64_t)a->M, 1, 1}, str[3] = {(uint64_t)a->a2_rs * ES, 0, 0};
    uint32_t box[4] = {(uint32_t)EPR, (uint32_t)kBM, 1, 1};
    rc = make_tensor_map(&tmA2, a->A2, ES, 4, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    uint64_t dimsb[4] = {(uint64_t)a->K2, (uint64_t)a->N, 1, 1}, strb[3] = {(uint64_t)a->b2_rs * ES, 0, 0};
    uint32_t boxb[4] = {(uint32_t)EPR, (uint32_t)bn_cta, 1, 1};
    rc = make_tensor_map(&tmB2, a->B2, ES, 4, dimsb, strb, boxb, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  } else {
    tmA2 = tmA;
    tmB2 = tmA;
  }
  W4Params wp;
  memset(&wp, 0, sizeof(wp));
  for (int i = 0; i < 8; ++i) {
    const int j = i < w->nshards ? i : 0;
    const long long rows_here = w->nshards > 1 ? rps : w->rows;
    uint64_t dims[2] = {(uint64_t)w->kw / 2, (uint64_t)rows_here}, str[1] = {(uint64_t)w->kw / 2};
    uint32_t box[2] = {a->b_mn ? 64u : 32u, a->b_mn ? 64u : 128u};
    rc = make_tensor_map(&wp.tmW[i], w->packed[j], 1, 2, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    wp.scale[i] = w->scale[j];
    wp.zs[i] = w->zs[j];
  }
  wp.lut = w->lut;
  wp.nshards = w->nshards;
  wp.rows_per_shard = (int)rps;
  wp.group = w->group;
  wp.kw = w->kw;
  wp.raw_stages = raw_stages;
  wp.lut_bytes = lut_bytes;
  wp.is_bf16 = a->in_dtype == B200_DT_BF16;
#ifdef B200_W4_EXPERIMENT
  { const char* d = getenv("B200_W4_DEBUG"); wp.debug = d ? atoi(d) : 0; }
#else
  wp.debug = 0;
#endif

  const int grid = (p.stream_k ? units : (int)(tiles < units ? tiles : units)) * ctas;
  const int smem = stages * stage_bytes + raw_stages * kRawStageBytes + lut_bytes + 8 * (2 * stages + 5 + 2 * raw_stages) + 1024;
  static bool configured = false;
  if (!configured) {
    B200_CHECK(cudaFuncSetAttribute(gemm_w4_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, g_max_smem));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kW4Threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = ctas;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CHECK(cudaLaunchKernelEx(&cfg, gemm_w4_tcgen05_kernel, tmA, tmA2, tmB2, wp, p));
  B200_LAUNCH_CHECK();
  return 0;
}
