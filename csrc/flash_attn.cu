// K4 — flash attention forward and backward over the fused qkv projection: softmax(Q K^T * scale + causal mask) V with
// grouped KV heads, no L x L tensor in memory and no materialised mask. The reference builds a boolean L x L mask on
// every forward (LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:93-95) and otherwise relies on
// SDPA / flash-attn behind AutoModelForCausalLM (Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:101-112).
//
// q, k and v are read in place from qkv [B*S, (H + 2 Hkv) * D] (row stride W), the context is written as [B*S, H*D],
// the row log-sum-exp as [B, H, S] fp32. Tiles are 64 queries x 64 keys; operands are staged in shared memory with
// cp.async (16-byte chunks, XOR-swizzled so every ldmatrix is conflict-free) and multiplied with
// mma.sync.m16n8k16 (bf16 / fp16 in, fp32 accumulate). At the sequence lengths of the training recipes (512) a
// (batch, head) problem is 512 x 512 x 128: eight 64-row query blocks, so the kernel is sized for many small CTAs
// (640 per layer for Qwen3-14B) rather than for one large TMEM-resident tile.
//
// Forward: one CTA (4 warps, 16 query rows each) per (query block, head, batch); K/V blocks double-buffered; online
// softmax in the exp2 domain; heavy (late) causal blocks are scheduled first.
// Backward: one CTA (8 warps) per (key block, head, batch) holds K_j, V_j and walks the query blocks i >= j:
//   S = Q K^T, P = exp2(S * c - lse), dP = dO V^T, dS = P (dP - delta) * scale,
//   dV += P^T dO, dK += dS^T Q (registers, flushed once), dQ_i += dS K (fp32 red.global into the accumulation buffer).
// dq/dk/dv accumulate in one fp32 buffer shaped like qkv (the q heads of a KV group add into the same dk/dv rows), which a
// final pass converts to the 16-bit gradient of the fused projection.
#include "common.cuh"
#include "../include/b200_ops.h"

#include <stdlib.h>

namespace b200 {

struct FlashParams {
  const void* qkv;     // [B*S, W] 16-bit
  void* out;           // fwd: [B*S, H*D] 16-bit
  float* lse;          // [B, H, S]
  const void* dout;    // bwd: [B*S, H*D]
  const float* delta;  // bwd: [B, H, S] rowsum(dO * O)
  float* dqkv32;       // bwd: [B*S, W] fp32, zeroed by the caller
  const int* cu;       // fwd only, may be null: packed sequences, sequence b = rows [cu[b], cu[b+1]) (then S = longest)
  int B, S, H, Hkv, D;
  long long W;
  float scale, scale_log2;
  unsigned long long seed;   // attention-probability dropout (nn.MultiheadAttention(dropout=p)): Philox, regenerated in backward
  uint32_t drop_thresh;      // 0 = no dropout; P(drop) = thresh / 2^32
  float inv_keep;
  const unsigned long long* seed_off;   // optional per-replay seed offset in device memory (CUDA-graph captured steps)
};

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// Philox4x32-10. One call covers the 2 x 2 patch a thread holds of a score tile: rows (r, r + 8), keys (2 kp, 2 kp + 1), so the
// mask of an element depends only on (seed, batch*head, row, key) and forward / backward regenerate it identically.
__device__ __forceinline__ uint4 flash_philox(unsigned long long seed, unsigned long long q) {
  constexpr uint32_t kA = 0xD2511F53u, kB = 0xCD9E8D57u, kW0 = 0x9E3779B9u, kW1 = 0xBB67AE85u;
  uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
  uint4 ctr = make_uint4((uint32_t)q, (uint32_t)(q >> 32), 0x6a09e667u, 0u);
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
// keep/scale factors of the patch (row_lo, row_lo + 8) x (key0, key0 + 1), key0 even: m[c] for fragment element c
__device__ __forceinline__ void flash_drop4(const FlashParams& p, long long bh, int row_lo, int key0, float (&m)[4]) {
  const unsigned long long q = ((unsigned long long)bh * (unsigned long long)p.S + (unsigned long long)row_lo) *
                               (unsigned long long)((p.S + 1) / 2) + (unsigned long long)(key0 >> 1);
  const uint4 r = flash_philox(p.seed + (p.seed_off ? __ldg(p.seed_off) : 0ull), q);
  m[0] = r.x >= p.drop_thresh ? p.inv_keep : 0.f;
  m[1] = r.y >= p.drop_thresh ? p.inv_keep : 0.f;
  m[2] = r.z >= p.drop_thresh ? p.inv_keep : 0.f;
  m[3] = r.w >= p.drop_thresh ? p.inv_keep : 0.f;
}

// byte offset of 16-byte chunk `c` of row `r` in a tile with CH chunks per row. Rows of 128 / 256 bytes (CH = 8, 16) XOR the
// chunk with the row's low 3 bits; rows of 64 / 192 bytes (CH = 4, 12: head_dim 32, 96) alternate between the two halves of
// a 128-byte line by themselves, so XOR-ing the low 2 bits with (r >> 1) & 3 is what spreads 8 consecutive rows over the 8
// 16-byte slots. Either way every ldmatrix (8 rows x 16 bytes) is conflict-free.
template <int CH>
__device__ __forceinline__ uint32_t tile_off(int r, int c) {
  if constexpr (CH % 8 == 0) return (uint32_t)((r * CH + (c ^ (r & 7))) << 4);
  else return (uint32_t)((r * CH + (c ^ ((r >> 1) & 3))) << 4);
}

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int n = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}

template <bool BF16>
__device__ __forceinline__ void mma_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  if (BF16)
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  else
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <bool BF16>
__device__ __forceinline__ uint32_t pack16(float a, float b) {
  if (BF16) return pack_bf16x2(a, b);
  const __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}

// rows [row0, row0 + 64) x D columns starting at `base` (element pointer of row 0, column 0 of this head) -> swizzled
// tile; rows >= S are zero-filled
template <int D, int NT>
__device__ __forceinline__ void load_tile64(uint32_t sdst, const uint16_t* base, long long stride, int row0, int S, int tid) {
  constexpr int CH = D / 8;
#pragma unroll
  for (int i = tid; i < 64 * CH; i += NT) {
    const int r = i / CH, c = i % CH;
    const bool ok = row0 + r < S;
    const uint16_t* src = base + (long long)(ok ? row0 + r : 0) * stride + c * 8;
    cp_async16(sdst + tile_off<CH>(r, c), src, ok);
  }
}

// A fragment (16 rows x 16 k) of a [rows][k] tile: rows r0.., k chunk pair 2*ks
template <int CH>
__device__ __forceinline__ void load_a(uint32_t tile, int r0, int ks, int lane, uint32_t (&a)[4]) {
  const int m = lane >> 3, r = lane & 7;
  ldsm_x4(tile + tile_off<CH>(r0 + (m & 1) * 8 + r, 2 * ks + (m >> 1)), a[0], a[1], a[2], a[3]);
}
// B fragments of two n-tiles (16 n rows) for one k-step from a [n][k] tile (k contiguous): b[0..1] n-tile 0, b[2..3] n-tile 1
template <int CH>
__device__ __forceinline__ void load_b_nk(uint32_t tile, int n0, int ks, int lane, uint32_t (&b)[4]) {
  const int m = lane >> 3, r = lane & 7;
  ldsm_x4(tile + tile_off<CH>(n0 + (m >> 1) * 8 + r, 2 * ks + (m & 1)), b[0], b[1], b[2], b[3]);
}
// B fragments of two n-tiles (n chunk c0, c0 + 1) for k rows k0..k0+15 from a [k][n] tile (n contiguous), transposing
template <int CH>
__device__ __forceinline__ void load_b_kn(uint32_t tile, int k0, int c0, int lane, uint32_t (&b)[4]) {
  const int m = lane >> 3, r = lane & 7;
  ldsm_x4_t(tile + tile_off<CH>(k0 + (m & 1) * 8 + r, c0 + (m >> 1)), b[0], b[1], b[2], b[3]);
}
// A fragment of the TRANSPOSE of a [k][m] tile (m contiguous): A rows m0..m0+15 = tile columns, A k = tile rows k0..k0+15
template <int CH>
__device__ __forceinline__ void load_a_t(uint32_t tile, int k0, int mc0, int lane, uint32_t (&a)[4]) {
  const int m = lane >> 3, r = lane & 7;
  ldsm_x4_t(tile + tile_off<CH>(k0 + (m >> 1) * 8 + r, mc0 + (m & 1)), a[0], a[1], a[2], a[3]);
}

// ------------------------------------------------------------------------------------------------ forward
template <bool BF16, int D, bool CAUSAL>
__global__ void __launch_bounds__(128)
flash_fwd_kernel(const FlashParams p) {
  constexpr int CH = D / 8, KS = D / 16, NT = 128;
  constexpr uint32_t kTile = 64 * D * 2;
  extern __shared__ __align__(128) uint8_t fsm[];
  const uint32_t sQ = smem_u32(fsm), sK = sQ + kTile, sV = sK + 2 * kTile;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int qb = CAUSAL ? (int)(gridDim.x - 1 - blockIdx.x) : (int)blockIdx.x;   // longest blocks first
  const int h = blockIdx.y, b = blockIdx.z, hk = h / (p.H / p.Hkv);
  const long long row_base = p.cu ? (long long)p.cu[b] : (long long)b * p.S;
  const int S = p.cu ? p.cu[b + 1] - p.cu[b] : p.S, q0 = qb * 64;
  if (q0 >= S) return;
  const uint16_t* base = reinterpret_cast<const uint16_t*>(p.qkv) + row_base * p.W;
  const uint16_t* gq = base + (long long)h * D;
  const uint16_t* gk = base + (long long)(p.H + hk) * D;
  const uint16_t* gv = base + (long long)(p.H + p.Hkv + hk) * D;
  const int nkb = CAUSAL ? min(qb + 1, (S + 63) / 64) : (S + 63) / 64;

  load_tile64<D, NT>(sQ, gq, p.W, q0, S, tid);
  load_tile64<D, NT>(sK, gk, p.W, 0, S, tid);
  load_tile64<D, NT>(sV, gv, p.W, 0, S, tid);
  cp_async_commit();

  uint32_t qf[KS][4];
  float o[D / 8][4];
#pragma unroll
  for (int i = 0; i < D / 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) o[i][j] = 0.f;
  float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
  const int rowA = q0 + warp * 16 + g, rowB = rowA + 8;

  for (int j = 0; j < nkb; ++j) {
    const uint32_t cK = sK + (j & 1) * kTile, cV = sV + (j & 1) * kTile;
    if (j + 1 < nkb) {
      load_tile64<D, NT>(sK + ((j + 1) & 1) * kTile, gk, p.W, (j + 1) * 64, S, tid);
      load_tile64<D, NT>(sV + ((j + 1) & 1) * kTile, gv, p.W, (j + 1) * 64, S, tid);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (j == 0) {
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) load_a<CH>(sQ, warp * 16, ks, lane, qf[ks]);
    }
    float s[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int c = 0; c < 4; ++c) s[i][c] = 0.f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        uint32_t bf[4];
        load_b_nk<CH>(cK, np * 16, ks, lane, bf);
        mma_16816<BF16>(s[2 * np], qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], bf[0], bf[1]);
        mma_16816<BF16>(s[2 * np + 1], qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], bf[2], bf[3]);
      }
    }
    // scale into the exp2 domain, mask, row max
    const int k0 = j * 64;
    const bool edge = (CAUSAL && j == qb) || (k0 + 64 > S);
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float v = s[i][c] * p.scale_log2;
        if (edge) {
          const int key = k0 + i * 8 + 2 * t + (c & 1), row = (c & 2) ? rowB : rowA;
          if (key >= S || (CAUSAL && key > row)) v = -INFINITY;
        }
        s[i][c] = v;
        mx[c >> 1] = fmaxf(mx[c >> 1], v);
      }
    }
    float alpha[2], mnew[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      mnew[r] = fmaxf(mrow[r], mx[r]);
      const float mref = (mnew[r] == -INFINITY) ? 0.f : mnew[r];
      alpha[r] = exp2f(mrow[r] - mref);
      mrow[r] = mnew[r];
      mnew[r] = mref;
    }
    float rs[2] = {0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float e = exp2f(s[i][c] - mnew[c >> 1]);
        s[i][c] = e;
        rs[c >> 1] += e;
      }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) lrow[r] = lrow[r] * alpha[r] + rs[r];
    if (p.drop_thresh) {   // the normaliser above is of the undropped probabilities; only the P that multiplies V is dropped
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float dm[4];
        flash_drop4(p, (long long)b * p.H + h, rowA, k0 + i * 8 + 2 * t, dm);
#pragma unroll
        for (int c = 0; c < 4; ++c) s[i][c] *= dm[c];
      }
    }
#pragma unroll
    for (int i = 0; i < D / 8; ++i) {
      o[i][0] *= alpha[0]; o[i][1] *= alpha[0]; o[i][2] *= alpha[1]; o[i][3] *= alpha[1];
    }
    // O += P V: the C fragments of S are the A fragments of P
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const uint32_t a0 = pack16<BF16>(s[2 * kk][0], s[2 * kk][1]), a1 = pack16<BF16>(s[2 * kk][2], s[2 * kk][3]);
      const uint32_t a2 = pack16<BF16>(s[2 * kk + 1][0], s[2 * kk + 1][1]), a3 = pack16<BF16>(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
      for (int dp = 0; dp < D / 16; ++dp) {
        uint32_t bf[4];
        load_b_kn<CH>(cV, kk * 16, 2 * dp, lane, bf);
        mma_16816<BF16>(o[2 * dp], a0, a1, a2, a3, bf[0], bf[1]);
        mma_16816<BF16>(o[2 * dp + 1], a0, a1, a2, a3, bf[2], bf[3]);
      }
    }
    __syncthreads();   // every warp is done with this K/V buffer before the load two blocks ahead overwrites it
  }
  // finish: O / l, lse = m ln2 + log l; stage the 16 rows of this warp in its own (dead) rows of the Q tile
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 1);
    lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 2);
  }
  const float inv[2] = {lrow[0] > 0.f ? 1.f / lrow[0] : 0.f, lrow[1] > 0.f ? 1.f / lrow[1] : 0.f};
  if (t == 0 && p.lse) {
    float* lse = p.lse + ((long long)b * p.H + h) * p.S;
    if (rowA < S) lse[rowA] = mrow[0] * kLn2 + logf(lrow[0]);
    if (rowB < S) lse[rowB] = mrow[1] * kLn2 + logf(lrow[1]);
  }
  uint8_t* sq = fsm;
#pragma unroll
  for (int i = 0; i < D / 8; ++i) {
    const int ra = warp * 16 + g, rb = ra + 8;
    *reinterpret_cast<uint32_t*>(sq + tile_off<CH>(ra, i) + t * 4) = pack16<BF16>(o[i][0] * inv[0], o[i][1] * inv[0]);
    *reinterpret_cast<uint32_t*>(sq + tile_off<CH>(rb, i) + t * 4) = pack16<BF16>(o[i][2] * inv[1], o[i][3] * inv[1]);
  }
  __syncwarp();
  uint16_t* go = reinterpret_cast<uint16_t*>(p.out) + row_base * p.H * D + (long long)h * D;
  for (int i = lane; i < 16 * CH; i += 32) {
    const int r = warp * 16 + i / CH, c = i % CH;
    if (q0 + r < S)
      *reinterpret_cast<uint4*>(go + (long long)(q0 + r) * p.H * D + c * 8) = *reinterpret_cast<const uint4*>(sq + tile_off<CH>(r, c));
  }
}

// ------------------------------------------------------------------------------------------------ backward
// delta[b, h, s] = sum_d dO[b, s, h, d] * O[b, s, h, d]: one warp per (token, head)
template <bool BF16, int D>
__global__ void __launch_bounds__(256)
flash_delta_kernel(const uint16_t* __restrict__ o, const uint16_t* __restrict__ dout, float* __restrict__ delta, int B, int S, int H) {
  const long long w = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (w >= (long long)B * S * H) return;
  const long long tok = w / H;
  const int h = (int)(w % H);
  const uint16_t* po = o + w * D;
  const uint16_t* pd = dout + w * D;
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < (D + 63) / 64; ++i) {
    if ((i * 32 + lane) * 2 >= D) break;
    const uint32_t a = *reinterpret_cast<const uint32_t*>(po + (i * 32 + lane) * 2);
    const uint32_t c = *reinterpret_cast<const uint32_t*>(pd + (i * 32 + lane) * 2);
    float2 fa, fc;
    if (BF16) { fa = unpack_bf16x2(a); fc = unpack_bf16x2(c); }
    else { fa = __half22float2(*reinterpret_cast<const __half2*>(&a)); fc = __half22float2(*reinterpret_cast<const __half2*>(&c)); }
    acc += fa.x * fc.x + fa.y * fc.y;
  }
  acc = warp_sum(acc);
  if (lane == 0) delta[((tok / S) * H + h) * S + (tok % S)] = acc;
}

__device__ __forceinline__ void red_add_v2(float* addr, float a, float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

template <bool BF16, int D, bool CAUSAL>
__global__ void __launch_bounds__(256, 1)
flash_bwd_kernel(const FlashParams p) {
  constexpr int CH = D / 8, KS = D / 16, NT = 256, DH = D / 2;   // each warp owns one half of the head dimension
  constexpr uint32_t kTile = 64 * D * 2, kPT = 64 * 64 * 2;
  extern __shared__ __align__(128) uint8_t fsm[];
  const uint32_t sK = smem_u32(fsm), sV = sK + kTile, sQ = sV + kTile, sdO = sQ + 2 * kTile, sP = sdO + 2 * kTile, sdS = sP + kPT;
  float* sL = reinterpret_cast<float*>(fsm + 6 * kTile + 2 * kPT);   // [2][64] lse * log2e
  float* sDl = sL + 128;                                              // [2][64] delta
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int rw = warp & 3, cw = warp >> 2;
  const int kb = blockIdx.x, h = blockIdx.y, b = blockIdx.z, hk = h / (p.H / p.Hkv);
  const int S = p.S, k0 = kb * 64, nqb = (S + 63) / 64;
  const uint16_t* base = reinterpret_cast<const uint16_t*>(p.qkv) + (long long)b * S * p.W;
  const uint16_t* gq = base + (long long)h * D;
  const uint16_t* gk = base + (long long)(p.H + hk) * D;
  const uint16_t* gv = base + (long long)(p.H + p.Hkv + hk) * D;
  const uint16_t* gdo = reinterpret_cast<const uint16_t*>(p.dout) + (long long)b * S * p.H * D + (long long)h * D;
  const float* glse = p.lse + ((long long)b * p.H + h) * S;
  const float* gdl = p.delta + ((long long)b * p.H + h) * S;
  const int qb0 = CAUSAL ? kb : 0;

  auto load_q = [&](int qb, int buf) {
    load_tile64<D, NT>(sQ + buf * kTile, gq, p.W, qb * 64, S, tid);
    load_tile64<D, NT>(sdO + buf * kTile, gdo, (long long)p.H * D, qb * 64, S, tid);
    if (tid < 64) {
      const int r = qb * 64 + tid;
      sL[buf * 64 + tid] = r < S ? glse[r] * kLog2e : 0.f;
      sDl[buf * 64 + tid] = r < S ? gdl[r] : 0.f;
    }
  };
  load_tile64<D, NT>(sK, gk, p.W, k0, S, tid);
  load_tile64<D, NT>(sV, gv, p.W, k0, S, tid);
  load_q(qb0, 0);
  cp_async_commit();

  float dk[DH / 8][4], dv[DH / 8][4];
#pragma unroll
  for (int i = 0; i < DH / 8; ++i)
#pragma unroll
    for (int c = 0; c < 4; ++c) { dk[i][c] = 0.f; dv[i][c] = 0.f; }

  for (int qb = qb0; qb < nqb; ++qb) {
    const int buf = (qb - qb0) & 1;
    const uint32_t cQ = sQ + buf * kTile, cdO = sdO + buf * kTile;
    if (qb + 1 < nqb) {
      load_q(qb + 1, buf ^ 1);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    // S and dP: query rows 16 rw .., keys 32 cw .. (4 n-tiles)
    float s[4][4], dp[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < 4; ++c) { s[i][c] = 0.f; dp[i][c] = 0.f; }
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      uint32_t aq[4], ad[4];
      load_a<CH>(cQ, rw * 16, ks, lane, aq);
      load_a<CH>(cdO, rw * 16, ks, lane, ad);
#pragma unroll
      for (int np = 0; np < 2; ++np) {
        uint32_t bk[4], bv[4];
        load_b_nk<CH>(sK, cw * 32 + np * 16, ks, lane, bk);
        load_b_nk<CH>(sV, cw * 32 + np * 16, ks, lane, bv);
        mma_16816<BF16>(s[2 * np], aq[0], aq[1], aq[2], aq[3], bk[0], bk[1]);
        mma_16816<BF16>(s[2 * np + 1], aq[0], aq[1], aq[2], aq[3], bk[2], bk[3]);
        mma_16816<BF16>(dp[2 * np], ad[0], ad[1], ad[2], ad[3], bv[0], bv[1]);
        mma_16816<BF16>(dp[2 * np + 1], ad[0], ad[1], ad[2], ad[3], bv[2], bv[3]);
      }
    }
    // P = exp2(S c - lse), dS = P (dP - delta) scale -> 16-bit tiles [query][key]
    const int lrA = rw * 16 + g, lrB = lrA + 8;
    const float lA = sL[buf * 64 + lrA], lB = sL[buf * 64 + lrB], dA = sDl[buf * 64 + lrA], dB = sDl[buf * 64 + lrB];
    const int qA = qb * 64 + lrA, qB = qA + 8;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float pv[4], dsv[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int key = k0 + cw * 32 + i * 8 + 2 * t + (c & 1), q = (c & 2) ? qB : qA;
        const bool ok = key < S && q < S && (!CAUSAL || key <= q);
        const float pr = ok ? exp2f(s[i][c] * p.scale_log2 - ((c & 2) ? lB : lA)) : 0.f;
        pv[c] = pr;
        dsv[c] = dp[i][c];
      }
      if (p.drop_thresh) {   // dV takes the dropped P, dS the undropped P times the masked dP
        float dm[4];
        flash_drop4(p, (long long)b * p.H + h, qA, k0 + cw * 32 + i * 8 + 2 * t, dm);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          dsv[c] = pv[c] * (dsv[c] * dm[c] - ((c & 2) ? dB : dA)) * p.scale;
          pv[c] *= dm[c];
        }
      } else {
#pragma unroll
        for (int c = 0; c < 4; ++c) dsv[c] = pv[c] * (dsv[c] - ((c & 2) ? dB : dA)) * p.scale;
      }
      const int chunk = cw * 4 + i;
      uint8_t* pP = fsm + 6 * kTile;
      uint8_t* pS = pP + kPT;
      *reinterpret_cast<uint32_t*>(pP + tile_off<8>(lrA, chunk) + t * 4) = pack16<BF16>(pv[0], pv[1]);
      *reinterpret_cast<uint32_t*>(pP + tile_off<8>(lrB, chunk) + t * 4) = pack16<BF16>(pv[2], pv[3]);
      *reinterpret_cast<uint32_t*>(pS + tile_off<8>(lrA, chunk) + t * 4) = pack16<BF16>(dsv[0], dsv[1]);
      *reinterpret_cast<uint32_t*>(pS + tile_off<8>(lrB, chunk) + t * 4) = pack16<BF16>(dsv[2], dsv[3]);
    }
    __syncthreads();
    // dV[keys 16 rw .., d half cw] += P^T dO;  dK += dS^T Q   (k = the 64 queries of this block)
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      uint32_t ap[4], as[4];
      load_a_t<8>(sP, kk * 16, rw * 2, lane, ap);
      load_a_t<8>(sdS, kk * 16, rw * 2, lane, as);
#pragma unroll
      for (int dp2 = 0; dp2 < DH / 16; ++dp2) {
        uint32_t bo[4], bq[4];
        load_b_kn<CH>(cdO, kk * 16, cw * (CH / 2) + 2 * dp2, lane, bo);
        load_b_kn<CH>(cQ, kk * 16, cw * (CH / 2) + 2 * dp2, lane, bq);
        mma_16816<BF16>(dv[2 * dp2], ap[0], ap[1], ap[2], ap[3], bo[0], bo[1]);
        mma_16816<BF16>(dv[2 * dp2 + 1], ap[0], ap[1], ap[2], ap[3], bo[2], bo[3]);
        mma_16816<BF16>(dk[2 * dp2], as[0], as[1], as[2], as[3], bq[0], bq[1]);
        mma_16816<BF16>(dk[2 * dp2 + 1], as[0], as[1], as[2], as[3], bq[2], bq[3]);
      }
    }
    // dQ[queries 16 rw .., d half cw] = dS K   (k = the 64 keys of this block), accumulated across key blocks in fp32
    float dq[DH / 8][4];
#pragma unroll
    for (int i = 0; i < DH / 8; ++i)
#pragma unroll
      for (int c = 0; c < 4; ++c) dq[i][c] = 0.f;
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      uint32_t as[4];
      load_a<8>(sdS, rw * 16, kk, lane, as);
#pragma unroll
      for (int dp2 = 0; dp2 < DH / 16; ++dp2) {
        uint32_t bk[4];
        load_b_kn<CH>(sK, kk * 16, cw * (CH / 2) + 2 * dp2, lane, bk);
        mma_16816<BF16>(dq[2 * dp2], as[0], as[1], as[2], as[3], bk[0], bk[1]);
        mma_16816<BF16>(dq[2 * dp2 + 1], as[0], as[1], as[2], as[3], bk[2], bk[3]);
      }
    }
    float* gdq = p.dqkv32 + (long long)b * S * p.W + (long long)h * D + cw * DH + 2 * t;
#pragma unroll
    for (int i = 0; i < DH / 8; ++i) {
      if (qA < S) red_add_v2(gdq + (long long)qA * p.W + i * 8, dq[i][0], dq[i][1]);
      if (qB < S) red_add_v2(gdq + (long long)qB * p.W + i * 8, dq[i][2], dq[i][3]);
    }
    __syncthreads();   // P / dS / this Q, dO buffer are free for the next block
  }
  const int keyA = k0 + rw * 16 + g, keyB = keyA + 8;
  float* gdk = p.dqkv32 + (long long)b * S * p.W + (long long)(p.H + hk) * D + cw * DH + 2 * t;
  float* gdv = p.dqkv32 + (long long)b * S * p.W + (long long)(p.H + p.Hkv + hk) * D + cw * DH + 2 * t;
#pragma unroll
  for (int i = 0; i < DH / 8; ++i) {
    if (keyA < S) {
      red_add_v2(gdk + (long long)keyA * p.W + i * 8, dk[i][0], dk[i][1]);
      red_add_v2(gdv + (long long)keyA * p.W + i * 8, dv[i][0], dv[i][1]);
    }
    if (keyB < S) {
      red_add_v2(gdk + (long long)keyB * p.W + i * 8, dk[i][2], dk[i][3]);
      red_add_v2(gdv + (long long)keyB * p.W + i * 8, dv[i][2], dv[i][3]);
    }
  }
}

// Backward with 128-query blocks: same math as flash_bwd_kernel, warp tiles of 32 x 32 (S, dP, dK, dV) and 32 x 64 (dQ) so
// every ldmatrix feeds twice as many MMAs (the 64-query kernel is bound by shared-memory bandwidth, not by the tensor pipe).
template <bool BF16, int D, bool CAUSAL>
__global__ void __launch_bounds__(256, 1)
flash_bwd128_kernel(const FlashParams p) {
  constexpr int CH = D / 8, KS = D / 16, NT = 256;
  constexpr uint32_t kTile = 64 * D * 2, kPT = 128 * 64 * 2;
  extern __shared__ __align__(128) uint8_t fsm[];
  // K | V | Q[2][128] | dO[2][128] | P[128][64] | dS[128][64] | lse[2][128] | delta[2][128]
  const uint32_t sK = smem_u32(fsm), sV = sK + kTile, sQ = sV + kTile, sdO = sQ + 4 * kTile, sP = sdO + 4 * kTile, sdS = sP + kPT;
  uint8_t* pP = fsm + 10 * kTile;
  uint8_t* pS = pP + kPT;
  float* sL = reinterpret_cast<float*>(fsm + 10 * kTile + 2 * kPT);
  float* sDl = sL + 256;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int kb = blockIdx.x, h = blockIdx.y, b = blockIdx.z, hk = h / (p.H / p.Hkv);
  const int S = p.S, k0 = kb * 64, nqb = (S + 127) / 128;
  const uint16_t* base = reinterpret_cast<const uint16_t*>(p.qkv) + (long long)b * S * p.W;
  const uint16_t* gq = base + (long long)h * D;
  const uint16_t* gk = base + (long long)(p.H + hk) * D;
  const uint16_t* gv = base + (long long)(p.H + p.Hkv + hk) * D;
  const uint16_t* gdo = reinterpret_cast<const uint16_t*>(p.dout) + (long long)b * S * p.H * D + (long long)h * D;
  const float* glse = p.lse + ((long long)b * p.H + h) * S;
  const float* gdl = p.delta + ((long long)b * p.H + h) * S;
  const int qb0 = CAUSAL ? (k0 >> 7) : 0;

  auto load_q = [&](int qb, int buf) {
    load_tile64<D, NT>(sQ + buf * 2 * kTile, gq, p.W, qb * 128, S, tid);
    load_tile64<D, NT>(sQ + buf * 2 * kTile + kTile, gq, p.W, qb * 128 + 64, S, tid);
    load_tile64<D, NT>(sdO + buf * 2 * kTile, gdo, (long long)p.H * D, qb * 128, S, tid);
    load_tile64<D, NT>(sdO + buf * 2 * kTile + kTile, gdo, (long long)p.H * D, qb * 128 + 64, S, tid);
    if (tid < 128) {
      const int r = qb * 128 + tid;
      sL[buf * 128 + tid] = r < S ? glse[r] * kLog2e : 0.f;
      sDl[buf * 128 + tid] = r < S ? gdl[r] : 0.f;
    }
  };
  load_tile64<D, NT>(sK, gk, p.W, k0, S, tid);
  load_tile64<D, NT>(sV, gv, p.W, k0, S, tid);
  load_q(qb0, 0);
  cp_async_commit();

  // dK / dV: warp tile = keys 32 kg .. x d columns 32 dq ..
  const int kg = warp & 1, dq4 = warp >> 1;
  float dk[2][4][4], dv[2][4][4];
#pragma unroll
  for (int m = 0; m < 2; ++m)
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < 4; ++c) { dk[m][i][c] = 0.f; dv[m][i][c] = 0.f; }
  const int rq = warp & 3, ck = warp >> 2;   // S / dP: queries 32 rq .., keys 32 ck ..;  dQ: queries 32 rq .., d 64 ck ..

  for (int qb = qb0; qb < nqb; ++qb) {
    const int buf = (qb - qb0) & 1;
    const uint32_t cQ = sQ + buf * 2 * kTile, cdO = sdO + buf * 2 * kTile;   // 128 rows: row r lives in tile r / 64
    if (qb + 1 < nqb) {
      load_q(qb + 1, buf ^ 1);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    {
      float s[2][4][4], dp[2][4][4];
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int c = 0; c < 4; ++c) { s[m][i][c] = 0.f; dp[m][i][c] = 0.f; }
      const uint32_t tq = cQ + (rq >> 1) * kTile, tdo = cdO + (rq >> 1) * kTile;
      const int r0 = (rq & 1) * 32;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        uint32_t aq[2][4], ad[2][4];
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          load_a<CH>(tq, r0 + m * 16, ks, lane, aq[m]);
          load_a<CH>(tdo, r0 + m * 16, ks, lane, ad[m]);
        }
#pragma unroll
        for (int np = 0; np < 2; ++np) {
          uint32_t bk[4], bv[4];
          load_b_nk<CH>(sK, ck * 32 + np * 16, ks, lane, bk);
          load_b_nk<CH>(sV, ck * 32 + np * 16, ks, lane, bv);
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            mma_16816<BF16>(s[m][2 * np], aq[m][0], aq[m][1], aq[m][2], aq[m][3], bk[0], bk[1]);
            mma_16816<BF16>(s[m][2 * np + 1], aq[m][0], aq[m][1], aq[m][2], aq[m][3], bk[2], bk[3]);
            mma_16816<BF16>(dp[m][2 * np], ad[m][0], ad[m][1], ad[m][2], ad[m][3], bv[0], bv[1]);
            mma_16816<BF16>(dp[m][2 * np + 1], ad[m][0], ad[m][1], ad[m][2], ad[m][3], bv[2], bv[3]);
          }
        }
      }
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        const int lrA = rq * 32 + m * 16 + g, lrB = lrA + 8;
        const float lA = sL[buf * 128 + lrA], lB = sL[buf * 128 + lrB], dA = sDl[buf * 128 + lrA], dB = sDl[buf * 128 + lrB];
        const int qA = qb * 128 + lrA, qB = qA + 8;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float pv[4], dsv[4];
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int key = k0 + ck * 32 + i * 8 + 2 * t + (c & 1), q = (c & 2) ? qB : qA;
            const bool ok = key < S && q < S && (!CAUSAL || key <= q);
            const float pr = ok ? exp2f(s[m][i][c] * p.scale_log2 - ((c & 2) ? lB : lA)) : 0.f;
            pv[c] = pr;
            dsv[c] = dp[m][i][c];
          }
          if (p.drop_thresh) {
            float dm[4];
            flash_drop4(p, (long long)b * p.H + h, qA, k0 + ck * 32 + i * 8 + 2 * t, dm);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              dsv[c] = pv[c] * (dsv[c] * dm[c] - ((c & 2) ? dB : dA)) * p.scale;
              pv[c] *= dm[c];
            }
          } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) dsv[c] = pv[c] * (dsv[c] - ((c & 2) ? dB : dA)) * p.scale;
          }
          const int chunk = ck * 4 + i;
          *reinterpret_cast<uint32_t*>(pP + tile_off<8>(lrA, chunk) + t * 4) = pack16<BF16>(pv[0], pv[1]);
          *reinterpret_cast<uint32_t*>(pP + tile_off<8>(lrB, chunk) + t * 4) = pack16<BF16>(pv[2], pv[3]);
          *reinterpret_cast<uint32_t*>(pS + tile_off<8>(lrA, chunk) + t * 4) = pack16<BF16>(dsv[0], dsv[1]);
          *reinterpret_cast<uint32_t*>(pS + tile_off<8>(lrB, chunk) + t * 4) = pack16<BF16>(dsv[2], dsv[3]);
        }
      }
    }
    __syncthreads();
    // dV += P^T dO, dK += dS^T Q over the 128 queries of this block
#pragma unroll
    for (int kk = 0; kk < 8; ++kk) {
      uint32_t ap[2][4], as[2][4];
      const uint32_t tq = cQ + (kk >> 2) * kTile, tdo = cdO + (kk >> 2) * kTile;
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        load_a_t<8>(sP, kk * 16, kg * 4 + m * 2, lane, ap[m]);
        load_a_t<8>(sdS, kk * 16, kg * 4 + m * 2, lane, as[m]);
      }
#pragma unroll
      for (int np = 0; np < 2; ++np) {
        uint32_t bo[4], bq[4];
        load_b_kn<CH>(tdo, (kk & 3) * 16, dq4 * 4 + np * 2, lane, bo);
        load_b_kn<CH>(tq, (kk & 3) * 16, dq4 * 4 + np * 2, lane, bq);
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          mma_16816<BF16>(dv[m][2 * np], ap[m][0], ap[m][1], ap[m][2], ap[m][3], bo[0], bo[1]);
          mma_16816<BF16>(dv[m][2 * np + 1], ap[m][0], ap[m][1], ap[m][2], ap[m][3], bo[2], bo[3]);
          mma_16816<BF16>(dk[m][2 * np], as[m][0], as[m][1], as[m][2], as[m][3], bq[0], bq[1]);
          mma_16816<BF16>(dk[m][2 * np + 1], as[m][0], as[m][1], as[m][2], as[m][3], bq[2], bq[3]);
        }
      }
    }
    // dQ[32 rq .., d half ck] = dS K
    {
      constexpr int NP = D / 32;   // n-tile pairs in this warp's half of the head dimension
      float dq[2][2 * NP][4];
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int i = 0; i < 2 * NP; ++i)
#pragma unroll
          for (int c = 0; c < 4; ++c) dq[m][i][c] = 0.f;
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        uint32_t as[2][4];
        load_a<8>(sdS, rq * 32, kk, lane, as[0]);
        load_a<8>(sdS, rq * 32 + 16, kk, lane, as[1]);
#pragma unroll
        for (int np = 0; np < NP; ++np) {
          uint32_t bk[4];
          load_b_kn<CH>(sK, kk * 16, ck * (CH / 2) + 2 * np, lane, bk);
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            mma_16816<BF16>(dq[m][2 * np], as[m][0], as[m][1], as[m][2], as[m][3], bk[0], bk[1]);
            mma_16816<BF16>(dq[m][2 * np + 1], as[m][0], as[m][1], as[m][2], as[m][3], bk[2], bk[3]);
          }
        }
      }
      float* gdq = p.dqkv32 + (long long)b * S * p.W + (long long)h * D + ck * (D / 2) + 2 * t;
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        const int qA = qb * 128 + rq * 32 + m * 16 + g, qB = qA + 8;
#pragma unroll
        for (int i = 0; i < 2 * NP; ++i) {
          if (qA < S) red_add_v2(gdq + (long long)qA * p.W + i * 8, dq[m][i][0], dq[m][i][1]);
          if (qB < S) red_add_v2(gdq + (long long)qB * p.W + i * 8, dq[m][i][2], dq[m][i][3]);
        }
      }
    }
    __syncthreads();
  }
  float* gdk = p.dqkv32 + (long long)b * S * p.W + (long long)(p.H + hk) * D + dq4 * 32 + 2 * t;
  float* gdv = p.dqkv32 + (long long)b * S * p.W + (long long)(p.H + p.Hkv + hk) * D + dq4 * 32 + 2 * t;
#pragma unroll
  for (int m = 0; m < 2; ++m) {
    const int keyA = k0 + kg * 32 + m * 16 + g, keyB = keyA + 8;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (keyA < S) {
        red_add_v2(gdk + (long long)keyA * p.W + i * 8, dk[m][i][0], dk[m][i][1]);
        red_add_v2(gdv + (long long)keyA * p.W + i * 8, dv[m][i][0], dv[m][i][1]);
      }
      if (keyB < S) {
        red_add_v2(gdk + (long long)keyB * p.W + i * 8, dk[m][i][2], dk[m][i][3]);
        red_add_v2(gdv + (long long)keyB * p.W + i * 8, dv[m][i][2], dv[m][i][3]);
      }
    }
  }
}

// dqkv (16-bit) = dqkv32
template <bool BF16>
__global__ void __launch_bounds__(256)
flash_cast_kernel(const float4* __restrict__ src, uint2* __restrict__ dst, long long n4) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    const float4 v = src[i];
    dst[i] = make_uint2(pack16<BF16>(v.x, v.y), pack16<BF16>(v.z, v.w));
  }
}

template <bool BF16, int D>
static int launch_fwd(const FlashParams& p, int causal, cudaStream_t st) {
  const int smem = 5 * 64 * D * 2;
  const dim3 grid((p.S + 63) / 64, p.H, p.B);
  if (causal) {
    B200_CHECK(cudaFuncSetAttribute(flash_fwd_kernel<BF16, D, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    flash_fwd_kernel<BF16, D, true><<<grid, 128, smem, st>>>(p);
  } else {
    B200_CHECK(cudaFuncSetAttribute(flash_fwd_kernel<BF16, D, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    flash_fwd_kernel<BF16, D, false><<<grid, 128, smem, st>>>(p);
  }
  B200_LAUNCH_CHECK();
  return 0;
}

template <bool BF16>
static int finish_bwd(const FlashParams& p, void* dqkv, cudaStream_t st) {
  const long long n4 = (long long)p.B * p.S * p.W / 4;
  const unsigned cg = (unsigned)((n4 + 255) / 256 < kNumSMs * 8 ? (n4 + 255) / 256 : kNumSMs * 8);
  flash_cast_kernel<BF16><<<cg, 256, 0, st>>>(reinterpret_cast<const float4*>(p.dqkv32), reinterpret_cast<uint2*>(dqkv), n4);
  B200_LAUNCH_CHECK();
  return 0;
}

template <bool BF16, int D>
static int launch_bwd(const FlashParams& p, const void* out, void* dqkv, int causal, cudaStream_t st) {
  const long long rows = (long long)p.B * p.S * p.H;
  flash_delta_kernel<BF16, D><<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(
      reinterpret_cast<const uint16_t*>(out), reinterpret_cast<const uint16_t*>(p.dout), const_cast<float*>(p.delta), p.B, p.S, p.H);
  B200_LAUNCH_CHECK();
  const dim3 grid((p.S + 63) / 64, p.H, p.B);
  // 128-query blocks halve the shared-memory traffic per MMA but measured 6 % slower than the 64-query kernel on B200
  // (0.163 vs 0.151 ms at B 2, S 512, 40 heads; 2.01 vs 1.88 ms at S 4096): fewer, longer CTAs and a coarser causal edge.
  static const bool use128 = getenv("B200_FLASH_BWD128") != nullptr;
  if constexpr (D == 128) {
    if (use128 && p.S > 64) {
      const int smem128 = 10 * 64 * D * 2 + 2 * 128 * 64 * 2 + 4 * 128 * 4;
      if (causal) {
        B200_CHECK(cudaFuncSetAttribute(flash_bwd128_kernel<BF16, D, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem128));
        flash_bwd128_kernel<BF16, D, true><<<grid, 256, smem128, st>>>(p);
      } else {
        B200_CHECK(cudaFuncSetAttribute(flash_bwd128_kernel<BF16, D, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem128));
        flash_bwd128_kernel<BF16, D, false><<<grid, 256, smem128, st>>>(p);
      }
      B200_LAUNCH_CHECK();
      return finish_bwd<BF16>(p, dqkv, st);
    }
  }
  const int smem = 6 * 64 * D * 2 + 2 * 64 * 64 * 2 + 4 * 64 * 4;
  if (causal) {
    B200_CHECK(cudaFuncSetAttribute(flash_bwd_kernel<BF16, D, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    flash_bwd_kernel<BF16, D, true><<<grid, 256, smem, st>>>(p);
  } else {
    B200_CHECK(cudaFuncSetAttribute(flash_bwd_kernel<BF16, D, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    flash_bwd_kernel<BF16, D, false><<<grid, 256, smem, st>>>(p);
  }
  B200_LAUNCH_CHECK();
  return finish_bwd<BF16>(p, dqkv, st);
}

static int flash_check(int B, int S, int H, int Hkv, int D, int dtype) {
  if (B < 1 || S < 1 || H < 1 || Hkv < 1 || H % Hkv) return B200_ERR_BAD_SHAPE;
  if (D != 32 && D != 64 && D != 96 && D != 128) return B200_ERR_UNSUPPORTED;
  if (dtype != B200_DT_BF16 && dtype != B200_DT_FP16) return B200_ERR_UNSUPPORTED;
  if (B > 65535 || H > 65535) return B200_ERR_BAD_SHAPE;
  return 0;
}

}  // namespace b200
using namespace b200;

static void set_dropout(FlashParams& p, float dropout_p, unsigned long long seed) {
  p.seed = seed;
  p.seed_off = b200_graph_seed_offset();
  if (dropout_p <= 0.f) { p.drop_thresh = 0; p.inv_keep = 1.f; return; }
  p.drop_thresh = dropout_p >= 1.f ? 0xFFFFFFFFu : (uint32_t)((double)dropout_p * 4294967296.0);
  p.inv_keep = dropout_p >= 1.f ? 0.f : 1.f / (1.f - dropout_p);
}

B200_API int b200_flash_attn_fwd(const void* qkv, void* out, float* lse, int B, int S, int H, int Hkv, int D, int causal,
                                 float scale, float dropout_p, unsigned long long seed, int dtype, void* stream) {
  if (int rc = flash_check(B, S, H, Hkv, D, dtype)) return rc;
  if ((reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(out)) & 15) return B200_ERR_BAD_ALIGN;
  FlashParams p = {};
  p.qkv = qkv; p.out = out; p.lse = lse; p.B = B; p.S = S; p.H = H; p.Hkv = Hkv; p.D = D;
  p.W = (long long)(H + 2 * Hkv) * D; p.scale = scale; p.scale_log2 = scale * kLog2e;
  set_dropout(p, dropout_p, seed);
  cudaStream_t st = (cudaStream_t)stream;
  const bool bf = dtype == B200_DT_BF16;
  if (D == 128) return bf ? launch_fwd<true, 128>(p, causal, st) : launch_fwd<false, 128>(p, causal, st);
  if (D == 96) return bf ? launch_fwd<true, 96>(p, causal, st) : launch_fwd<false, 96>(p, causal, st);
  if (D == 32) return bf ? launch_fwd<true, 32>(p, causal, st) : launch_fwd<false, 32>(p, causal, st);
  return bf ? launch_fwd<true, 64>(p, causal, st) : launch_fwd<false, 64>(p, causal, st);
}

B200_API int b200_flash_attn_varlen_fwd(const void* qkv, void* out, const int* cu_seqlens, int n_seqs, int max_seqlen, int H,
                                        int Hkv, int D, int causal, float scale, int dtype, void* stream) {
  if (int rc = flash_check(n_seqs, max_seqlen, H, Hkv, D, dtype)) return rc;
  if (!cu_seqlens) return B200_ERR_BAD_SHAPE;
  if ((reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(out)) & 15) return B200_ERR_BAD_ALIGN;
  FlashParams p = {};
  p.qkv = qkv; p.out = out; p.cu = cu_seqlens; p.B = n_seqs; p.S = max_seqlen; p.H = H; p.Hkv = Hkv; p.D = D;
  p.W = (long long)(H + 2 * Hkv) * D; p.scale = scale; p.scale_log2 = scale * kLog2e;
  cudaStream_t st = (cudaStream_t)stream;
  const bool bf = dtype == B200_DT_BF16;
  if (D == 128) return bf ? launch_fwd<true, 128>(p, causal, st) : launch_fwd<false, 128>(p, causal, st);
  if (D == 96) return bf ? launch_fwd<true, 96>(p, causal, st) : launch_fwd<false, 96>(p, causal, st);
  if (D == 32) return bf ? launch_fwd<true, 32>(p, causal, st) : launch_fwd<false, 32>(p, causal, st);
  return bf ? launch_fwd<true, 64>(p, causal, st) : launch_fwd<false, 64>(p, causal, st);
}

B200_API int b200_flash_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                                 float* dqkv32, void* dqkv, int B, int S, int H, int Hkv, int D, int causal, float scale,
                                 float dropout_p, unsigned long long seed, int dtype, void* stream) {
  if (int rc = flash_check(B, S, H, Hkv, D, dtype)) return rc;
  if ((reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(dout) | reinterpret_cast<uintptr_t>(dqkv32) |
       reinterpret_cast<uintptr_t>(dqkv)) & 15) return B200_ERR_BAD_ALIGN;
  FlashParams p = {};
  p.qkv = qkv; p.dout = dout; p.lse = const_cast<float*>(lse); p.delta = delta; p.dqkv32 = dqkv32;
  p.B = B; p.S = S; p.H = H; p.Hkv = Hkv; p.D = D;
  p.W = (long long)(H + 2 * Hkv) * D; p.scale = scale; p.scale_log2 = scale * kLog2e;
  set_dropout(p, dropout_p, seed);
  cudaStream_t st = (cudaStream_t)stream;
  B200_CHECK(cudaMemsetAsync(dqkv32, 0, (size_t)B * S * p.W * sizeof(float), st));
  const bool bf = dtype == B200_DT_BF16;
  if (D == 128) return bf ? launch_bwd<true, 128>(p, out, dqkv, causal, st) : launch_bwd<false, 128>(p, out, dqkv, causal, st);
  if (D == 96) return bf ? launch_bwd<true, 96>(p, out, dqkv, causal, st) : launch_bwd<false, 96>(p, out, dqkv, causal, st);
  if (D == 32) return bf ? launch_bwd<true, 32>(p, out, dqkv, causal, st) : launch_bwd<false, 32>(p, out, dqkv, causal, st);
  return bf ? launch_bwd<true, 64>(p, out, dqkv, causal, st) : launch_bwd<false, 64>(p, out, dqkv, causal, st);
}
