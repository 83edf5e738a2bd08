// K5 — RMSNorm / LayerNorm forward and backward, with the residual add fused in.
//
// One CTA per row; 128-bit loads; statistics in fp32. The forward optionally adds a residual
// stream first and writes the sum back (h = x + res; y = norm(h)), so a transformer block touches
// the hidden state once per norm instead of three times. The backward produces dx (plus the
// gradient flowing into the residual stream) and per-CTA partial dw/db rows in a caller-provided
// scratch buffer that a second kernel reduces — no atomics, bitwise reproducible.
//
// Replaces the ATen kernels behind nn.LayerNorm in GPTLike
// (LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:99-116) and behind HF
// Qwen3RMSNorm in every Fine-Tuning/*.py script (fp32 statistics, eps 1e-6, weight scale).
#include "common.cuh"
#include "../include/b200_ops.h"

namespace b200 {

constexpr int kNormThreads = 256;

template <typename T> struct Vec8 { static constexpr int N = 16 / sizeof(T); };

template <typename T, int N>
__device__ __forceinline__ void load_vec(const T* p, float (&v)[N]) {
  uint4 raw = *reinterpret_cast<const uint4*>(p);
  const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
  for (int i = 0; i < N; ++i) v[i] = to_f32<T>(e[i]);
}
template <typename T, int N>
__device__ __forceinline__ void store_vec(T* p, const float (&v)[N]) {
  uint4 raw;
  T* e = reinterpret_cast<T*>(&raw);
#pragma unroll
  for (int i = 0; i < N; ++i) e[i] = from_f32<T>(v[i]);
  *reinterpret_cast<uint4*>(p) = raw;
}

// ---------------------------------------------------------------------------------------------
// forward:  h = x (+ res);  y = (h - mean) * rstd * w (+ b)      [mean = 0 for RMSNorm]
// x, res, y, h_out: [rows, cols] of T;  w, b: [cols] of WT;  stats: [rows, 2] fp32 (mean, rstd)
// ---------------------------------------------------------------------------------------------
template <typename T, typename WT, bool kLayerNorm>
__global__ void __launch_bounds__(kNormThreads)
norm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ res, const WT* __restrict__ w,
                const WT* __restrict__ b, T* __restrict__ y, T* __restrict__ h_out,
                float* __restrict__ stats, int cols, float eps) {
  constexpr int V = 16 / sizeof(T);
  __shared__ float red[32];
  pdl_launch_dependents();
  pdl_wait();
  const long long row = blockIdx.x;
  const T* xr = x + row * cols;
  const T* rr = res ? res + row * cols : nullptr;
  T* yr = y + row * cols;
  T* hr = h_out ? h_out + row * cols : nullptr;
  const int nvec = cols / V;

  float s1 = 0.f, s2 = 0.f;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    float v[V];
    load_vec<T, V>(xr + i * V, v);
    if (rr) {
      float r[V];
      load_vec<T, V>(rr + i * V, r);
#pragma unroll
      for (int j = 0; j < V; ++j) v[j] += r[j];
      if (hr) {
        // round the sum to T first so forward and backward see the same hidden state
        store_vec<T, V>(hr + i * V, v);
      }
    }
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float t = to_f32<T>(from_f32<T>(v[j]));
      s1 += t;
      s2 += t * t;
    }
  }
  s1 = block_sum(s1, red);
  s2 = block_sum(s2, red);
  float mean = 0.f, var;
  if (kLayerNorm) {
    mean = s1 / cols;
    var = fmaxf(s2 / cols - mean * mean, 0.f);
  } else {
    var = s2 / cols;
  }
  const float rstd = rsqrtf(var + eps);
  if (threadIdx.x == 0 && stats) {
    stats[row * 2] = mean;
    stats[row * 2 + 1] = rstd;
  }
  const T* src = (rr && hr) ? hr : xr;  // re-read (L1/L2 hit) instead of keeping the row in registers
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    float v[V];
    load_vec<T, V>(src + i * V, v);
    if (rr && !hr) {
      float r[V];
      load_vec<T, V>(rr + i * V, r);
#pragma unroll
      for (int j = 0; j < V; ++j) v[j] = to_f32<T>(from_f32<T>(v[j] + r[j]));
    }
    float o[V];
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const int c = i * V + j;
      float t = (v[j] - mean) * rstd * to_f32<WT>(w[c]);
      if (b) t += to_f32<WT>(b[c]);
      o[j] = t;
    }
    store_vec<T, V>(yr + i * V, o);
  }
}

// ---------------------------------------------------------------------------------------------
// backward. Each CTA walks rows blockIdx.x, +gridDim.x, ... and keeps its dw/db partial sums in
// registers (thread t owns column vectors t, t+blockDim, ...), so the partials need one store
// per CTA. dx = rstd * (g - mean(g) - xhat * mean(g * xhat)) with g = dy * w (RMSNorm drops the
// mean(g) term). dres_in (optional) is added to dx: the gradient arriving on the residual path.
// ---------------------------------------------------------------------------------------------
template <typename T, typename WT, bool kLayerNorm, int kMaxVecPerThread>
__global__ void __launch_bounds__(kNormThreads)
norm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ h, const WT* __restrict__ w,
                const float* __restrict__ stats, const T* __restrict__ dres_in, T* __restrict__ dx,
                float* __restrict__ dw_part, float* __restrict__ db_part, int rows, int cols) {
  constexpr int V = 16 / sizeof(T);
  __shared__ float red[32];
  const int nvec = cols / V;
  float dw_acc[kMaxVecPerThread][V];
  float db_acc[kMaxVecPerThread][V];
#pragma unroll
  for (int k = 0; k < kMaxVecPerThread; ++k)
#pragma unroll
    for (int j = 0; j < V; ++j) dw_acc[k][j] = db_acc[k][j] = 0.f;

  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* dyr = dy + row * cols;
    const T* hr = h + row * cols;
    const float mean = stats[row * 2], rstd = stats[row * 2 + 1];
    float sg = 0.f, sgx = 0.f;
#pragma unroll
    for (int k = 0; k < kMaxVecPerThread; ++k) {
      const int i = threadIdx.x + k * kNormThreads;
      if (i < nvec) {
        float g[V], xv[V];
        load_vec<T, V>(dyr + i * V, g);
        load_vec<T, V>(hr + i * V, xv);
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float xh = (xv[j] - mean) * rstd;
          const float gw = g[j] * to_f32<WT>(w[i * V + j]);
          sg += gw;
          sgx += gw * xh;
          if (dw_part) dw_acc[k][j] += g[j] * xh;
          if (db_part) db_acc[k][j] += g[j];
        }
      }
    }
    sg = block_sum(sg, red);
    sgx = block_sum(sgx, red);
    const float mg = kLayerNorm ? sg / cols : 0.f;
    const float mgx = sgx / cols;
#pragma unroll
    for (int k = 0; k < kMaxVecPerThread; ++k) {
      const int i = threadIdx.x + k * kNormThreads;
      if (i < nvec) {
        float g[V], xv[V], o[V];
        load_vec<T, V>(dyr + i * V, g);
        load_vec<T, V>(hr + i * V, xv);
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float xh = (xv[j] - mean) * rstd;
          const float gw = g[j] * to_f32<WT>(w[i * V + j]);
          o[j] = rstd * (gw - mg - xh * mgx);
        }
        if (dres_in) {
          float r[V];
          load_vec<T, V>(dres_in + row * cols + i * V, r);
#pragma unroll
          for (int j = 0; j < V; ++j) o[j] += r[j];
        }
        store_vec<T, V>(dx + row * cols + i * V, o);
      }
    }
  }
  if (dw_part || db_part) {
#pragma unroll
    for (int k = 0; k < kMaxVecPerThread; ++k) {
      const int i = threadIdx.x + k * kNormThreads;
      if (i < nvec) {
#pragma unroll
        for (int j = 0; j < V; ++j) {
          if (dw_part) dw_part[(long long)blockIdx.x * cols + i * V + j] = dw_acc[k][j];
          if (db_part) db_part[(long long)blockIdx.x * cols + i * V + j] = db_acc[k][j];
        }
      }
    }
  }
}

// sum the per-CTA partial rows: out[c] (+)= sum_p part[p, c]
template <typename WT>
__global__ void reduce_partials_kernel(const float* __restrict__ part, WT* __restrict__ out, int nparts, int cols,
                                       int accumulate) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float s = 0.f;
  for (int p = 0; p < nparts; ++p) s += part[(long long)p * cols + c];
  if (accumulate) s += to_f32<WT>(out[c]);
  out[c] = from_f32<WT>(s);
}

template <typename T, typename WT>
static int launch_norm_fwd(int layer_norm, const void* x, const void* res, const void* w, const void* b, void* y,
                           void* h_out, float* stats, int rows, int cols, float eps, cudaStream_t st) {
  if (layer_norm)
    return (int)launch_pdl(norm_fwd_kernel<T, WT, true>, rows, kNormThreads, 0, st, (const T*)x, (const T*)res, (const WT*)w,
                           (const WT*)b, (T*)y, (T*)h_out, stats, cols, eps);
  return (int)launch_pdl(norm_fwd_kernel<T, WT, false>, rows, kNormThreads, 0, st, (const T*)x, (const T*)res, (const WT*)w,
                         (const WT*)b, (T*)y, (T*)h_out, stats, cols, eps);
}

template <typename T, typename WT, int MV>
static int launch_norm_bwd(int layer_norm, const void* dy, const void* h, const void* w, const float* stats,
                           const void* dres, void* dx, float* dwp, float* dbp, int rows, int cols, int grid,
                           cudaStream_t st) {
  if (layer_norm)
    norm_bwd_kernel<T, WT, true, MV><<<grid, kNormThreads, 0, st>>>((const T*)dy, (const T*)h, (const WT*)w, stats,
                                                                   (const T*)dres, (T*)dx, dwp, dbp, rows, cols);
  else
    norm_bwd_kernel<T, WT, false, MV><<<grid, kNormThreads, 0, st>>>((const T*)dy, (const T*)h, (const WT*)w, stats,
                                                                    (const T*)dres, (T*)dx, dwp, dbp, rows, cols);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // namespace b200
using namespace b200;

// dtype: activations; wdtype: weight/bias/grad-of-weight dtype. cols must be a multiple of 8 (bf16/fp16)
// or 4 (fp32) and at most 8192 * (4 / sizeof(T)) ... see the checks below.
B200_API int b200_norm_fwd(int layer_norm, const void* x, const void* res, const void* w, const void* b, void* y,
                           void* h_out, float* stats, int rows, int cols, float eps, int dtype, int wdtype,
                           void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const int V = dtype == B200_DT_FP32 ? 4 : 8;
  if (rows <= 0 || cols <= 0 || cols % V) return B200_ERR_BAD_SHAPE;
  if (dtype == B200_DT_BF16 && wdtype == B200_DT_BF16)
    return launch_norm_fwd<__nv_bfloat16, __nv_bfloat16>(layer_norm, x, res, w, b, y, h_out, stats, rows, cols, eps, st);
  if (dtype == B200_DT_BF16 && wdtype == B200_DT_FP32)
    return launch_norm_fwd<__nv_bfloat16, float>(layer_norm, x, res, w, b, y, h_out, stats, rows, cols, eps, st);
  if (dtype == B200_DT_FP32 && wdtype == B200_DT_FP32)
    return launch_norm_fwd<float, float>(layer_norm, x, res, w, b, y, h_out, stats, rows, cols, eps, st);
  if (dtype == B200_DT_FP16 && wdtype == B200_DT_FP16)
    return launch_norm_fwd<__half, __half>(layer_norm, x, res, w, b, y, h_out, stats, rows, cols, eps, st);
  return B200_ERR_UNSUPPORTED;
}

// Number of CTAs (= partial rows) the backward uses for `rows` rows; scratch must hold
// 2 * b200_norm_bwd_parts(rows) * cols floats when dw/db are requested.
B200_API int b200_norm_bwd_parts(int rows) {
  int g = 2 * kNumSMs;
  return rows < g ? rows : g;
}

B200_API int b200_norm_bwd(int layer_norm, const void* dy, const void* h, const void* w, const float* stats,
                           const void* dres_in, void* dx, void* dw, void* db, float* scratch, int accumulate,
                           int rows, int cols, int dtype, int wdtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const int V = dtype == B200_DT_FP32 ? 4 : 8;
  if (rows <= 0 || cols <= 0 || cols % V) return B200_ERR_BAD_SHAPE;
  const int nvec = cols / V;
  const int mv = ceil_div(nvec, kNormThreads);
  if (mv > 4) return B200_ERR_TOO_LARGE;  // cols <= 8192 (16-bit) / 4096 (fp32)
  if ((dw || db) && !scratch) return B200_ERR_BAD_SHAPE;
  const int grid = b200_norm_bwd_parts(rows);
  float* dwp = dw ? scratch : nullptr;
  float* dbp = db ? scratch + (long long)grid * cols : nullptr;
  int rc = B200_ERR_UNSUPPORTED;
#define B200_NORM_BWD(T, WT)                                                                                      \
  rc = mv <= 1   ? launch_norm_bwd<T, WT, 1>(layer_norm, dy, h, w, stats, dres_in, dx, dwp, dbp, rows, cols, grid, st) \
       : mv <= 2 ? launch_norm_bwd<T, WT, 2>(layer_norm, dy, h, w, stats, dres_in, dx, dwp, dbp, rows, cols, grid, st) \
                 : launch_norm_bwd<T, WT, 4>(layer_norm, dy, h, w, stats, dres_in, dx, dwp, dbp, rows, cols, grid, st)
  if (dtype == B200_DT_BF16 && wdtype == B200_DT_BF16) { B200_NORM_BWD(__nv_bfloat16, __nv_bfloat16); }
  else if (dtype == B200_DT_BF16 && wdtype == B200_DT_FP32) { B200_NORM_BWD(__nv_bfloat16, float); }
  else if (dtype == B200_DT_FP32 && wdtype == B200_DT_FP32) { B200_NORM_BWD(float, float); }
  else if (dtype == B200_DT_FP16 && wdtype == B200_DT_FP16) { B200_NORM_BWD(__half, __half); }
#undef B200_NORM_BWD
  if (rc) return rc;
  const int tb = 256, gb = ceil_div(cols, tb);
  if (dw) {
    if (wdtype == B200_DT_FP32) reduce_partials_kernel<float><<<gb, tb, 0, st>>>(dwp, (float*)dw, grid, cols, accumulate);
    else if (wdtype == B200_DT_BF16) reduce_partials_kernel<__nv_bfloat16><<<gb, tb, 0, st>>>(dwp, (__nv_bfloat16*)dw, grid, cols, accumulate);
    else reduce_partials_kernel<__half><<<gb, tb, 0, st>>>(dwp, (__half*)dw, grid, cols, accumulate);
  }
  if (db) {
    if (wdtype == B200_DT_FP32) reduce_partials_kernel<float><<<gb, tb, 0, st>>>(dbp, (float*)db, grid, cols, accumulate);
    else if (wdtype == B200_DT_BF16) reduce_partials_kernel<__nv_bfloat16><<<gb, tb, 0, st>>>(dbp, (__nv_bfloat16*)db, grid, cols, accumulate);
    else reduce_partials_kernel<__half><<<gb, tb, 0, st>>>(dbp, (__half*)db, grid, cols, accumulate);
  }
  B200_LAUNCH_CHECK();
  return 0;
}

// =================================================================================================
// b200_norm_bwd_noalloc — the same backward, for callers that must not allocate.
//
// Who that is: anything that reaches this library through the C ABI without an allocator it is
// allowed to use at call time — an inference/serving process that handed every free byte of HBM to
// its KV-cache pool at start-up, a plugin host or real-time loop that bans cudaMalloc after
// initialisation because it serialises the whole device, or a small C/C++ harness that simply has
// no caching allocator behind it (the Python side of this repository borrows torch's). The regular
// entry point asks such callers for a `scratch` buffer they have no good way to produce.
//
// This variant keeps the per-CTA dw/db partial rows in a fixed array that lives in the library
// image itself: kNoAllocScratchFloats floats = 8 MiB, present on every GPU that loads
// libb200ops.so from load until the process exits. To stay inside it the kernel runs with at most
// kNoAllocParts CTAs, and `cols` may not exceed kNoAllocMaxCols; anything larger is refused with
// B200_ERR_TOO_LARGE and a message on stderr — nothing is launched and nothing is allocated.
//
// Host code cannot hand the array's name to a kernel (on the host it is only a stand-in symbol),
// so the device address is looked up with cudaGetSymbolAddress, once per device, with that device
// current, the first time the entry point is called there; the result is cached. That first call
// therefore does a little extra driver work: make it before starting any stream capture.
//
// The array is shared by every call and keeps whatever the last call left in it. Nothing has to
// be cleared: each call rewrites every partial row its own reduction reads before reading it.
// The flip side is that two calls whose kernels may overlap on the GPU (different streams, or
// different host threads on the same device) race on the array — serialise them.
//
// No cudaMalloc / cudaFree / allocating library routine is called here, on any path.
// =================================================================================================
#include <mutex>

namespace b200 {
constexpr int kNoAllocMaxCols = 8192;
constexpr int kNoAllocParts = 128;
constexpr long long kNoAllocScratchFloats = 2ll * kNoAllocParts * kNoAllocMaxCols;  // dw rows, then db rows
static __device__ __align__(256) float g_norm_bwd_scratch[kNoAllocScratchFloats];

constexpr int kNoAllocMaxDevices = 64;
static float* g_norm_bwd_scratch_addr[kNoAllocMaxDevices] = {nullptr};
static std::mutex g_norm_bwd_scratch_mu;

// device address of the scratch array on the current device (looked up once per device)
static int norm_bwd_noalloc_scratch(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kNoAllocMaxDevices) return B200_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_norm_bwd_scratch_mu);
  if (g_norm_bwd_scratch_addr[dev] == nullptr) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_norm_bwd_scratch);
    if (e != cudaSuccess) return (int)e;
    if (p == nullptr) return B200_ERR_UNSUPPORTED;
    g_norm_bwd_scratch_addr[dev] = p;
  }
  *out = g_norm_bwd_scratch_addr[dev];
  return 0;
}
}  // namespace b200

B200_API int b200_norm_bwd_noalloc_max_cols(void) { return b200::kNoAllocMaxCols; }

B200_API int b200_norm_bwd_noalloc(int layer_norm, const void* dy, const void* h, const void* w, const float* stats,
                                   const void* dres_in, void* dx, void* dw, void* db, int accumulate, int rows,
                                   int cols, int dtype, int wdtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const int V = dtype == B200_DT_FP32 ? 4 : 8;
  if (rows <= 0 || cols <= 0 || cols % V) return B200_ERR_BAD_SHAPE;
  if (cols > kNoAllocMaxCols) {
    fprintf(stderr,
            "b200_norm_bwd_noalloc: cols = %d exceeds the built-in scratch (max %d columns, %lld bytes); "
            "use b200_norm_bwd with caller-provided scratch instead\n",
            cols, kNoAllocMaxCols, kNoAllocScratchFloats * 4);
    return B200_ERR_TOO_LARGE;
  }
  const int nvec = cols / V;
  const int mv = ceil_div(nvec, kNormThreads);
  if (mv > 4) {
    // fp32 activations: a thread holds 4 columns per vector instead of 8, so the row limit of the kernel is half
    fprintf(stderr,
            "b200_norm_bwd_noalloc: cols = %d exceeds the limit for this activation type (%d columns for fp32, %d for "
            "bf16 / fp16); use b200_norm_bwd with caller-provided scratch instead\n",
            cols, kNoAllocMaxCols / 2, kNoAllocMaxCols);
    return B200_ERR_TOO_LARGE;
  }
  float* scratch = nullptr;
  int rc = norm_bwd_noalloc_scratch(&scratch);
  if (rc) return rc;
  const int grid = rows < kNoAllocParts ? rows : kNoAllocParts;
  float* dwp = dw ? scratch : nullptr;
  float* dbp = db ? scratch + (long long)kNoAllocParts * kNoAllocMaxCols : nullptr;
  rc = B200_ERR_UNSUPPORTED;
#define B200_NORM_BWD_NA(T, WT)                                                                                      \
  rc = mv <= 1   ? launch_norm_bwd<T, WT, 1>(layer_norm, dy, h, w, stats, dres_in, dx, dwp, dbp, rows, cols, grid, st) \
       : mv <= 2 ? launch_norm_bwd<T, WT, 2>(layer_norm, dy, h, w, stats, dres_in, dx, dwp, dbp, rows, cols, grid, st) \
                 : launch_norm_bwd<T, WT, 4>(layer_norm, dy, h, w, stats, dres_in, dx, dwp, dbp, rows, cols, grid, st)
  if (dtype == B200_DT_BF16 && wdtype == B200_DT_BF16) { B200_NORM_BWD_NA(__nv_bfloat16, __nv_bfloat16); }
  else if (dtype == B200_DT_BF16 && wdtype == B200_DT_FP32) { B200_NORM_BWD_NA(__nv_bfloat16, float); }
  else if (dtype == B200_DT_FP32 && wdtype == B200_DT_FP32) { B200_NORM_BWD_NA(float, float); }
  else if (dtype == B200_DT_FP16 && wdtype == B200_DT_FP16) { B200_NORM_BWD_NA(__half, __half); }
#undef B200_NORM_BWD_NA
  if (rc) return rc;
  const int tb = 256, gb = ceil_div(cols, tb);
  if (dw) {
    if (wdtype == B200_DT_FP32) reduce_partials_kernel<float><<<gb, tb, 0, st>>>(dwp, (float*)dw, grid, cols, accumulate);
    else if (wdtype == B200_DT_BF16) reduce_partials_kernel<__nv_bfloat16><<<gb, tb, 0, st>>>(dwp, (__nv_bfloat16*)dw, grid, cols, accumulate);
    else reduce_partials_kernel<__half><<<gb, tb, 0, st>>>(dwp, (__half*)dw, grid, cols, accumulate);
  }
  if (db) {
    if (wdtype == B200_DT_FP32) reduce_partials_kernel<float><<<gb, tb, 0, st>>>(dbp, (float*)db, grid, cols, accumulate);
    else if (wdtype == B200_DT_BF16) reduce_partials_kernel<__nv_bfloat16><<<gb, tb, 0, st>>>(dbp, (__nv_bfloat16*)db, grid, cols, accumulate);
    else reduce_partials_kernel<__half><<<gb, tb, 0, st>>>(dbp, (__half*)db, grid, cols, accumulate);
  }
  B200_LAUNCH_CHECK();
  return 0;
}
