// Micro-probe (not part of the library): how fast does one SM retire tcgen05.mma in SS mode, and
// how fast does TMA fill shared memory, each in isolation and together? Used to find out why the
// GEMM main loop sat at ~50% of the tensor peak (see profiles/gemm_notes.md).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I csrc -o build/umma_probe bench/umma_probe.cu
#include "common.cuh"
#include "tma_host.h"
#include <stdio.h>
#include <stdlib.h>

using namespace b200;

// mode 0: UMMA only (operands = whatever is in smem), iters x 4 k-steps of N = bn, one commit per 4
// mode 1: TMA only (ring of `stages` stages, consumer just frees them)
// mode 2: both, like the GEMM main loop
__global__ void __launch_bounds__(192, 1)
probe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int mode, int bn,
             int stages, int iters, int ksteps, long long* cycles_out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t stage_b = (uint32_t)bn * 128u;
  const uint32_t sA0 = base, sB0 = base + stages * 16384u;
  const uint32_t bar0 = sB0 + stages * stage_b;
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (stages + s); };
  const uint32_t done_bar = bar0 + 8u * (2 * stages);
  const uint32_t slot = bar0 + 8u * (2 * stages + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(done_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) { tmem_alloc(slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(slot));
  const uint32_t idesc = make_idesc_f16(128, bn, true, false, false);
  long long t0 = 0;
  if (warp == 0 && lane == 0 && mode >= 1) {
    int stage = 0; uint32_t phase = 0;
    for (int it = 0; it < iters; ++it) {
      mbar_wait(empty_bar(stage), phase ^ 1u);
      mbar_arrive_expect_tx(full_bar(stage), 16384u + stage_b);
      tma_load_4d(sA0 + stage * 16384u, &tmA, full_bar(stage), (it * 64) % 4096, (blockIdx.x % 8) * 128, 0, 0);
      tma_load_4d(sB0 + stage * stage_b, &tmB, full_bar(stage), (it * 64) % 4096, (blockIdx.x % 16) * bn, 0, 0);
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1 && lane == 0) {
    int stage = 0; uint32_t phase = 0;
    t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (mode >= 1) { mbar_wait(full_bar(stage), phase); tc_fence_after(); }
      if (mode != 1) {
        for (int k = 0; k < ksteps; ++k) {
          const uint64_t ad = make_smem_desc(sA0 + stage * 16384u + (k & 3) * 32, 16, 1024, kSwizzle128B);
          const uint64_t bd = make_smem_desc(sB0 + stage * stage_b + (k & 3) * 32, 16, 1024, kSwizzle128B);
          umma_f16(tmem, ad, bd, idesc, 1u);
        }
        if (mode == 2) umma_commit(empty_bar(stage));
      } else {
        mbar_arrive(empty_bar(stage));
      }
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
    if (mode != 1) {
      umma_commit(done_bar);
      mbar_wait(done_bar, 0);
    }
    cycles_out[blockIdx.x] = clock64() - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main(int argc, char** argv) {
  const int nsm = 148;
  void *A, *B;
  cudaMalloc(&A, 1024ull * 4096 * 2);
  cudaMalloc(&B, 4096ull * 4096 * 2);
  cudaMemset(A, 0, 1024ull * 4096 * 2);
  cudaMemset(B, 0, 4096ull * 4096 * 2);
  long long* cyc;
  cudaMallocManaged(&cyc, nsm * sizeof(long long));
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
  const int iters = 2000;
  for (int bn : {256, 128}) {
    CUtensorMap tmA, tmB;
    uint64_t da[4] = {4096, 1024, 1, 1}, sa[3] = {8192, 0, 0};
    uint32_t ba[4] = {64, 128, 1, 1};
    if (make_tensor_map(&tmA, A, 2, 4, da, sa, ba, CU_TENSOR_MAP_SWIZZLE_128B)) { printf("map A failed\n"); return 1; }
    uint64_t db[4] = {4096, 4096, 1, 1}, sb[3] = {8192, 0, 0};
    uint32_t bb[4] = {64, (uint32_t)bn, 1, 1};
    if (make_tensor_map(&tmB, B, 2, 4, db, sb, bb, CU_TENSOR_MAP_SWIZZLE_128B)) { printf("map B failed\n"); return 1; }
    const int stage_bytes = 16384 + bn * 128;
    for (int stages : {2, 4, 6}) {
      if (stages * stage_bytes > 220000) continue;
      for (int mode = 0; mode < 3; ++mode) {
        for (int ksteps : {4, 8}) {
          if (mode == 1 && ksteps != 4) continue;
          const int smem = stages * stage_bytes + 8 * (2 * stages + 4) + 1024;
          probe_kernel<<<nsm, 192, smem>>>(tmA, tmB, mode, bn, stages, iters, ksteps, cyc);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
          long long mx = 0, mn = 1ll << 60;
          for (int i = 0; i < nsm; ++i) { mx = cyc[i] > mx ? cyc[i] : mx; mn = cyc[i] < mn ? cyc[i] : mn; }
          const char* names[3] = {"umma-only", "tma-only ", "tma+umma "};
          printf("bn=%3d stages=%d %s ksteps/iter=%d : %8.1f cycles/iter (max over SMs; min %.1f)  ideal umma %d\n", bn,
                 stages, names[mode], ksteps, (double)mx / iters, (double)mn / iters, mode == 1 ? 0 : ksteps * bn / 2);
        }
      }
    }
  }
  return 0;
}
