// Cross-GPU signalling over symmetric memory, shared by the collective kernels (comm.cu) and by the compute kernels that
// end in a collective (gemv.cu: row-parallel GEMV + all-reduce + residual + norm in one launch).
// Flags layout per rank: uint32 flags[channel][kMaxRanks]; a rank publishes epoch e on every peer with st.release.sys and
// waits for all peers with ld.acquire.sys. Epochs only grow; every spin is bounded and traps naming the missing rank.
#pragma once
#include "common.cuh"
#include <stdio.h>

namespace b200 {

constexpr int kMaxRanks = 8;
constexpr int kCommThreads = 256;

struct PeerPtrs {
  void* p[kMaxRanks];
};

__device__ __forceinline__ void st_release_sys(uint32_t* addr, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* addr) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}

// flags layout per rank: uint32 flags[channel][kMaxRanks]
__device__ __forceinline__ void signal_all(const PeerPtrs& flags, int channel, int world, int rank, uint32_t epoch) {
  if ((int)threadIdx.x < world) {
    uint32_t* f = reinterpret_cast<uint32_t*>(flags.p[threadIdx.x]) + channel * kMaxRanks + rank;
    st_release_sys(f, epoch);
  }
}
__device__ __forceinline__ void wait_all(const PeerPtrs& flags, int channel, int world, int rank, uint32_t epoch) {
  if ((int)threadIdx.x < world) {
    const uint32_t* f = reinterpret_cast<const uint32_t*>(flags.p[rank]) + channel * kMaxRanks + threadIdx.x;
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
      if (clock64() - t0 > 20000000000ll) {  // ~10 s
        printf("b200 comm: rank %d timed out waiting for rank %d (channel %d, epoch %u)\n", rank, (int)threadIdx.x,
               channel, epoch);
        __trap();
      }
    }
  }
  __syncthreads();
}


}  // namespace b200
