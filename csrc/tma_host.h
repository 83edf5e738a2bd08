// Host-side construction of TMA tensor maps without linking libcuda: the encode entry point
// is resolved through the runtime (cudaGetDriverEntryPoint) the first time it is needed, so the
// library loads on a CPU-only box (where only the build and the CPU tests run).
#pragma once
#include <stdio.h>

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <mutex>

namespace b200 {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

// dims/strides innermost-first. strides_bytes[i] is the stride of dim i+1 (dim 0 is contiguous).
// elem_bytes: 1 (uint8), 2 (bf16/fp16 — encoded as bf16, the copy is bit-exact) or 4 (fp32).
inline int make_tensor_map(CUtensorMap* map, const void* base, int elem_bytes, int rank, const uint64_t* dims,
                           const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swizzle) {
  PFN_encodeTiled enc = get_encode_tiled();
  if (!enc) return 10003;  // B200_ERR_NO_DRIVER
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    estr[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) {
    uint64_t s = strides_bytes[i];
    // size-1 dims come in with stride 0: give them the dense value so the map stays well-formed
    if (s == 0) s = (i == 0 ? (uint64_t)elem_bytes : gstr[i - 1]) * dims[i];
    if (s % 16) s = (s + 15) / 16 * 16;
    gstr[i] = s;
  }
  CUtensorMapDataType dt = elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                           : elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                             : CU_TENSOR_MAP_DATA_TYPE_UINT8;
  CUresult r = enc(map, dt, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "b200: cuTensorMapEncodeTiled failed (CUresult %d): base %p elem %d rank %d dims [%llu, %llu] stride0 %llu box [%u, %u]\n",
            (int)r, base, elem_bytes, rank, (unsigned long long)gdim[0], (unsigned long long)(rank > 1 ? gdim[1] : 0),
            (unsigned long long)(rank > 1 ? gstr[0] : 0), bx[0], rank > 1 ? bx[1] : 0u);
    return 10004;  // B200_ERR_TENSORMAP
  }
  return 0;
}

}  // namespace b200
