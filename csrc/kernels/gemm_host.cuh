// Pieces shared by the tcgen05 GEMM kernels (gemm_tcgen05.cu, convbn_tcgen05.cu): the UMMA
// shared-memory descriptor, the driver entry point for tensor-map encoding and 2-D / N-d tensor-map
// builders (128-byte swizzle throughout), and the cached SM count.
#pragma once

#include <cuda.h>

#include <mutex>

#include "common.cuh"
#include "../../include/cloud_b200_kernels.h"

namespace cb2 {

__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3FFFu);
  d |= static_cast<uint64_t>(lbo & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>(sbo & 0x3FFFu) << 32;
  d |= 1ull << 46;  // descriptor version: Blackwell
  d |= 2ull << 61;  // SWIZZLE_128B
  return d;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  // The driver call wants a current context on the calling thread.  A thread that has only ever
  // been handed device pointers (autograd's backward thread running the first backward of the
  // process) has none yet: a runtime call binds the primary context (seen as error 201).
  static thread_local bool bound = false;
  if (!bound) {
    cudaFree(nullptr);
    bound = true;
  }
  return fn;
}

// 2-D tensor map over a row-major [outer, inner] view with row pitch `ld` elements.
inline int make_tmap(CUtensorMap* tm, const void* base, uint64_t inner, uint64_t outer, uint64_t ld,
                     uint32_t box_inner, uint32_t box_outer, bool f32 = false) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_last_error("cuTensorMapEncodeTiled entry point not available");
    return CB2_ERR_CUDA;
  }
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld * (f32 ? 4 : 2)};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  const_cast<void*>(base), dims, strides,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled failed (%d): inner=%llu outer=%llu ld=%llu box=%ux%u",
                   static_cast<int>(r), (unsigned long long)inner, (unsigned long long)outer,
                   (unsigned long long)ld, box_inner, box_outer);
    return CB2_ERR_CUDA;
  }
  return 0;
}

// N-d bf16 / fp32 tensor map with explicit byte strides (dims[0] is the contiguous one).
inline int make_tmap_nd(CUtensorMap* tm, const void* base, int rank, const cuuint64_t* dims,
                        const cuuint64_t* strides_bytes /*[rank-1]*/, const cuuint32_t* box,
                        bool f32 = false) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_last_error("cuTensorMapEncodeTiled entry point not available");
    return CB2_ERR_CUDA;
  }
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(tm, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank,
                  const_cast<void*>(base), dims, strides_bytes, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled (rank %d) failed (%d)", rank, static_cast<int>(r));
    return CB2_ERR_CUDA;
  }
  return 0;
}

inline int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!cached[dev]) cudaDeviceGetAttribute(&cached[dev], cudaDevAttrMultiProcessorCount, dev);
  return cached[dev];
}


}  // namespace cb2
