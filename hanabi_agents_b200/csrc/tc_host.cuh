// Host side of the tcgen05 GEMM tile: tensor-map encoding and launch.
#pragma once
#include "common.cuh"
#include "tc_gemm.cuh"

namespace tc {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda).
inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// fp16 matrix [rows, cols] with leading dimension ld (elements), K-major; box = 128 rows x 64 cols,
// 128-byte swizzle; out-of-range parts of a box read as zero.
inline int make_tmap(CUtensorMap* map, const __half* ptr, int64_t rows, int64_t cols, int64_t ld) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) { hb::set_error("cuTensorMapEncodeTiled is not available"); return HB_ERR_CUDA; }
  if (((uintptr_t)ptr & 15) != 0 || (ld * 2) % 16 != 0 || rows < 1 || cols < 1) {
    hb::set_error("make_tmap: unaligned operand (ptr %p ld %lld)", (const void*)ptr, (long long)ld);
    return HB_ERR_INVALID;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)BM};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (void*)ptr, dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { hb::set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return HB_ERR_CUDA; }
  return HB_OK;
}

template <int EPI>
inline int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, int M, int N, int K,
                       const EpiParams& ep, cudaStream_t stream) {
  static bool configured[64] = {false};
  int dev = 0;
  HB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    HB_CUDA(cudaFuncSetAttribute(gemm_f16_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 kSmemBytes));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  dim3 grid((unsigned)((N + BN - 1) / BN), (unsigned)((M + BM - 1) / BM));
  gemm_f16_kernel<EPI><<<grid, kThreads, kSmemBytes, stream>>>(ta, tb, K, ep);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

// D = A . B^T with freshly encoded maps (acting path: row count changes call to call).
template <int EPI>
inline int gemm(const __half* A, int64_t lda, const __half* B, int64_t ldb, int M, int N, int K,
                EpiParams ep, cudaStream_t stream) {
  CUtensorMap ta, tb;
  int rc = make_tmap(&ta, A, M, K, lda);
  if (rc) return rc;
  rc = make_tmap(&tb, B, N, K, ldb);
  if (rc) return rc;
  ep.M = M;
  ep.N = N;
  return launch_gemm<EPI>(ta, tb, M, N, K, ep, stream);
}

}  // namespace tc
