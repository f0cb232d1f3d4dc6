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

// fp16 matrix [rows, cols] with leading dimension ld (elements), K-major; box = box_rows x 64 cols,
// 128-byte swizzle; out-of-range parts of a box read as zero.
inline int make_tmap(CUtensorMap* map, const __half* ptr, int64_t rows, int64_t cols, int64_t ld,
                     int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) { hb::set_error("cuTensorMapEncodeTiled is not available"); return HB_ERR_CUDA; }
  if (((uintptr_t)ptr & 15) != 0 || (ld * 2) % 16 != 0 || rows < 1 || cols < 1) {
    hb::set_error("make_tmap: unaligned operand (ptr %p ld %lld)", (const void*)ptr, (long long)ld);
    return HB_ERR_INVALID;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (void*)ptr, dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { hb::set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return HB_ERR_CUDA; }
  return HB_OK;
}

inline int sm_count() {
  static int n[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return 148;
  if (n[dev] == 0) n[dev] = hb::num_sms(dev);
  return n[dev];
}

template <int EPI, int BN_, bool ABITS = false>
inline int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tout, int M, int N,
                       int K, const EpiParams& ep, cudaStream_t stream) {
  static bool configured[64] = {false};
  int dev = 0;
  HB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    HB_CUDA(cudaFuncSetAttribute(gemm_f16_kernel<EPI, BN_, ABITS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 smem_bytes_for(BN_)));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  const long long tiles = (long long)((N + BN_ - 1) / BN_) * ((M + BM - 1) / BM);
  int sms = sm_count();
  if (ep.max_ctas > 0 && ep.max_ctas < sms) sms = ep.max_ctas;
  const unsigned grid = (unsigned)(tiles < sms ? tiles : sms);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = ep.pdl ? 1 : 0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(threads_for(ABITS));
  cfg.dynamicSmemBytes = smem_bytes_for(BN_);
  cfg.stream = stream;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  HB_CUDA(cudaLaunchKernelEx(&cfg, gemm_f16_kernel<EPI, BN_, ABITS>, ta, tb, tout, K, ep));
  return HB_OK;
}

// Output map of the fp16 epilogue (box = 32 rows x 64 columns, what one epilogue warp stores at a
// time); falls back to direct stores when the output is not 16-byte aligned.
template <int EPI>
inline int out_map(CUtensorMap* tout, const CUtensorMap& dflt, EpiParams& ep) {
  *tout = dflt;
  ep.tma_out = 0;
  if (EPI != EPI_F16 || ep.out16 == nullptr) return HB_OK;
  if (((uintptr_t)ep.out16 & 15) != 0 || ((long long)ep.ld_out * 2) % 16 != 0) return HB_OK;
  const int rc = make_tmap(tout, ep.out16, ep.M, ep.N, ep.ld_out, 32);
  if (rc) return rc;
  ep.tma_out = 1;
  return HB_OK;
}

// D = A . B^T with freshly encoded maps.  Tile width: 256 when the output is at least that wide.
template <int EPI>
inline int gemm(const __half* A, int64_t lda, const __half* B, int64_t ldb, int M, int N, int K,
                EpiParams ep, cudaStream_t stream) {
  // wide tiles only when there are enough of them to occupy the machine: a small problem is
  // latency bound and finishes sooner spread over twice as many CTAs
  const int bn = (N >= 256 && (long long)((N + 255) / 256) * ((M + BM - 1) / BM) >= 64) ? 256 : 128;
  CUtensorMap ta, tb;
  int rc = make_tmap(&ta, A, M, K, lda, BM);
  if (rc) return rc;
  rc = make_tmap(&tb, B, N, K, ldb, bn);
  if (rc) return rc;
  ep.M = M;
  ep.N = N;
  CUtensorMap tout;
  rc = out_map<EPI>(&tout, tb, ep);
  if (rc) return rc;
  if (bn == 256) return launch_gemm<EPI, 256>(ta, tb, tout, M, N, K, ep, stream);
  return launch_gemm<EPI, 128>(ta, tb, tout, M, N, K, ep, stream);
}

// Same with the A operand given as bit-packed rows [M, a_stride words] (K = 32 * a_stride bits,
// a_stride even): the kernel expands the bits itself, no fp16 copy of A ever exists in HBM.
template <int EPI>
inline int gemm_bits(const uint32_t* A_bits, int a_stride, const __half* B, int64_t ldb, int M, int N,
                     int K, EpiParams ep, cudaStream_t stream) {
  if (a_stride % 2 != 0 || K > 32 * a_stride || ((uintptr_t)A_bits & 7) != 0) {
    hb::set_error("gemm_bits: packed rows must be 8-byte aligned with an even word stride");
    return HB_ERR_INVALID;
  }
  // wide tiles only when there are enough of them to occupy the machine: a small problem is
  // latency bound and finishes sooner spread over twice as many CTAs
  const int bn = (N >= 256 && (long long)((N + 255) / 256) * ((M + BM - 1) / BM) >= 64) ? 256 : 128;
  CUtensorMap tb;
  int rc = make_tmap(&tb, B, N, K, ldb, bn);
  if (rc) return rc;
  ep.M = M;
  ep.N = N;
  ep.a_bits = A_bits;
  ep.a_stride = a_stride;
  CUtensorMap tout;
  rc = out_map<EPI>(&tout, tb, ep);
  if (rc) return rc;
  if (bn == 256) return launch_gemm<EPI, 256, true>(tb, tb, tout, M, N, K, ep, stream);
  return launch_gemm<EPI, 128, true>(tb, tb, tout, M, N, K, ep, stream);
}

}  // namespace tc
