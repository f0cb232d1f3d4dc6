// Q-network (NoisyMLP / MLP), policy and learner step on tcgen05 GEMM tiles, and their C ABI.
#include "tc_host.cuh"
#include <new>
#include <vector>

extern "C" {

// Test / microbenchmark entry: D = epilogue(alpha * A . B^T + bias), A [M, K] and B [N, K] fp16
// row-major device matrices.  mode 0: fp16 out (+relu), 1: fp32 out, 2: categorical head
// (out32 = logits [M, N], q = [M, N/64]).
int hb_gemm_f16(const void* d_a, int64_t lda, const void* d_b, int64_t ldb, int M, int N, int K,
                int mode, float alpha, const float* d_bias, int relu, void* d_out, int64_t ld_out,
                void* d_out_t, int64_t ld_out_t, float* d_q, int n_atoms, float z0, float dz,
                float q_scale, void* stream) {
  HB_REQUIRE(d_a && d_b && M >= 1 && N >= 1 && K >= 1, "hb_gemm_f16: bad arguments");
  HB_REQUIRE(N % 8 == 0 && K % 8 == 0, "hb_gemm_f16: N and K must be multiples of 8");
  tc::EpiParams ep{};
  ep.alpha = alpha;
  ep.beta = 0.f;
  ep.bias = d_bias;
  ep.relu = relu;
  ep.ld_out = (int)ld_out;
  ep.ld_out_t = (int)ld_out_t;
  cudaStream_t s = (cudaStream_t)stream;
  const __half* A = (const __half*)d_a;
  const __half* B = (const __half*)d_b;
  switch (mode) {
    case 0:
      ep.out16 = (__half*)d_out;
      ep.out16_t = (__half*)d_out_t;
      return tc::gemm<tc::EPI_F16>(A, lda, B, ldb, M, N, K, ep, s);
    case 1:
      ep.out32 = (float*)d_out;
      return tc::gemm<tc::EPI_F32>(A, lda, B, ldb, M, N, K, ep, s);
    case 2:
      HB_REQUIRE(N % 64 == 0 && n_atoms >= 1 && n_atoms <= 64, "hb_gemm_f16: bad categorical head");
      ep.out32 = (float*)d_out;
      ep.q = d_q;
      ep.ld_q = N / 64;
      ep.n_atoms = n_atoms;
      ep.z0 = z0;
      ep.dz = dz;
      ep.q_scale = q_scale;
      return tc::gemm<tc::EPI_C51>(A, lda, B, ldb, M, N, K, ep, s);
    default:
      hb::set_error("hb_gemm_f16: unknown mode %d", mode);
      return HB_ERR_INVALID;
  }
}

}  // extern "C"
