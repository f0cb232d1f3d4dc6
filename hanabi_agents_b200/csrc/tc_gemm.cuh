// tcgen05 / TMA GEMM tile for the Q-network (sm_100a only).
//
//   D[M, N] = A[M, K] . B[N, K]^T      A, B fp16, both K-major (row-major with K contiguous),
//                                      fp32 accumulation in tensor memory.
//
// Persistent kernel, one CTA per SM, tiles of 128 x BN (BN = 128 or 256):
//   warp 0     TMA producer: cp.async.bulk.tensor boxes of 128 (A) / BN (B) rows x 64 fp16 (= one
//              128-byte swizzle span) into a 6 x 32 KB or 4 x 48 KB shared-memory ring that keeps
//              running across tiles
//   warp 1     MMA issuer: one lane issues tcgen05.mma.cta_group::1.kind::f16 (M128 x N BN x K16),
//              four per k-block, into one of TWO TMEM accumulators (2 x BN columns);
//              tcgen05.commit releases the ring slot / hands the accumulator to the epilogue
//   warp 2     allocates / frees tensor memory
//   warps 4-11 epilogue (two warps per TMEM lane quarter, half the columns each) of the previous tile while the next one is being multiplied:
//              tcgen05.ld 32 lanes x 32 columns at a time, fused bias / ReLU / gate / transposed
//              copy / categorical head; fp16 outputs leave through swizzled shared memory and
//              cp.async.bulk.tensor stores, the others with vectorised global stores
// BN = 256 halves the shared-memory fill traffic per flop (A 16 KB + B 32 KB per 128x256x64
// block = 87 flop/B against 64 flop/B at BN = 128): the v1 one-tile-per-CTA kernel sat at 16-22 %
// tensor-pipe activity, bound by CTA set-up and L2 -> SM fill (profiles/r01_gemm_acting_v1_*).
//
// Replaces the jnp.dot calls of NoisyLinear (hanabi_agents/rlax_dqn/noisy_mlp.py:62,80) and their
// jax.grad transposes (rlax_rainbow.py:203-213).
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace tc {

constexpr int BM = 128, BK = 64;                  // CTA tile rows; BK fp16 = 128 B = swizzle span
constexpr int UMMA_K = 16;
constexpr int kThreads = 384;                     // 4 control warps + 8 epilogue warps
// tile width BN_ is 128 or 256 (template parameter): 6 x 32 KB or 4 x 48 KB of operand ring
__host__ __device__ constexpr int stages_for(int bn) { return bn == 256 ? 4 : 6; }
constexpr int kEpiWarps = 8;
constexpr int kOutStage = 32 * 64 * 2;            // per epilogue warp: 32 rows x 64 fp16, TMA-store box
__host__ __device__ constexpr int smem_bytes_for(int bn) {
  return stages_for(bn) * (BM + bn) * BK * 2 + kEpiWarps * kOutStage + 1024 /*align slack*/ +
         256 /*barriers*/;
}

enum Epilogue {
  EPI_F16 = 0,   // out16[m, n] = act(alpha * acc + bias[n]) (* gate(m, n) > 0), optional transposed copy
  EPI_F32 = 1,   // out32[m, n] = alpha * acc + bias[n] (+ beta * out32[m, n])
  EPI_C51 = 2,   // categorical head: columns are [action][64] (51 atoms + pad):
                 //   logits32[m, n] (optional), q[m, a] = mean_j softmax_j(acc + bias) * z_j
};

struct EpiParams {
  int M, N;                  // valid rows / columns of D
  float alpha;
  float beta;                // EPI_F32 only: accumulate into existing out32
  const float* bias;         // [N] or nullptr
  int relu;                  // EPI_F16: apply max(., 0)
  const __half* gate;        // EPI_F16: multiply by (gate[m, n] > 0) (ReLU backward), or nullptr
  int ld_gate;
  __half* out16;             // EPI_F16: [M, ld_out]
  int tma_out;               // EPI_F16: out16 goes out through the TMA-store map (staged in smem)
  __half* out16_t;           // EPI_F16: optional transposed copy [N, ld_out_t]
  float* out32;              // EPI_F32 / EPI_C51 (logits, optional)
  int ld_out, ld_out_t;
  // EPI_C51
  float* q;                  // [M, ld_q] expected value per action
  int ld_q;
  int n_atoms;               // 51
  float z0, dz;              // atoms z_j = z0 + j * dz
  float q_scale;             // 1/n_atoms (rlax_rainbow.py:118 uses mean) or 1 (sum)
  // ABITS kernels: the A operand is bit-packed rows (bit k of row m at word k>>5, position k&31)
  const uint32_t* a_bits;    // [M, a_stride] words
  int a_stride;
  // host side only: launch with programmatic stream serialization, i.e. this kernel's prologue
  // (barrier init, TMEM allocation, tensor-map prefetch) may run while the previous kernel of the
  // stream is still finishing; every global access sits behind griddepcontrol.wait
  int pdl;
  // the B operand does not depend on the kernel this launch waits for (it was complete before that kernel
  // could start): the first ring-full of B tiles is requested in front of griddepcontrol.wait.  Only for
  // operands written before the launch's whole dependency chain began (weights of a previous update); the
  // learner does not use it (no measurable gain).
  int b_early;
  int a_early;               // the same for the A operand (not both)
  int max_ctas;              // host side only: upper bound of the persistent grid (0: one CTA per SM)
  // EPI_F16 through the tensor store only: row_part[(n >> 6) * ld_part + m] = sum of the 64 stored (fp16
  // rounded) values of row m in columns n .. n + 63, in column order -- the pieces of a bias gradient
  float* row_part;
  int ld_part;
};

// Programmatic dependent launch (PDL).  wait: returns once every grid this one depends on has
// completed and its memory is visible (a no-op for a launch without the attribute).  launch_dependents:
// the next kernel of the stream may start its own prologue now.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// ---------------------------------------------------------------- PTX wrappers ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.b32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must become a trap (an error the host sees), never a hung GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000ll) __trap();
  }
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar,
                                            int x, int y) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"((uint64_t)map), "r"(smem_u32(bar)), "r"(x), "r"(y)
      : "memory");
}
// shared -> global tile store (bulk async-group completion)
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* smem_src, int x, int y) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   (uint64_t)map),
               "r"(smem_u32(smem_src)), "r"(x), "r"(y)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// the staging buffer of every committed store has been read (it may be overwritten)
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
// every committed store is complete
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)map) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_dst)),
               "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                         uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// tcgen05.commit: arrives on the mbarrier when all MMAs issued so far by this thread are done
// (implies tcgen05.fence::before_thread_sync).
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (thread = TMEM lane = D row)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
        "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor, K-major, SWIZZLE_128B (cute::UMMA::SmemDescriptor):
//   [13:0] start address >> 4, [29:16] leading byte offset >> 4 (unused for swizzled K-major),
//   [45:32] stride byte offset >> 4 (8 rows x 128 B = 1024 B between 8-row groups),
//   [47:46] version = 1 (Blackwell), [63:61] layout type 2 = SWIZZLE_128B.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// Instruction descriptor (cute::UMMA::InstrDescriptor), kind::f16: D fp32, A/B fp16, both K-major.
__host__ __device__ constexpr uint32_t make_idesc_f16(int m, int n) {
  return (1u << 4)                     // c_format = F32
         | (0u << 7) | (0u << 10)      // a_format = b_format = F16
         | (0u << 15) | (0u << 16)     // a_major = b_major = K
         | ((uint32_t)(n >> 3) << 17)  // n_dim
         | ((uint32_t)(m >> 4) << 24); // m_dim
}

// bias[nb .. nb+W) -> x (zero where there is no bias / beyond N).  N is a multiple of 4 and the
// array is 16-byte aligned, so whole float4 groups are either inside or outside.  The loads carry
// no dependence on one another: the v1 epilogue's per-element predicated loads serialised on the
// load latency and cost 24k cycles per 128 x 256 tile (profiles/r01_gemm_epilogue.md).
template <int W>
__device__ __forceinline__ void load_bias(const float* __restrict__ bias, int nb, int N, float (&x)[W]) {
  float4 b[W / 4];
#pragma unroll
  for (int j = 0; j < W / 4; ++j) {
    b[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (bias != nullptr && nb + 4 * j < N) b[j] = __ldg(reinterpret_cast<const float4*>(bias + nb) + j);
  }
#pragma unroll
  for (int j = 0; j < W / 4; ++j) {
    x[4 * j] = b[j].x; x[4 * j + 1] = b[j].y; x[4 * j + 2] = b[j].z; x[4 * j + 3] = b[j].w;
  }
}

// den = sum_j exp(x_j - max), num = sum_j exp(x_j - max) z_j over the 64 columns of one action (pad columns
// hold -inf).  ex2.approx.ftz instead of exp2f: the arguments are <= 0, a result that flushes to zero adds
// nothing to sums that hold at least one 1.0, and exp2f's range handling was four instructions around every
// MUFU -- with the masks and the 64-column loops 4200 cycles of a head tile's epilogue.  The common case
// n_atoms <= 52 walks 52 columns.
__device__ __forceinline__ float ex2_approx(float v) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
  return r;
}
template <int W>
__device__ __forceinline__ void softmax_mean_w(const float (&x)[64], float z0, float dz, float& den, float& num) {
  float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
  for (int j = 0; j < W; ++j) m4[j & 3] = fmaxf(m4[j & 3], x[j]);
  const float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
  const float kLog2e = 1.4426950408889634f;
  const float mxs = mx * kLog2e;
  float d4[4] = {0.f, 0.f, 0.f, 0.f}, n4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int j = 0; j < W; ++j) {
    const float e = ex2_approx(fmaf(x[j], kLog2e, -mxs));
    d4[j & 3] += e;
    n4[j & 3] = fmaf(e, z0 + (float)j * dz, n4[j & 3]);
  }
  den = (d4[0] + d4[1]) + (d4[2] + d4[3]);
  num = (n4[0] + n4[1]) + (n4[2] + n4[3]);
}
__device__ __forceinline__ void softmax_mean64(const float (&x)[64], int n_atoms, float z0, float dz, float& den,
                                               float& num) {
  if (n_atoms <= 52) softmax_mean_w<52>(x, z0, dz, den, num);
  else softmax_mean_w<64>(x, z0, dz, den, num);
}

// ---------------------------------------------------------------------- epilogue ----
// One epilogue warp: rows m0 + ew*32 + lane of the tile whose accumulator starts at TMEM column
// `acc_col` (thread = TMEM lane = D row, 32 fp32 columns per tcgen05.ld).
template <int EPI, int BN_>
__device__ __forceinline__ void epilogue_tile(uint32_t tmem_base, uint32_t acc_col, int m0, int n0,
                                              int ew, int half, int lane, const EpiParams& ep,
                                              const CUtensorMap* tmap_out, uint8_t* stage) {
  const uint32_t lane_addr = tmem_base + ((uint32_t)(ew * 32) << 16) + acc_col;
  const bool row_ok = (m0 + ew * 32 + lane) < ep.M;
  const int m = m0 + ew * 32 + lane;
  if (EPI == EPI_C51) {
    // BN_/64 actions per tile, 64 columns each (n_atoms real + pad)
#pragma unroll 1
    for (int h = half * (BN_ / 128); h < (half + 1) * (BN_ / 128); ++h) {
      uint32_t v0[32], v1[32];
      tmem_ld32(lane_addr + h * 64, v0);
      tmem_ld32(lane_addr + h * 64 + 32, v1);
      const int nb = n0 + h * 64;
      if (nb >= ep.N) continue;
      float x[64];
      load_bias<64>(ep.bias, nb, ep.N, x);           // vector loads issued together
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        x[j] += ep.alpha * __uint_as_float(v0[j]);
        x[32 + j] += ep.alpha * __uint_as_float(v1[j]);
      }
      if (row_ok && ep.out32 != nullptr) {
        float4* dst = reinterpret_cast<float4*>(ep.out32 + (size_t)m * ep.ld_out + nb);
#pragma unroll
        for (int j = 0; j < 16; ++j) dst[j] = make_float4(x[4 * j], x[4 * j + 1], x[4 * j + 2], x[4 * j + 3]);
      }
      if (row_ok && ep.q != nullptr) {
        // pad columns (j >= n_atoms) become -inf: they add exp(-inf) = 0 to both sums, so the
        // loops below run over all 64 columns without per-element predicates; four partial
        // accumulators break the dependent-add chains
#pragma unroll
        for (int j = 32; j < 64; ++j)
          if (j >= ep.n_atoms) x[j] = -INFINITY;
        if (ep.n_atoms < 32) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j >= ep.n_atoms) x[j] = -INFINITY;
        }
        float den, num;
        softmax_mean64(x, ep.n_atoms, ep.z0, ep.dz, den, num);
        ep.q[(size_t)m * ep.ld_q + (nb >> 6)] = ep.q_scale * num / den;
      }
    }
  } else if (EPI == EPI_F16 && ep.tma_out) {
    // 64 columns at a time: the warp's 32 x 64 fp16 block is laid out in shared memory as the
    // 128-byte-swizzled box of the output tensor map and leaves with ONE bulk tensor store --
    // whole 128-byte lines, asynchronous, clipped at M and N by the map.  (Direct stores had every
    // thread write its own 64-byte pieces of a row, 1 KB apart from its neighbours': the fp16
    // layers ran 20-25 % behind the same tile with a small output.)
#pragma unroll 1
    for (int c = half * (BN_ / 128); c < (half + 1) * (BN_ / 128); ++c) {
      uint32_t v0[32], v1[32];
      tmem_ld32(lane_addr + c * 64, v0);
      tmem_ld32(lane_addr + c * 64 + 32, v1);
      const int nb = n0 + c * 64;
      if (nb >= ep.N) continue;                    // warp-uniform
      float x[64];
      load_bias<64>(ep.bias, nb, ep.N, x);
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        x[j] += ep.alpha * __uint_as_float(v0[j]);
        x[32 + j] += ep.alpha * __uint_as_float(v1[j]);
      }
      if (ep.relu) {
#pragma unroll
        for (int j = 0; j < 64; ++j) x[j] = fmaxf(x[j], 0.f);
      }
      if (ep.gate != nullptr && row_ok) {
        const uint4* g = reinterpret_cast<const uint4*>(ep.gate + (size_t)m * ep.ld_gate + nb);
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          if (nb + 8 * j4 < ep.N) {
            const uint4 gv = __ldg(g + j4);
            const __half2* gh = reinterpret_cast<const __half2*>(&gv);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float2 f = __half22float2(gh[j]);
              if (!(f.x > 0.f)) x[8 * j4 + 2 * j] = 0.f;
              if (!(f.y > 0.f)) x[8 * j4 + 2 * j + 1] = 0.f;
            }
          }
        }
      }
      uint32_t p[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const __half2 h2 = __floats2half2_rn(x[2 * j], x[2 * j + 1]);
        p[j] = *reinterpret_cast<const uint32_t*>(&h2);
      }
      if (ep.row_part != nullptr && row_ok) {
        float rs = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&p[j]));
          if (nb + 2 * j < ep.N) rs += f.x;
          if (nb + 2 * j + 1 < ep.N) rs += f.y;
        }
        ep.row_part[(size_t)(nb >> 6) * ep.ld_part + m] = rs;
      }
      if (lane == 0) tma_store_wait_read();        // the previous block has left the buffer
      __syncwarp();
#pragma unroll
      for (int j4 = 0; j4 < 8; ++j4)               // 16-byte chunk j4 of row `lane`, swizzled
        *reinterpret_cast<uint4*>(stage + lane * 128 + ((j4 ^ (lane & 7)) << 4)) =
            make_uint4(p[4 * j4], p[4 * j4 + 1], p[4 * j4 + 2], p[4 * j4 + 3]);
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        tma_store_2d(tmap_out, stage, nb, m0 + ew * 32);
        tma_store_commit();
      }
      if (ep.out16_t != nullptr && row_ok) {
#pragma unroll
        for (int j = 0; j < 64; ++j)
          if (nb + j < ep.N)
            ep.out16_t[(size_t)(nb + j) * ep.ld_out_t + m] = __float2half_rn(x[j]);
      }
    }
  } else {
#pragma unroll 1
    for (int c = half * (BN_ / 64); c < (half + 1) * (BN_ / 64); ++c) {
      uint32_t v[32];
      tmem_ld32(lane_addr + c * 32, v);
      const int nb = n0 + c * 32;
      if (nb >= ep.N) continue;                    // warp-uniform
      float x[32];
      load_bias<32>(ep.bias, nb, ep.N, x);           // vector loads issued together
#pragma unroll
      for (int j = 0; j < 32; ++j) x[j] += ep.alpha * __uint_as_float(v[j]);
      if (EPI == EPI_F16) {
        if (ep.relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
        }
        if (ep.gate != nullptr && row_ok) {
          const uint4* g = reinterpret_cast<const uint4*>(ep.gate + (size_t)m * ep.ld_gate + nb);
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4) {
            if (nb + 8 * j4 < ep.N) {
              const uint4 gv = __ldg(g + j4);
              const __half2* gh = reinterpret_cast<const __half2*>(&gv);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float2 f = __half22float2(gh[j]);
                if (!(f.x > 0.f)) x[8 * j4 + 2 * j] = 0.f;
                if (!(f.y > 0.f)) x[8 * j4 + 2 * j + 1] = 0.f;
              }
            }
          }
        }
        uint32_t p[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const __half2 h2 = __floats2half2_rn(x[2 * j], x[2 * j + 1]);
          p[j] = *reinterpret_cast<const uint32_t*>(&h2);
        }
        if (row_ok && ep.out16 != nullptr) {
          uint4* dst = reinterpret_cast<uint4*>(ep.out16 + (size_t)m * ep.ld_out + nb);
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4)
            if (nb + 8 * j4 < ep.N) dst[j4] = make_uint4(p[4 * j4], p[4 * j4 + 1], p[4 * j4 + 2], p[4 * j4 + 3]);
        }
        if (ep.out16_t != nullptr && row_ok) {
          // transposed copy: for a fixed column the warp's 32 rows are contiguous
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (nb + j < ep.N)
              ep.out16_t[(size_t)(nb + j) * ep.ld_out_t + m] = __float2half_rn(x[j]);
        }
      } else {  // EPI_F32
        if (row_ok) {
          float4* dst = reinterpret_cast<float4*>(ep.out32 + (size_t)m * ep.ld_out + nb);
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            if (nb + 4 * j4 < ep.N) {
              float4 o = make_float4(x[4 * j4], x[4 * j4 + 1], x[4 * j4 + 2], x[4 * j4 + 3]);
              if (ep.beta != 0.f) {
                const float4 old = dst[j4];
                o.x += ep.beta * old.x; o.y += ep.beta * old.y;
                o.z += ep.beta * old.z; o.w += ep.beta * old.w;
              }
              dst[j4] = o;
            }
          }
        }
      }
    }
  }
}

// ---------------------------------------------------------------------- kernel ----
// Persistent: grid = min(#tiles, #SMs); CTA c computes tiles c, c + grid, ... (n fastest, so the
// CTAs working at the same time share A tiles through L2).  The TMEM accumulator is double
// buffered: the MMA warp fills buffer (i & 1) of its i-th tile while the epilogue warps drain the
// other one.
// ABITS kernels carry more warps behind the epilogue's: further pairs of expanding warps, each taking every
// kAbitsPairs-th k-block.
constexpr int kAbitsPairs = 2;   // pairs of expanding warps in an ABITS kernel (warps 2-3, then 12-13, 14-15)
__host__ __device__ constexpr int threads_for(bool abits) { return abits ? kThreads + 64 * (kAbitsPairs - 1) : kThreads; }

template <int EPI, int BN_, bool ABITS>
__global__ void __launch_bounds__(threads_for(ABITS), 1)
gemm_f16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                const __grid_constant__ CUtensorMap tmap_out, int K, EpiParams ep) {
  constexpr int kSt = stages_for(BN_);
  constexpr int kStB = (BM + BN_) * BK * 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* out_stage = smem + kSt * kStB;            // kEpiWarps x 4 KB, 1024-byte aligned
  uint64_t* full_bar = (uint64_t*)(out_stage + kEpiWarps * kOutStage);
  uint64_t* empty_bar = full_bar + kSt;
  uint64_t* acc_full = empty_bar + kSt;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_slot = (uint32_t*)(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb = (K + BK - 1) / BK;
  const int tiles_n = (ep.N + BN_ - 1) / BN_;
  const int tiles_m = (ep.M + BM - 1) / BM;
  const int total = tiles_m * tiles_n;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_b);
    if (EPI == EPI_F16 && ep.tma_out) prefetch_tmap(&tmap_out);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kSt; ++s) {
      mbar_init(&full_bar[s], ABITS ? 3 : 1);        // TMA thread (+ the two unpacking warps)
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], 8);                   // one arrival per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 2 * BN_);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // everything above touched no global memory; the operands may still be in the making
  pdl_launch_dependents();
  const bool a_first = !ABITS && ep.a_early && !ep.b_early;
  const int early = ((ep.b_early || a_first) && (int)blockIdx.x < total) ? (num_kb < kSt ? num_kb : kSt) : 0;
  if (warp == 0 && lane == 0) {
    const int m0 = ((int)blockIdx.x / tiles_n) * BM, n0 = ((int)blockIdx.x % tiles_n) * BN_;
    for (int kb = 0; kb < early; ++kb) {             // first tile's first k-blocks: one operand now,
      mbar_expect_tx(&full_bar[kb], ABITS ? BN_ * BK * 2 : kStB);   // the other behind the wait
      if (a_first) tma_load_2d(smem + kb * kStB, &tmap_a, &full_bar[kb], kb * BK, m0);
      else tma_load_2d(smem + kb * kStB + BM * BK * 2, &tmap_b, &full_bar[kb], kb * BK, n0);
    }
  }
  pdl_wait();

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;                                    // k-blocks issued so far (ring position)
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const int m0 = (t / tiles_n) * BM, n0 = (t % tiles_n) * BN_;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kSt;
          const uint32_t ph = (it / kSt) & 1;
          uint8_t* a_dst = smem + s * kStB;
          uint8_t* b_dst = a_dst + BM * BK * 2;
          if (it < early) {                          // one operand is on its way already
            if (a_first) tma_load_2d(b_dst, &tmap_b, &full_bar[s], kb * BK, n0);
            else if (!ABITS) tma_load_2d(a_dst, &tmap_a, &full_bar[s], kb * BK, m0);
            continue;
          }
          mbar_wait(&empty_bar[s], ph ^ 1);          // slot free (first pass: returns at once)
          if (ABITS) {
            mbar_expect_tx(&full_bar[s], BN_ * BK * 2);
          } else {
            mbar_expect_tx(&full_bar[s], kStB);
            tma_load_2d(a_dst, &tmap_a, &full_bar[s], kb * BK, m0);
          }
          tma_load_2d(b_dst, &tmap_b, &full_bar[s], kb * BK, n0);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_f16(BM, BN_);
      int it = 0, i = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x, ++i) {
        const int buf = i & 1;
        mbar_wait(&acc_empty[buf], ((i >> 1) & 1) ^ 1);   // epilogue has drained this buffer
        tc_fence_after();
        const uint32_t acc = tmem_base + buf * BN_;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kSt;
          const uint32_t ph = (it / kSt) & 1;
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + s * kStB);
          const uint32_t b_addr = a_addr + BM * BK * 2;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t adesc = make_smem_desc(a_addr + k * UMMA_K * 2);
            const uint64_t bdesc = make_smem_desc(b_addr + k * UMMA_K * 2);
            umma_f16(acc, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);                // frees the slot when these MMAs retire
        }
        umma_commit(&acc_full[buf]);                 // accumulator complete
      }
    }
  } else if (ABITS && (warp == 2 || warp == 3 || warp >= 12)) {
    // A operand straight from the bit-packed observation rows: thread p owns rows p and p + 64 of
    // the tile, keeps their packed words in registers and expands 8 bits -> 8 fp16 (one 16-byte
    // chunk) at a time into the 128-byte-swizzled layout the MMA descriptor expects
    // (chunk c of row r at r*128 + ((c ^ (r & 7)) << 4)).
    // Two pairs of warps, each expanding every other k-block of the ring: one k-block costs a pair a serial
    // chain (load the packed words, wait for the slot, expand, proxy fence, arrive) of ~0.8 us whatever is
    // done to shorten its parts, against 0.27 us of MMA at BN = 256.
    const int pair = warp >= 12 ? 1 + ((warp - 12) >> 1) : 0;
    const int p = (warp >= 12 ? (warp - 12) & 1 : warp - 2) * 32 + lane;
    int it = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
      const int m0 = (t / tiles_n) * BM;
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        if (it % kAbitsPairs != pair) continue;
        const int s = it % kSt;
        const uint32_t ph = (it / kSt) & 1;
        uint2 w[2];
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
          const int m = m0 + p + 64 * rr;
          w[rr] = make_uint2(0u, 0u);
          if (m < ep.M && 2 * kb + 1 < ep.a_stride)
            w[rr] = __ldg(reinterpret_cast<const uint2*>(ep.a_bits + (size_t)m * ep.a_stride) + kb);
        }
        mbar_wait(&empty_bar[s], ph ^ 1);
        uint8_t* a_dst = smem + s * kStB;
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
          const int r = p + 64 * rr;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const uint32_t byte = ((c < 4 ? w[rr].x : w[rr].y) >> (8 * (c & 3))) & 0xffu;
            uint32_t h[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint32_t two = (byte >> (2 * i)) & 3u;
              h[i] = ((two | (two << 15)) & 0x00010001u) * 0x3c00u;   // half(1.0) = 0x3c00
            }
            *reinterpret_cast<uint4*>(a_dst + r * 128 + ((c ^ (r & 7)) << 4)) = make_uint4(h[0], h[1], h[2], h[3]);
          }
        }
        fence_proxy_async();                         // generic-proxy writes -> visible to the MMA
        __syncwarp();
        if (lane == 0) mbar_arrive(&full_bar[s]);
      }
    }
  } else if (warp >= 4 && warp < 12) {
    const int ew = warp & 3;                         // the TMEM lane quarter this warp may read
    const int half = (warp - 4) >> 2;                // which half of the tile's columns it drains
    int i = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x, ++i) {
      const int buf = i & 1;
      const int m0 = (t / tiles_n) * BM, n0 = (t % tiles_n) * BN_;
      mbar_wait(&acc_full[buf], (i >> 1) & 1);
      tc_fence_after();
      epilogue_tile<EPI, BN_>(tmem_base, (uint32_t)(buf * BN_), m0, n0, ew, half, lane, ep, &tmap_out,
                              out_stage + (warp - 4) * kOutStage);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
    }
    if (EPI == EPI_F16 && ep.tma_out && lane == 0) tma_store_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN_);
  }
}

}  // namespace tc
