// A whole multi-layer forward pass (several networks, several layers) as ONE launch of the tcgen05 GEMM
// tile: one CTA per output tile of every layer, all resident at once; a tile of layer l + 1 starts its
// weight (B operand) loads at once and its activation (A operand) loads as soon as the row block it
// reads is complete in layer l -- a per-row-block counter in global memory, released by the epilogue
// warps of the producing tiles and acquired by the consuming tile's TMA thread.
//
// Why: at batch 256 a layer is 8-40 tiles of ~3 us of work; as six launches (online and target network,
// three layers) the forward pass took 31 us behind the sample, most of it launch hand-over, prologue and
// the first round trip of every kernel (profiles/r02_learner_timeline_*.txt).  Replaces the same jnp.dot
// calls as tc_gemm.cuh (noisy_mlp.py:62,80) for rlax_rainbow.py:154-217's loss inputs.
//
// Restrictions (the caller falls back to one launch per layer otherwise): every tile of every job fits
// on the device at once (tiles <= SMs: a consumer CTA spins while it waits, so its producers must be
// resident or finished -- CTAs are numbered producers first), at most kChainMaxJobs jobs.
#pragma once
#include "tc_gemm.cuh"

namespace tc {

constexpr int kChainMaxJobs = 6;
constexpr int kChainBN = 128;

struct ChainJob {
  CUtensorMap ta, tb, tout;   // A [M, K] (unused with abits), B [N, K], fp16 output (EPI_F16 + tma_out)
  EpiParams ep;               // M, N, epilogue operands
  int K;
  int epi;                    // EPI_F16 / EPI_F32 / EPI_C51
  int abits;                  // A operand expanded from ep.a_bits by the epilogue warps
  int tile_base;              // first CTA of the job
  int tiles_n;
  int bn;                     // tile width of this job: 128, or 64 (half the weight bytes per CTA)
  int dep;                    // job that produces this job's A operand (same row blocks), or -1
  int dep_arrivals;           // arrivals per row block of `dep` and launch (its tiles_n x epilogue warps)
  int flag_base;              // this job's row-block counters: flags[flag_base + row block]
  int has_consumer;
  // categorical head of the online network: logits are only read for the action taken of the first
  // sel_rows rows (the loss), expected values only for the rows behind them (the double-Q argmax)
  const int* sel_action;      // [sel_rows] or nullptr: store everything
  int sel_rows;
  // Backward tail (dZ_0^T -> dW_0): the B operand of the first job is what the launch waits for, its A
  // operand (weights) is not
  int b_late;                 // request A ahead of the wait and B behind it (default: the other way round)
  int no_early;               // nothing ahead of the wait: the operand requested first comes from this launch's
                              // own dependencies as well
  // The consumer job's tiles of column block 0 also finish the bias gradient from the row sums the
  // producer's epilogue left (EpiParams::row_part): bias_out[m] = bias_alpha * sum_p bias_in[p * ld + m]
  const float* bias_in;
  int bias_parts, bias_ld;
  float* bias_out;
  float bias_alpha;
};

struct ChainParams {
  int n_jobs, total;
  unsigned int* flags;        // monotonic arrival counters (never reset)
  unsigned int* state;        // [0] launches completed, [1] CTAs of the current launch that are done
  unsigned long long* trace;  // debugging (HB_CHAIN_TRACE): 8 globaltimer stamps per CTA, or nullptr
  int compact;                // HB_CHAIN_COMPACT: 0 tc_gemm.cuh's unrolled epilogues, 1 (default) fp32 tiles through
                              // epilogue_f32_staged, 2 fp16 tiles through epilogue_f16_compact as well
  ChainJob job[kChainMaxJobs];
};

__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define HB_CHAIN_STAMP(k) do { if (p.trace != nullptr) p.trace[blockIdx.x * 16 + (k)] = globaltimer_ns(); } while (0)
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// The categorical head's epilogue (tc_gemm.cuh, EPI_C51) with the logits leaving through shared memory:
// a thread owns a row, so direct stores had every lane of a store instruction in a different 5 KB-apart row
// (32 lines per instruction, 16 instructions per action: 6-7 us of the head's tiles).  Staged as a swizzled
// [32 rows][64 floats] block per warp, a row goes out as one 256-byte store instruction.
// `stage`: 8 KB of the operand ring (free once the accumulator is complete).
// x[64] arrives holding the bias values and my_act the row's action (or -1): both are loaded before the
// accumulator is complete -- behind it they were two L2 round trips in every head tile's epilogue.
__device__ __forceinline__ void epilogue_c51_staged(uint32_t tmem_base, int m0, int n0, int ew, int half, int lane,
                                                    const EpiParams& ep, uint8_t* stage, float (&x)[64], int my_act,
                                                    const int* sel_action, int sel_rows, unsigned long long* tr) {
#define HB_EPI_STAMP(k) do { if (tr != nullptr) tr[k] = (unsigned long long)clock64(); } while (0)
  HB_EPI_STAMP(8);
  const uint32_t lane_addr = tmem_base + ((uint32_t)(ew * 32) << 16);
  const int m = m0 + ew * 32 + lane;
  const bool row_ok = m < ep.M;
  const int nb = n0 + half * 64;
  if (nb >= ep.N) return;                            // warp-uniform
  uint32_t v0[32], v1[32];
  tmem_ld32(lane_addr + half * 64, v0);
  tmem_ld32(lane_addr + half * 64 + 32, v1);
  HB_EPI_STAMP(9);
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    x[j] += ep.alpha * __uint_as_float(v0[j]);
    x[32 + j] += ep.alpha * __uint_as_float(v1[j]);
  }
  HB_EPI_STAMP(10);
  // rows whose logits somebody reads (all of them without sel_action)
  const bool mine = sel_action == nullptr || my_act == (nb >> 6);
  const unsigned want = __ballot_sync(0xffffffffu, mine && row_ok);
  if (ep.out32 != nullptr && want != 0u) {
#pragma unroll
    for (int j4 = 0; j4 < 16; ++j4)
      *reinterpret_cast<float4*>(stage + lane * 256 + ((j4 ^ (lane & 15)) << 4)) =
          make_float4(x[4 * j4], x[4 * j4 + 1], x[4 * j4 + 2], x[4 * j4 + 3]);
    __syncwarp();
    HB_EPI_STAMP(11);
    const int rows = ep.M - (m0 + ew * 32);
#pragma unroll 4
    for (int i = 0; i < 16; ++i) {                   // two rows per instruction, 16 bytes per lane
      const int r = 2 * i + (lane >> 4), c4 = lane & 15;
      if (r < rows && ((want >> r) & 1u)) {
        const float4 o = *reinterpret_cast<const float4*>(stage + r * 256 + ((c4 ^ (r & 15)) << 4));
        *reinterpret_cast<float4*>(ep.out32 + (size_t)(m0 + ew * 32 + r) * ep.ld_out + nb + 4 * c4) = o;
      }
    }
  }
  HB_EPI_STAMP(12);
  if (row_ok && ep.q != nullptr && (sel_action == nullptr || m >= sel_rows)) {
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
  HB_EPI_STAMP(13);
#undef HB_EPI_STAMP
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// out32[m, n] = alpha * acc (+ bias) leaving through shared memory: as with the head's logits, a thread owns a
// row, so direct stores put every lane of an instruction in a different row (the first layer's weight
// gradient: 64 KB per tile, 4.6 us of epilogue).  Same staging as epilogue_c51_staged.
__device__ __forceinline__ void epilogue_f32_staged(uint32_t tmem_base, int m0, int n0, int ew, int half, int lane,
                                                    const EpiParams& ep, uint8_t* stage,
                                                    unsigned long long* tr = nullptr) {
#define HB_EPI_STAMP(k) do { if (tr != nullptr) tr[k] = (unsigned long long)clock64(); } while (0)
  HB_EPI_STAMP(8);
  const uint32_t lane_addr = tmem_base + ((uint32_t)(ew * 32) << 16);
  const int nb = n0 + half * 64;
  if (nb >= ep.N) return;                            // warp-uniform
  HB_EPI_STAMP(9);
  // sixteen columns at a time in a loop that is NOT unrolled: a CTA of this kernel runs its epilogue once, so
  // every instruction of it is an instruction-cache miss; as 64 unrolled columns the accumulator took 3100
  // cycles to reach shared memory
#pragma unroll 1
  for (int c = 0; c < 4; ++c) {
    uint32_t v[16];
    tmem_ld16(lane_addr + half * 64 + c * 16, v);
    float x[16];
    load_bias<16>(ep.bias, nb + c * 16, ep.N, x);
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] += ep.alpha * __uint_as_float(v[j]);
#pragma unroll
    for (int j4 = 0; j4 < 4; ++j4)
      *reinterpret_cast<float4*>(stage + lane * 256 + (((4 * c + j4) ^ (lane & 15)) << 4)) =
          make_float4(x[4 * j4], x[4 * j4 + 1], x[4 * j4 + 2], x[4 * j4 + 3]);
  }
  __syncwarp();
  HB_EPI_STAMP(10);
  HB_EPI_STAMP(11);
  const int rows = ep.M - (m0 + ew * 32);
#pragma unroll 4
  for (int i = 0; i < 16; ++i) {                     // two rows per instruction, 16 bytes per lane
    const int r = 2 * i + (lane >> 4), c4 = lane & 15;
    if (r < rows && nb + 4 * c4 < ep.N) {
      const float4 o = *reinterpret_cast<const float4*>(stage + r * 256 + ((c4 ^ (r & 15)) << 4));
      *reinterpret_cast<float4*>(ep.out32 + (size_t)(m0 + ew * 32 + r) * ep.ld_out + nb + 4 * c4) = o;
    }
  }
  HB_EPI_STAMP(12);
  HB_EPI_STAMP(13);
#undef HB_EPI_STAMP
}

// fp16 epilogue (bias, ReLU or ReLU gate, optional row sums) through the tile's tensor store, sixteen columns
// per iteration of a loop that is not unrolled (see epilogue_f32_staged: code size is what these one-shot
// epilogues pay for).  The warp's 32 x 64 block goes to `stage` as the 128-byte-swizzled box of the output
// map.  Requires ep.tma_out; no transposed copy.
__device__ __forceinline__ void epilogue_f16_compact(uint32_t tmem_base, int m0, int n0, int ew, int half, int lane,
                                                     const EpiParams& ep, const CUtensorMap* tmap_out,
                                                     uint8_t* stage) {
  const uint32_t lane_addr = tmem_base + ((uint32_t)(ew * 32) << 16);
  const int m = m0 + ew * 32 + lane;
  const bool row_ok = m < ep.M;
  const int nb = n0 + half * 64;
  if (nb >= ep.N) return;                            // warp-uniform
  float rs = 0.f;
#pragma unroll 1
  for (int c = 0; c < 4; ++c) {
    uint32_t v[16];
    tmem_ld16(lane_addr + half * 64 + c * 16, v);
    const int n16 = nb + c * 16;
    float x[16];
    load_bias<16>(ep.bias, n16, ep.N, x);
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] += ep.alpha * __uint_as_float(v[j]);
    if (ep.relu) {
#pragma unroll
      for (int j = 0; j < 16; ++j) x[j] = fmaxf(x[j], 0.f);
    }
    if (ep.gate != nullptr && row_ok) {
      const uint4* g = reinterpret_cast<const uint4*>(ep.gate + (size_t)m * ep.ld_gate + n16);
#pragma unroll
      for (int j4 = 0; j4 < 2; ++j4) {
        if (n16 + 8 * j4 < ep.N) {
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
    uint32_t pk[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const __half2 h2 = __floats2half2_rn(x[2 * j], x[2 * j + 1]);
      pk[j] = *reinterpret_cast<const uint32_t*>(&h2);
      if (ep.row_part != nullptr) {
        const float2 f = __half22float2(h2);
        if (n16 + 2 * j < ep.N) rs += f.x;
        if (n16 + 2 * j + 1 < ep.N) rs += f.y;
      }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k)                      // 16-byte chunk 2c + k of row `lane`, swizzled
      *reinterpret_cast<uint4*>(stage + lane * 128 + (((2 * c + k) ^ (lane & 7)) << 4)) =
          make_uint4(pk[4 * k], pk[4 * k + 1], pk[4 * k + 2], pk[4 * k + 3]);
  }
  if (ep.row_part != nullptr && row_ok) ep.row_part[(size_t)(nb >> 6) * ep.ld_part + m] = rs;
  fence_proxy_async();
  __syncwarp();
  if (lane == 0) {
    tma_store_2d(tmap_out, stage, nb, m0 + ew * 32);
    tma_store_commit();
  }
}

__global__ void __launch_bounds__(kThreads, 1) mlp_chain_kernel(const __grid_constant__ ChainParams p) {
  constexpr int BN_ = kChainBN;
  constexpr int kSt = stages_for(BN_);
  constexpr int kStB = (BM + BN_) * BK * 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* out_stage = smem + kSt * kStB;
  uint64_t* full_bar = (uint64_t*)(out_stage + kEpiWarps * kOutStage);
  uint64_t* empty_bar = full_bar + kSt;
  uint64_t* acc_full = empty_bar + kSt;
  uint32_t* tmem_slot = (uint32_t*)(acc_full + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int j = 0;
#pragma unroll
  for (int k = 1; k < kChainMaxJobs; ++k)
    if (k < p.n_jobs && (int)blockIdx.x >= p.job[k].tile_base) j = k;
  const ChainJob& job = p.job[j];
  const EpiParams& ep = job.ep;
  const int t = (int)blockIdx.x - job.tile_base;
  const int mblk = t / job.tiles_n;
  const int m0 = mblk * BM, n0 = (t % job.tiles_n) * job.bn;
  const uint32_t stage_bytes = (uint32_t)(BM + job.bn) * BK * 2;   // A + B bytes of one k-block
  const int num_kb = (job.K + BK - 1) / BK;
  const bool abits = job.abits != 0;

  if (warp == 0 && lane == 0) {
    if (!abits) prefetch_tmap(&job.ta);
    prefetch_tmap(&job.tb);
    if (job.epi == EPI_F16 && ep.tma_out) prefetch_tmap(&job.tout);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kSt; ++s) {
      mbar_init(&full_bar[s], abits ? 3 : 1);               // TMA thread (+ the pair of expanding warps)
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&acc_full[0], 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, BN_);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // The weights do not depend on the launch in front of this one (they were written by the previous
  // update): the first ring-full of B tiles is requested before the wait, while the sample kernel runs.
  const int pre = num_kb < kSt ? num_kb : kSt;
  if (warp == 0 && lane == 0 && !job.no_early) {
    for (int kb = 0; kb < pre; ++kb) {
      uint8_t* b_dst = smem + kb * kStB + BM * BK * 2;
      mbar_expect_tx(&full_bar[kb], abits ? (uint32_t)job.bn * BK * 2 : stage_bytes);
      if (job.b_late) tma_load_2d(smem + kb * kStB, &job.ta, &full_bar[kb], kb * BK, m0);
      else tma_load_2d(b_dst, &job.tb, &full_bar[kb], kb * BK, n0);
    }
  }
  // dependents are released only once this grid's own dependencies are complete: what the kernel in front
  // of this one wrote (the sampled actions, rewards, weights) is then visible to the next kernel's
  // prologue too, which starts its own loads there (c51_loss_kernel)
  pdl_wait();
  pdl_launch_dependents();

  if (warp == 0) {
    if (lane == 0) {
      HB_CHAIN_STAMP(0);                                    // dependencies of the launch are complete
      if (job.no_early) {
        for (int kb = 0; kb < pre; ++kb) {
          mbar_expect_tx(&full_bar[kb], abits ? (uint32_t)job.bn * BK * 2 : stage_bytes);
          if (job.b_late) tma_load_2d(smem + kb * kStB, &job.ta, &full_bar[kb], kb * BK, m0);
          else tma_load_2d(smem + kb * kStB + BM * BK * 2, &job.tb, &full_bar[kb], kb * BK, n0);
        }
      }
      if (job.dep >= 0) {
        const unsigned int launches = *reinterpret_cast<volatile const unsigned int*>(p.state);
        const unsigned int target = (launches + 1u) * (unsigned int)job.dep_arrivals;
        const unsigned int* flag = p.flags + p.job[job.dep].flag_base + mblk;
        const long long t0 = clock64();
        while ((int)(ld_acquire_gpu(flag) - target) < 0) {
          if (clock64() - t0 > 4000000000ll) __trap();     // a protocol bug: an error, never a hung GPU
        }
        fence_proxy_async_all();                           // the rows were written by other CTAs; read by TMA
      }
      HB_CHAIN_STAMP(1);                                    // input rows complete
      if (job.b_late) {
        for (int kb = 0; kb < pre; ++kb)
          tma_load_2d(smem + kb * kStB + BM * BK * 2, &job.tb, &full_bar[kb], kb * BK, n0);
      } else if (!abits) {
        for (int kb = 0; kb < pre; ++kb)
          tma_load_2d(smem + kb * kStB, &job.ta, &full_bar[kb], kb * BK, m0);
      }
      for (int kb = pre; kb < num_kb; ++kb) {
        const int s = kb % kSt;
        const uint32_t ph = (kb / kSt) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1);
        uint8_t* a_dst = smem + s * kStB;
        uint8_t* b_dst = a_dst + BM * BK * 2;
        if (abits) {
          mbar_expect_tx(&full_bar[s], (uint32_t)job.bn * BK * 2);
        } else {
          mbar_expect_tx(&full_bar[s], stage_bytes);
          tma_load_2d(a_dst, &job.ta, &full_bar[s], kb * BK, m0);
        }
        tma_load_2d(b_dst, &job.tb, &full_bar[s], kb * BK, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc_f16(BM, job.bn);
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % kSt;
        const uint32_t ph = (kb / kSt) & 1;
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + s * kStB);
        const uint32_t b_addr = a_addr + BM * BK * 2;
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k) {
          const uint64_t adesc = make_smem_desc(a_addr + k * UMMA_K * 2);
          const uint64_t bdesc = make_smem_desc(b_addr + k * UMMA_K * 2);
          umma_f16(tmem_base, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);
      }
      umma_commit(&acc_full[0]);
      HB_CHAIN_STAMP(2);                                    // last MMA issued
    }
  } else if (warp >= 4) {
    const int ew = warp & 3;                         // the TMEM lane quarter this warp may read
    const int half = (warp - 4) >> 2;                // which half of the tile's columns it drains
    if (abits) {
      // A operand straight from the bit-packed rows (tc_gemm.cuh, ABITS) by the eight epilogue warps, which
      // have nothing else to do until the accumulator is complete.  One k-block costs a warp a serial chain
      // (wait for the slot, expand, proxy fence, arrive) of ~0.5 us whatever its share of the rows, so the
      // warps do not share k-blocks: pair g takes k-blocks g, g + 4, ... and four are in the making at once.
      // Thread p of a pair owns rows p and p + 64 (chunk c of row r at r*128 + ((c ^ (r & 7)) << 4)).
      const int g = (warp - 4) >> 1;
      const int pr = ((warp - 4) & 1) * 32 + lane;
      auto load_words = [&](int kb, int rr) -> uint2 {
        const int m = m0 + pr + 64 * rr;
        if (kb < num_kb && m < ep.M && 2 * kb + 1 < ep.a_stride)
          return __ldg(reinterpret_cast<const uint2*>(ep.a_bits + (size_t)m * ep.a_stride) + kb);
        return make_uint2(0u, 0u);
      };
      auto expand = [&](int kb, const uint2 (&w)[2]) {
        const int s = kb % kSt;
        const uint32_t ph = (kb / kSt) & 1;
        const bool st = p.trace != nullptr && warp == 4 && lane == 0 && kb == 0;
        if (st) p.trace[blockIdx.x * 16 + 8] = (unsigned long long)clock64();
        mbar_wait(&empty_bar[s], ph ^ 1);
        uint8_t* a_dst = smem + s * kStB;
        if (st) p.trace[blockIdx.x * 16 + 9] = (unsigned long long)clock64();
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
          const int r = pr + 64 * rr;
          // (a 256-entry table of chunks in shared memory was tried instead of the arithmetic: twice as slow)
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
        if (st) p.trace[blockIdx.x * 16 + 10] = (unsigned long long)clock64();
        fence_proxy_async();                         // generic-proxy writes -> visible to the MMA
        if (st) p.trace[blockIdx.x * 16 + 11] = (unsigned long long)clock64();
        __syncwarp();
        if (lane == 0) mbar_arrive(&full_bar[s]);
        if (st) p.trace[blockIdx.x * 16 + 12] = (unsigned long long)clock64();
      };
      uint2 wv[4][2];                                // the pair's first four k-blocks, requested together
#pragma unroll
      for (int jj = 0; jj < 4; ++jj)
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) wv[jj][rr] = load_words(g + 4 * jj, rr);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj)
        if (g + 4 * jj < num_kb) expand(g + 4 * jj, wv[jj]);
      for (int kb = g + 16; kb < num_kb; kb += 4) {
        uint2 w[2] = {load_words(kb, 0), load_words(kb, 1)};
        expand(kb, w);
      }
    }
    // the tile's bias values pulled into L1 while the MMAs run (the epilogue's __ldg then hits)
    if (ep.bias != nullptr && lane < 16) {
      const int nb = n0 + half * 64 + lane * 4;
      if (nb < ep.N) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(ep.bias + nb));
        asm volatile("" ::"f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w));
      }
    }
    const bool drains = job.bn == BN_ || half == 0;       // a 64-wide tile is the first half's alone
    if (drains) {
    if (job.epi == EPI_F16 && ep.gate != nullptr) {
      // the ReLU gate of this thread's row pulled into L1 while the MMAs run (the epilogue's __ldg then hits)
      const int m = m0 + ew * 32 + lane, nb = n0 + half * 64;
      if (m < ep.M) {
        const uint4* g = reinterpret_cast<const uint4*>(ep.gate + (size_t)m * ep.ld_gate + nb);
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4)
          if (nb + 8 * j4 < ep.N) {
            const uint4 gv = __ldg(g + j4);
            asm volatile("" ::"r"(gv.x), "r"(gv.y), "r"(gv.z), "r"(gv.w));
          }
      }
    }
    float xb[64];
    int my_act = -1;
    if (job.epi == EPI_C51) {
      load_bias<64>(ep.bias, n0 + half * 64, ep.N, xb);
      const int m = m0 + ew * 32 + lane;
      if (job.sel_action != nullptr && m < job.sel_rows && m < ep.M) my_act = job.sel_action[m];
    }
    mbar_wait(&acc_full[0], 0);
    tc_fence_after();
    if (warp == 4 && lane == 0) HB_CHAIN_STAMP(3);         // accumulator complete
    if (warp == 4 && lane == 0 && p.trace != nullptr) p.trace[blockIdx.x * 16 + 6] = (unsigned long long)clock64();
    uint8_t* stage = out_stage + (warp - 4) * kOutStage;
    if (job.epi == EPI_F16 && ep.tma_out && ep.out16_t == nullptr && p.compact >= 2)
      epilogue_f16_compact(tmem_base, m0, n0, ew, half, lane, ep, &job.tout, stage);
    else if (job.epi == EPI_F16)
      epilogue_tile<EPI_F16, BN_>(tmem_base, 0u, m0, n0, ew, half, lane, ep, &job.tout, stage);
    else if (job.epi == EPI_C51)
      epilogue_c51_staged(tmem_base, m0, n0, ew, half, lane, ep, smem + (warp - 4) * 8192, xb, my_act,
                          job.sel_action, job.sel_rows,
                          (p.trace != nullptr && warp == 4 && lane == 0) ? p.trace + blockIdx.x * 16 : nullptr);
    else if (ep.beta == 0.f && (ep.ld_out & 3) == 0 && (ep.N & 3) == 0 && p.compact >= 1)
      epilogue_f32_staged(tmem_base, m0, n0, ew, half, lane, ep, smem + (warp - 4) * 8192,
                          (p.trace != nullptr && warp == 4 && lane == 0) ? p.trace + blockIdx.x * 16 : nullptr);
    else
      epilogue_tile<EPI_F32, BN_>(tmem_base, 0u, m0, n0, ew, half, lane, ep, &job.tout, stage);
    if (job.bias_out != nullptr && n0 == 0 && half == 0) {
      // (what the producer's epilogue warps wrote is visible: the TMA thread acquired their counter before
      // the loads whose arrival, through the MMAs, released this warp)
      const int m = m0 + ew * 32 + lane;
      if (m < ep.M) {
        float sum = 0.f;
        for (int q = 0; q < job.bias_parts; ++q) sum += job.bias_in[(size_t)q * job.bias_ld + m];
        job.bias_out[m] = job.bias_alpha * sum;
      }
    }
    if (warp == 4 && lane == 0) HB_CHAIN_STAMP(4);         // epilogue stores issued
    if (job.epi == EPI_F16 && ep.tma_out && lane == 0) tma_store_wait_all();
    if (warp == 4 && lane == 0) HB_CHAIN_STAMP(5);         // stores complete
    if (warp == 4 && lane == 0 && p.trace != nullptr) p.trace[blockIdx.x * 16 + 7] = (unsigned long long)clock64();
    if (job.has_consumer) {
      // this warp's part of the row block is in global memory: one arrival on the row block's counter
      __syncwarp();
      if (lane == 0) {
        fence_proxy_async_all();
        __threadfence();
        red_release_gpu_add(p.flags + job.flag_base + mblk, 1u);
      }
    }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BN_);
  }
  if (threadIdx.x == 0) {
    // the last CTA of the launch closes it: every CTA has read state[0] by then
    __threadfence();
    const unsigned int done = atomicAdd(p.state + 1, 1u);
    if (done == (unsigned int)p.total - 1u) {
      p.state[1] = 0u;
      __threadfence();
      atomicAdd(p.state, 1u);
    }
  }
}

}  // namespace tc
