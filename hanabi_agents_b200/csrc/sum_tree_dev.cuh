// Device side of the sum tree shared by sum_tree.cu and the fused learner sample (replay.cu).
//
// Flat 1-indexed heap: heap[1] root, level l = [2^l, 2^(l+1)), leaves [cap, 2 cap).
// getIndex (sum_tree/sum_tree/include/sum_tree.h:46-50, 92-105) is a chain of `depth` dependent
// decisions, each needing the LEFT child of the current node:
//     q < left ? go left : (q -= left, go right)
// On the device a dependent L2 load is ~0.7 us, so 22 levels would be 15 us.  Two things shorten it:
//   * levels 0..10 (heap[0..2047]) are staged in shared memory by the CTA;
//   * below that, the descendants of a node `n` at relative depth j are the CONTIGUOUS block
//     heap[n << j .. (n << j) + 2^j): 8, 16, 32 and 64 bytes for j = 1..4.  All four blocks are fetched
//     at once (their addresses depend only on n) and then four levels are decided locally --
//     one L2 round trip per FOUR levels.  The comparisons and subtractions are the reference's, in the
//     reference's order, on the same values: results are bit-identical.
// The arithmetic uses __fmul_rn / __fsub_rn so that no translation unit can contract
// `quantile * root - left` into a fused multiply-add, whatever its -fmad setting.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace hbst {

constexpr int kTopLevels = 11;                // levels 0..10 staged in smem
constexpr int kTopNodes = 1 << kTopLevels;    // heap[0 .. 2047]

__device__ __forceinline__ float ld_cg(const float* p) { return __ldcg(p); }
__device__ __forceinline__ void st_cg(float* p, float v) { __stcg(p, v); }
// Descendant blocks travel global -> shared memory with cp.async: the copies of one round are issued
// back to back whatever the compiler does with the decisions that follow (with register loads, nvcc and
// ptxas split the vector loads, predicated them on the decisions, or re-used their registers so that
// half of them were issued after the first decision -- the round trips ended up in series again), and
// the decisions then index shared memory instead of walking select chains.
constexpr int kScratchBytesPerThread = 128;   // 8 chunks of 16 bytes: block sizes 8 + 16 + 32 + 64 bytes
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(const void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async8(const void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_addr(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// heap[0 .. min(2048, 2 cap)) -> shared memory; returns the number of staged entries.  Ends with a
// CTA barrier.
__device__ __forceinline__ int stage_top(const float* __restrict__ heap, int64_t cap, float* top) {
  const int staged = (int)(2 * cap < (int64_t)kTopNodes ? 2 * cap : (int64_t)kTopNodes);
  for (int i = threadIdx.x; i < staged; i += blockDim.x) top[i] = ld_cg(heap + i);
  __syncthreads();
  return staged;
}

__device__ __forceinline__ void decide(float& q, float left, int& off) {
  if (q < left) {                 // sum_tree.h:99
    off = 2 * off;
  } else {                        // sum_tree.h:104
    q = __fsub_rn(q, left);
    off = 2 * off + 1;
  }
}

// One level per L2 round trip, nothing but registers: for launches with so many keys that the
// machine is full anyway (latency is hidden by other warps; the scratch of `descend` would cost occupancy).
__device__ __forceinline__ int64_t descend_throughput(const float* __restrict__ heap, const float* top,
                                                      int staged, int depth, float q, float* leaf_value) {
  int64_t node = 1;
  for (int l = 0; l < depth; ++l) {
    const int64_t left = node << 1;
    const float lv = left < staged ? top[left] : ld_cg(heap + left);
    if (q < lv) node = left;
    else { q = __fsub_rn(q, lv); node = left + 1; }
  }
  *leaf_value = node < staged ? top[node] : ld_cg(heap + node);
  return node;
}

// Descends from the root to a leaf.  q = quantile * root on entry (already multiplied).  Returns the
// heap position of the leaf (leaf index = position - cap) and its value.
// scratch: this thread's first 16-byte chunk of a [8 chunks][nthr threads] float4 array in shared
// memory (chunk c of thread t at scratch_base + (c * nthr + t) * 16 bytes).
__device__ __forceinline__ int64_t descend(const float* __restrict__ heap, const float* top, int staged,
                                           int depth, float q, float* leaf_value, float* scratch, int nthr) {
  int64_t node = 1;
  int l = 0;
  for (; l < depth && (node << 1) < staged; ++l) {
    const int64_t left = node << 1;
    const float lv = top[left];
    if (q < lv) node = left;
    else { q = __fsub_rn(q, lv); node = left + 1; }
  }
  float value = node < staged ? top[node] : 0.f;
  const int cstride = nthr * 4;                       // floats between consecutive chunks of one thread
  while (l < depth) {
    // up to four levels per L2 round trip (k is the same for every thread of the launch)
    const int k = depth - l < 4 ? depth - l : 4;
    cp_async8(scratch, heap + (node << 1));                                        // chunk 0: level +1
    if (k >= 2) cp_async16(scratch + cstride, heap + (node << 2));                 // chunk 1: level +2
    if (k >= 3) {                                                                  // chunks 2-3: level +3
      cp_async16(scratch + 2 * cstride, heap + (node << 3));
      cp_async16(scratch + 3 * cstride, heap + (node << 3) + 4);
    }
    if (k >= 4) {                                                                  // chunks 4-7: level +4
#pragma unroll
      for (int c = 0; c < 4; ++c) cp_async16(scratch + (4 + c) * cstride, heap + (node << 4) + 4 * c);
    }
    cp_async_wait_all();
    int off = 0;
    float lv = 0.f, rv = 0.f;
    int before = 0;
#pragma unroll
    for (int j = 1; j <= 4; ++j) {
      if (j <= k) {
        const int e = 2 * off;                        // entry of the left child in the level's block
        const int chunk0 = j == 1 ? 0 : (j == 2 ? 1 : (j == 3 ? 2 : 4));
        const float* p = scratch + (chunk0 + (e >> 2)) * cstride + (e & 3);
        lv = p[0];
        rv = p[1];
        before = off;
        decide(q, lv, off);
      }
    }
    value = off == 2 * before ? lv : rv;              // the node this round ends on
    node = (node << k) + off;
    l += k;
  }
  *leaf_value = value;
  return node;
}

}  // namespace hbst
