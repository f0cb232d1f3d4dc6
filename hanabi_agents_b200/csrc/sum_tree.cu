// Batched prioritized-replay sum tree for sm_100a.
//
// Replaces the pointer-chasing SumTree<float> of the reference
// (sum_tree/sum_tree/include/sum_tree.h:22-130) with a flat 1-indexed heap in
// HBM (it is 32 MiB at capacity 2^22, so it lives in the 126 MB L2):
//   heap[1] = root, level l = [2^l, 2^(l+1)), leaves = [cap, 2*cap).
//
// Arithmetic contract (bit-exact against the reference run with one thread):
//   * get_index:  q = quantile * root (one fp32 multiply), then per level
//     `q < left ? go left : (q -= left, go right)`          sum_tree.h:46-50,92-105
//   * update:     diff = value - leaf; every node on the leaf..root path does
//     `node += diff` in fp32, elements applied in batch order.  Internal nodes
//     are never recomputed from their children.               sum_tree.h:32-35,80-90
// The update kernel keeps that order without serialising the batch: one warp
// per tree level, 32 batch elements at a time; elements that hit the same node
// are grouped with __match_any_sync and their diffs are added in lane (= batch)
// order by the group's first lane, so duplicate leaves/ancestors are resolved
// deterministically (the reference races on them under OpenMP).
//
// This file is compiled with -fmad=false: no multiply-add contraction may
// change the reference's rounding.
#include "common.cuh"
#include "sum_tree_dev.cuh"
#include <math.h>
#include <new>

struct hb_sumtree {
  int device = 0;
  int depth = 0;
  int64_t cap = 0;       // 2^depth leaves
  float* heap = nullptr; // 2*cap floats
  int* err = nullptr;    // sticky out-of-range flag (device)
  float* scratch = nullptr;  // diffs for the range update
  int64_t scratch_n = 0;
  // staging for the *_host entry points
  void* stage = nullptr;
  size_t stage_bytes = 0;
};

namespace {

using hbst::kTopNodes;
using hbst::ld_cg;
using hbst::st_cg;
constexpr int kDescendThreads = 256;
constexpr int64_t kLatencyKeys = 16384;       // up to this many keys per launch: the latency-optimised descent
constexpr int kPass = 1024;                   // batch elements per update pass

// ---- stratified prefix-sum descent (ST-2) + leaf gather (ST-3) -------------
// hbst::descend: the top 11 levels from shared memory, then four levels per L2 round trip
// (sum_tree_dev.cuh); the leaf value comes out of the last fetched block, so ST-3 costs no load.
// kLatency: few keys (a learner batch): four levels per round trip through shared-memory scratch;
// otherwise (thousands of keys in flight) the plain one-level-per-load walk, which keeps more warps resident.
template <bool kWithValues, bool kLatency>
__global__ void __launch_bounds__(kDescendThreads)
st_descend_kernel(const float* __restrict__ heap, int depth, int64_t cap,
                  const float* __restrict__ q_in, int64_t* __restrict__ idx_out,
                  float* __restrict__ val_out, float* __restrict__ total_out, int64_t n) {
  __shared__ float top[kTopNodes];
  __shared__ float4 scratch[kLatency ? 8 * kDescendThreads : 1];
  const int64_t k = (int64_t)blockIdx.x * kDescendThreads + threadIdx.x;
  const float quant = k < n ? q_in[k] : 0.f;      // in flight while the top levels are staged
  const int staged = hbst::stage_top(heap, cap, top);
  const float root = top[1];
  if (kWithValues && total_out != nullptr && k == 0) *total_out = root;
  if (k >= n) return;
  const float q = __fmul_rn(quant, root);         // getIndex: quantile * root_->value   sum_tree.h:47
  float leaf_value;
  const int64_t node = kLatency
      ? hbst::descend(heap, top, staged, depth, q, &leaf_value, reinterpret_cast<float*>(scratch + threadIdx.x),
                      kDescendThreads)
      : hbst::descend_throughput(heap, top, staged, depth, q, &leaf_value);
  idx_out[k] = node - cap;
  if (kWithValues) val_out[k] = leaf_value;
}

__global__ void st_gather_kernel(const float* __restrict__ heap, int64_t cap,
                                 const int64_t* __restrict__ idx, float* __restrict__ out,
                                 int64_t n, int* err) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int64_t i = idx[k];
  if (i < 0 || i >= cap) {  // leaves_.at(idx) throws   sum_tree.h:62
    *err = 1;
    out[k] = 0.0f;
    return;
  }
  out[k] = ld_cg(heap + cap + i);
}

__global__ void st_total_kernel(const float* __restrict__ heap, float* out) { *out = ld_cg(heap + 1); }

// ---- ordered scatter-update (ST-4), general batch ---------------------------
// One CTA.  Warp 0 first walks the batch at the leaf level (it needs the old
// leaf to form diff), publishing diffs in shared memory; then warp w owns tree
// level(s) depth-1-w, depth-1-w-nwarps, ... and replays the diffs there.
__global__ void __launch_bounds__(1024)
st_update_kernel(float* __restrict__ heap, int depth, int64_t cap,
                 const int64_t* __restrict__ idx, const float* __restrict__ val, int64_t n,
                 int* err) {
  __shared__ float s_diff[kPass];
  __shared__ int64_t s_leaf[kPass];  // heap position of the leaf, or -1
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int nwarps = blockDim.x >> 5;
  const unsigned full = 0xffffffffu;

  for (int64_t base = 0; base < n; base += kPass) {
    const int pass_n = (int)min((int64_t)kPass, n - base);
    // ---- fast path: non-decreasing, in-range indices (what stratified sampling produces).
    // Elements that share a node are then a contiguous run of the batch, so one thread per
    // (level, run) adds the run's diffs in batch order -- the same per-node order of additions
    // as the sequential reference -- and all runs of all levels proceed in parallel.
    {
      bool ok = true;
      for (int b = threadIdx.x; b < pass_n; b += blockDim.x) {
        const int64_t li = idx[base + b];
        const bool in_range = li >= 0 && li < cap;
        s_leaf[b] = in_range ? cap + li : -1;
        if (!in_range || (b > 0 && idx[base + b - 1] > li)) ok = false;
      }
      if (__syncthreads_and(ok)) {
        for (int b = threadIdx.x; b < pass_n; b += blockDim.x) {
          const int64_t pos = s_leaf[b];
          if (b == 0 || s_leaf[b - 1] != pos) {
            float L = ld_cg(heap + pos);
            for (int j = b; j < pass_n && s_leaf[j] == pos; ++j) {
              const float d = val[base + j] - L;   // updateValue: value - leaf   sum_tree.h:33
              L = L + d;                           // updateValue_: node->value += diff   :84
              s_diff[j] = d;
            }
            st_cg(heap + pos, L);
          }
        }
        __syncthreads();
        // (level, element) items, eight per thread at a time: first the loads of every run head among
        // them (independent of one another -- a node is the head of exactly one run -- so the L2 round
        // trips overlap instead of queueing behind each other's stores), then the ordered adds.
        const int items = depth * pass_n;
        for (int it0 = threadIdx.x; it0 < items; it0 += 8 * blockDim.x) {
          float sacc[8];
          int64_t nodes[8];
          int shifts[8], firsts[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int it = it0 + u * blockDim.x;
            nodes[u] = -1;
            if (it < items) {
              const int l = it / pass_n, b = it - l * pass_n;    // level l: shift = depth - l
              const int shift = depth - l;
              const int64_t node = s_leaf[b] >> shift;
              if (b == 0 || (s_leaf[b - 1] >> shift) != node) {
                nodes[u] = node; shifts[u] = shift; firsts[u] = b;
                sacc[u] = ld_cg(heap + node);
              }
            }
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            if (nodes[u] < 0) continue;
            const int64_t node = nodes[u];
            const int shift = shifts[u];
            float acc = sacc[u];
            int j = firsts[u];
            // whole groups of 8 inside the run: the loads are independent, the adds stay ordered
            while (j + 8 <= pass_n && (s_leaf[j + 7] >> shift) == node) {
              float d[8];
#pragma unroll
              for (int v = 0; v < 8; ++v) d[v] = s_diff[j + v];
#pragma unroll
              for (int v = 0; v < 8; ++v) acc = acc + d[v];
              j += 8;
            }
            for (; j < pass_n && (s_leaf[j] >> shift) == node; ++j) acc = acc + s_diff[j];
            st_cg(heap + node, acc);
          }
        }
        __syncthreads();
        continue;
      }
    }
    if (warp == 0) {
      for (int c0 = 0; c0 < pass_n; c0 += 32) {
        const int cnt = min(32, pass_n - c0);
        const bool active = lane < cnt;
        int64_t li = -1;
        float v = 0.0f;
        if (active) {
          li = idx[base + c0 + lane];
          v = val[base + c0 + lane];
          if (li < 0 || li >= cap) { *err = 1; li = -1; }
        }
        const int64_t pos = li >= 0 ? cap + li : -1;
        const long long key = li >= 0 ? (long long)pos : -(long long)(lane + 1);
        const unsigned m = __match_any_sync(full, key);
        const bool leader = (li >= 0) && (lane == __ffs(m) - 1);
        float L = leader ? ld_cg(heap + pos) : 0.0f;
        s_leaf[c0 + lane] = pos;
        if (li < 0) s_diff[c0 + lane] = 0.0f;
        for (int b = 0; b < cnt; ++b) {
          const float vb = __shfl_sync(full, v, b);
          if (leader && ((m >> b) & 1u)) {
            const float d = vb - L;  // updateValue: value - leaf   sum_tree.h:33
            L = L + d;               // updateValue_: node->value += diff   :84
            s_diff[c0 + b] = d;
          }
        }
        if (leader) st_cg(heap + pos, L);
        __syncwarp();
      }
    }
    __syncthreads();
    // internal levels, deepest first is not required: levels are independent.
    for (int l = depth - 1 - warp; l >= 0; l -= nwarps) {
      const int shift = depth - l;
      for (int c0 = 0; c0 < pass_n; c0 += 32) {
        const int cnt = min(32, pass_n - c0);
        const int64_t pos = lane < cnt ? s_leaf[c0 + lane] : -1;
        const int64_t node = pos >= 0 ? (pos >> shift) : -1;
        const long long key = pos >= 0 ? (long long)node : -(long long)(lane + 1);
        const unsigned m = __match_any_sync(full, key);
        const bool leader = (pos >= 0) && (lane == __ffs(m) - 1);
        float s = leader ? ld_cg(heap + node) : 0.0f;
        for (int b = 0; b < cnt; ++b) {
          const float db = s_diff[c0 + b];
          if (leader && ((m >> b) & 1u)) s = s + db;
        }
        if (leader) st_cg(heap + node, s);
        __syncwarp();
      }
    }
    __syncthreads();
  }
}

// ---- ordered update of a contiguous ascending leaf range (PB-3 inserts) ----
__global__ void st_range_leaf_kernel(float* __restrict__ heap, int64_t cap, int64_t start,
                                     int64_t n, const float* __restrict__ val, float fill,
                                     const float* __restrict__ fill_dev, float* __restrict__ diff,
                                     const int64_t* __restrict__ rng = nullptr) {
  if (rng != nullptr) { start = rng[0]; n = rng[1]; }    // bounds from device memory (hb_st_update_range_indirect)
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  float* leaf = heap + cap + start + k;
  const float old = *leaf;
  const float v = val ? val[k] : (fill_dev ? *fill_dev : fill);
  const float d = v - old;
  *leaf = old + d;
  diff[k] = d;
}

// blockIdx.y = internal level (0 = root); one thread per covered node adds the
// diffs of its covered leaves in ascending (= batch) order.
__global__ void st_range_internal_kernel(float* __restrict__ heap, int depth, int64_t cap,
                                         int64_t start, int64_t n,
                                         const float* __restrict__ diff, int level0,
                                         const int64_t* __restrict__ rng = nullptr) {
  if (rng != nullptr) { start = rng[0]; n = rng[1]; if (n <= 0) return; }
  const int l = blockIdx.y + level0;
  const int shift = depth - l;
  const int64_t first = (cap + start) >> shift;
  const int64_t last = (cap + start + n - 1) >> shift;
  const int64_t node = first + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (node > last) return;
  int64_t lo = (node << shift) - cap;            // first leaf under node
  int64_t hi = lo + ((int64_t)1 << shift);       // one past last leaf
  lo = max(lo, start);
  hi = min(hi, start + n);
  float s = heap[node];
  const float* d = diff + (lo - start);
  const int64_t cnt = hi - lo;
  int64_t k = 0;
  for (; k + 8 <= cnt; k += 8) {
    float t[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) t[j] = __ldg(d + k + j);
#pragma unroll
    for (int j = 0; j < 8; ++j) s = s + t[j];
  }
  for (; k < cnt; ++k) s = s + __ldg(d + k);
  heap[node] = s;
}

// The levels near the root cover long runs of leaves (the root: all n of them) and the additions
// of one node must stay in batch order, so a node is one dependent chain of fp32 adds.  Here a
// WARP owns a node: the lanes fetch 32 diffs at a time (coalesced, the next 32 prefetched) and
// every lane replays the same ordered chain from shuffles -- the chain runs at add latency instead
// of load latency (a 65,536-row insert: 1.0 ms -> ~0.2 ms).  blockIdx.y = level (0 = root).
__global__ void st_range_internal_warp_kernel(float* __restrict__ heap, int depth, int64_t cap,
                                              int64_t start, int64_t n,
                                              const float* __restrict__ diff, int level0,
                                              const int64_t* __restrict__ rng = nullptr) {
  if (rng != nullptr) { start = rng[0]; n = rng[1]; if (n <= 0) return; }
  const int l = blockIdx.y + level0;
  const int shift = depth - l;
  const int lane = threadIdx.x & 31;
  const int64_t first = (cap + start) >> shift;
  const int64_t last = (cap + start + n - 1) >> shift;
  const int64_t node = first + (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (node > last) return;
  int64_t lo = (node << shift) - cap;
  int64_t hi = lo + ((int64_t)1 << shift);
  lo = max(lo, start);
  hi = min(hi, start + n);
  const float* d = diff + (lo - start);
  const int64_t cnt = hi - lo;
  float s = heap[node];
  float next = lane < cnt ? __ldg(d + lane) : 0.f;
  for (int64_t k = 0; k < cnt; k += 32) {
    const float cur = next;
    if (k + 32 + lane < cnt) next = __ldg(d + k + 32 + lane);
    if (cnt - k >= 32) {
#pragma unroll
      for (int j = 0; j < 32; ++j) s = s + __shfl_sync(0xffffffffu, cur, j);
    } else {
      const int m = (int)(cnt - k);
      for (int j = 0; j < m; ++j) s = s + __shfl_sync(0xffffffffu, cur, j);
    }
  }
  if (lane == 0) heap[node] = s;
}

// ---- the levels next to the root: an ORDERED fp32 sum as a parallel scan ------------------------
// A node near the root adds thousands of diffs (the root: all n) and the reference adds them one
// by one in batch order, each add rounded to fp32 (sum_tree.h:80-90) -- one dependent chain, at
// best 4 cycles per element (65,536 rows: 0.14 ms; the shuffle replay above: 0.32 ms, half of a
// self-play training tick).  The chain is only apparent.  While the running sum s stays inside one
// binade [2^e, 2^(e+1)) it is an integer S in [2^23, 2^24) times u = 2^(e-23), and
//     fl32(s + d) = (S + RN(d / u)) * u
// where RN rounds to nearest and the exact tie d/u = m + 1/2 goes to the EVEN neighbour of S + m.
// So one element is the map  S -> S + a  or (tie)  S -> even_up(S + m); a run of elements composes
// to  S -> S + A  (no tie in the run)  or  S -> even_up(S + A) + R  (A: everything up to the first
// tie, R: everything after it, which no longer depends on S because the value after a tie is even).
// These maps compose associatively: every thread folds its slice of the run, one block scan gives
// each thread its incoming S, and a second pass over the slice checks that every exact partial sum
// stayed inside [2^23, 2^24) -- the premise.  Where it did not (the sum crossed a binade, a diff was
// as large as the sum, NaN ...), the verified prefix is kept, the offending slice is added the slow
// way in fp32, and the rest is retried in the new binade.  Bit-identical to the chain by
// construction; tests/test_sum_tree_gpu.py drives ties, crossings, cancellations and NaNs through it.
constexpr int kScanThreads = 1024;
constexpr int kScanMinRun = 512;     // shorter remainders: plain ordered chain by one warp
constexpr int kScanSlice = 8;        // elements per thread and window
constexpr int kScanSpanLog2 = 11;    // levels whose nodes cover >= 2048 leaves use the scan

struct SegFn {
  long long A, R;
  int tie;
};

__device__ __forceinline__ SegFn seg_compose(const SegFn& f, const SegFn& g) {   // f first, then g
  SegFn o;
  if (!f.tie) {
    o.tie = g.tie; o.A = f.A + g.A; o.R = g.R;
  } else if (!g.tie) {
    o.tie = 1; o.A = f.A; o.R = f.R + g.A;
  } else {
    long long t = f.R + g.A;
    t += t & 1;
    o.tie = 1; o.A = f.A; o.R = t + g.R;
  }
  return o;
}

__device__ __forceinline__ long long seg_apply(const SegFn& f, long long S) {
  long long t = S + f.A;
  if (!f.tie) return t;
  t += t & 1;
  return t + f.R;
}

__device__ __forceinline__ SegFn seg_shfl_up(const SegFn& v, int delta) {
  SegFn o;
  o.A = __shfl_up_sync(0xffffffffu, v.A, delta);
  o.R = __shfl_up_sync(0xffffffffu, v.R, delta);
  o.tie = __shfl_up_sync(0xffffffffu, v.tie, delta);
  return o;
}

// ordered chain s = s + d[0] + d[1] ... by one warp (every lane returns the result)
__device__ __forceinline__ float warp_chain(float s, const float* __restrict__ d, int64_t cnt, int lane) {
  float next = lane < cnt ? __ldg(d + lane) : 0.f;
  for (int64_t k = 0; k < cnt; k += 32) {
    const float cur = next;
    if (k + 32 + lane < cnt) next = __ldg(d + k + 32 + lane);
    const int m = cnt - k >= 32 ? 32 : (int)(cnt - k);
    for (int j = 0; j < m; ++j) s = s + __shfl_sync(0xffffffffu, cur, j);
  }
  return s;
}

__global__ void __launch_bounds__(kScanThreads)
st_range_internal_scan_kernel(float* __restrict__ heap, int depth, int64_t cap, int64_t start,
                              int64_t n, const float* __restrict__ diff,
                              const int64_t* __restrict__ rng = nullptr) {
  if (rng != nullptr) { start = rng[0]; n = rng[1]; if (n <= 0) return; }
  const int l = blockIdx.y;
  const int shift = depth - l;
  const int64_t first = (cap + start) >> shift;
  const int64_t last = (cap + start + n - 1) >> shift;
  const int64_t node = first + blockIdx.x;
  if (node > last) return;
  int64_t lo = (node << shift) - cap;
  int64_t hi = lo + ((int64_t)1 << shift);
  lo = max(lo, start);
  hi = min(hi, start + n);
  const float* d = diff + (lo - start);
  const int64_t cnt = hi - lo;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ float sh_s;
  __shared__ int sh_bad;
  __shared__ SegFn sh_warp[32];
  constexpr long long kLo = 1ll << 23, kHi = (1ll << 24) - 1;

  float s = heap[node];
  int64_t pos = 0;
  while (pos < cnt) {                         // every branch below is uniform over the CTA
    const int64_t rem = cnt - pos;
    const uint32_t bits = __float_as_uint(s);
    const int eb = (int)((bits >> 23) & 0xffu);
    const bool speculate = rem >= kScanMinRun && s > 0.f && eb >= 1 && eb <= 254;
    if (!speculate) {
      const int64_t take = rem < kScanMinRun ? rem : (int64_t)kScanMinRun;
      if (warp == 0) {
        const float r = warp_chain(s, d + pos, take, lane);
        if (lane == 0) sh_s = r;
      }
      __syncthreads();
      s = sh_s;
      __syncthreads();
      pos += take;
      continue;
    }
    // one window of at most kScanThreads * kScanSlice elements: a failed premise costs one window,
    // not the whole remainder (a node that starts empty crosses a binade every doubling)
    const int64_t win = rem < (int64_t)kScanThreads * kScanSlice ? rem : (int64_t)kScanThreads * kScanSlice;
    const int64_t seg = (win + kScanThreads - 1) / kScanThreads;      // <= kScanSlice
    const int64_t wend = pos + win;
    int64_t b = pos + (int64_t)tid * seg;
    if (b > wend) b = wend;
    const int64_t e = b + seg < wend ? b + seg : wend;
    // d / u with u = 2^(eb-127-23): a power-of-two scaling, exact in double
    const double scale = __longlong_as_double((long long)(1173 - eb) << 52);
    if (tid == 0) sh_bad = kScanThreads;
    // the slice in registers; 0 past its end (d = 0 is the identity map)
    double Dv[kScanSlice];
    bool fin = true;
#pragma unroll
    for (int j = 0; j < kScanSlice; ++j) {
      const float x = b + j < e ? __ldg(d + b + j) : 0.f;
      Dv[j] = (double)x * scale;
      fin = fin && (fabs(Dv[j]) < 16777216.0);
    }
    // pass 1: this slice as a map on S (a slice with a non-finite / oversized element is
    // flagged below; nothing after it is kept)
    SegFn f;
    f.A = 0; f.R = 0; f.tie = 0;
#pragma unroll
    for (int j = 0; j < kScanSlice; ++j) {
      const double D = fin ? Dv[j] : 0.0;
      const double fl = floor(D);
      const double fr = D - fl;
      long long m = (long long)fl;
      const bool tie = fr == 0.5;
      if (fr > 0.5) m += 1;
      if (!f.tie) {
        f.A += m;
        f.tie = tie ? 1 : 0;
      } else if (!tie) {
        f.R += m;
      } else {
        long long t = f.R + m;
        t += t & 1;
        f.R = t;
      }
    }
    // block-wide exclusive scan under composition
    SegFn inc = f;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const SegFn up = seg_shfl_up(inc, o);
      if (lane >= o) inc = seg_compose(up, inc);
    }
    if (lane == 31) sh_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      SegFn w = sh_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const SegFn up = seg_shfl_up(w, o);
        if (lane >= o) w = seg_compose(up, w);
      }
      sh_warp[lane] = w;
    }
    __syncthreads();
    SegFn pre = seg_shfl_up(inc, 1);
    if (lane == 0) { pre.A = 0; pre.R = 0; pre.tie = 0; }
    if (warp > 0) pre = seg_compose(sh_warp[warp - 1], pre);
    const long long S0 = (long long)((bits & 0x7fffffu) | 0x800000u);
    const long long S_in = seg_apply(pre, S0);
    // pass 2: replay the slice from its true incoming S and check the premise at every element
    long long S = S_in;
    bool ok = fin;
#pragma unroll
    for (int j = 0; j < kScanSlice; ++j) {
      const double D = fin ? Dv[j] : 0.0;
      const double fl = floor(D);
      const double fr = D - fl;
      const long long m = (long long)fl;
      if (S + m < kLo) ok = false;                     // exact sum below 2^e: finer grid there
      long long t = S + m;
      if (fr == 0.5) t += t & 1;
      else if (fr > 0.5) t += 1;
      if (t > kHi) ok = false;                         // reached 2^(e+1): coarser grid from here
      S = t;
    }
    if (!ok) atomicMin(&sh_bad, tid);
    __syncthreads();
    const int bad = sh_bad;
    if (bad == kScanThreads) {
      if (tid == kScanThreads - 1) sh_s = __uint_as_float(((uint32_t)eb << 23) | ((uint32_t)S & 0x7fffffu));
      __syncthreads();
      s = sh_s;
      __syncthreads();
      pos = wend;
    } else {
      if (tid == bad) sh_s = __uint_as_float(((uint32_t)eb << 23) | ((uint32_t)S_in & 0x7fffffu));
      __syncthreads();
      const float sb = sh_s;
      __syncthreads();
      int64_t bb = pos + (int64_t)bad * seg;
      if (bb > wend) bb = wend;
      const int64_t be = bb + seg < wend ? bb + seg : wend;
      if (warp == 0) {
        const float r = warp_chain(sb, d + bb, be - bb, lane);
        if (lane == 0) sh_s = r;
      }
      __syncthreads();
      s = sh_s;
      __syncthreads();
      pos = be;
    }
  }
  if (tid == 0) heap[node] = s;
}

int ensure_stage(hb_sumtree* t, size_t bytes) {
  if (t->stage_bytes >= bytes) return HB_OK;
  if (t->stage) cudaFree(t->stage);
  t->stage = nullptr;
  t->stage_bytes = 0;
  size_t want = bytes < 4096 ? 4096 : bytes * 2;
  HB_CUDA(cudaMalloc(&t->stage, want));
  t->stage_bytes = want;
  return HB_OK;
}

}  // namespace

extern "C" {

int hb_st_create(int64_t capacity, int device, hb_sumtree** out) {
  HB_REQUIRE(out != nullptr, "hb_st_create: out is NULL");
  HB_REQUIRE(capacity >= 1, "hb_st_create: capacity must be >= 1 (got %lld)", (long long)capacity);
  // depth_ = ceil(log2(capacity))   sum_tree.h:28
  const int depth = (int)ceil(log2((double)capacity));
  HB_REQUIRE(depth <= 30, "hb_st_create: capacity %lld too large", (long long)capacity);
  hb::DeviceGuard g(device);
  if (!g.ok) { hb::set_error("hb_st_create: cannot select CUDA device %d", device); return HB_ERR_CUDA; }
  hb_sumtree* t = new (std::nothrow) hb_sumtree();
  HB_REQUIRE(t != nullptr, "hb_st_create: out of host memory");
  t->device = device;
  t->depth = depth;
  t->cap = (int64_t)1 << depth;
  cudaError_t e = cudaMalloc(&t->heap, sizeof(float) * 2 * t->cap);
  if (e == cudaSuccess) e = cudaMemset(t->heap, 0, sizeof(float) * 2 * t->cap);
  if (e == cudaSuccess) e = cudaMalloc(&t->err, sizeof(int));
  if (e == cudaSuccess) e = cudaMemset(t->err, 0, sizeof(int));
  if (e != cudaSuccess) {
    if (t->heap) cudaFree(t->heap);
    if (t->err) cudaFree(t->err);
    delete t;
    return hb::cuda_fail(e, "hb_st_create alloc", __FILE__, __LINE__);
  }
  *out = t;
  return HB_OK;
}

int hb_st_destroy(hb_sumtree* t) {
  if (!t) return HB_OK;
  hb::DeviceGuard g(t->device);
  if (t->heap) cudaFree(t->heap);
  if (t->err) cudaFree(t->err);
  if (t->scratch) cudaFree(t->scratch);
  if (t->stage) cudaFree(t->stage);
  delete t;
  return HB_OK;
}

int64_t hb_st_capacity(const hb_sumtree* t) { return t ? t->cap : 0; }
int hb_st_depth(const hb_sumtree* t) { return t ? t->depth : -1; }
float* hb_st_heap(hb_sumtree* t) { return t ? t->heap : nullptr; }

int hb_st_update(hb_sumtree* t, const int64_t* d_idx, const float* d_val, int64_t n, void* stream) {
  HB_REQUIRE(t && n >= 0, "hb_st_update: bad arguments");
  if (n == 0) return HB_OK;
  HB_REQUIRE(d_idx && d_val, "hb_st_update: NULL buffer");
  hb::DeviceGuard g(t->device);
  const int warps = 32;   // the sorted fast path spreads (level, run) items over the whole CTA
  st_update_kernel<<<1, warps * 32, 0, (cudaStream_t)stream>>>(t->heap, t->depth, t->cap, d_idx,
                                                                d_val, n, t->err);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

static int st_update_range_impl(hb_sumtree* t, int64_t start, int64_t n, const float* d_val,
                                float fill, const float* d_fill, void* stream, const int64_t* d_rng = nullptr);

int hb_st_update_range(hb_sumtree* t, int64_t start, int64_t n, const float* d_val, float fill,
                       void* stream) {
  return st_update_range_impl(t, start, n, d_val, fill, nullptr, stream);
}

int hb_st_update_range_dev(hb_sumtree* t, int64_t start, int64_t n, const float* d_fill,
                           void* stream) {
  HB_REQUIRE(d_fill != nullptr, "hb_st_update_range_dev: NULL fill");
  return st_update_range_impl(t, start, n, nullptr, 0.f, d_fill, stream);
}

// d_rng != nullptr: the range is {start, n} in DEVICE memory (n <= the argument n, which then only sizes the
// grids and the scratch; inside the capacity by the caller's construction).
/* hb_st_update_range_dev with the range itself in device memory: d_rng = {start, n}, n <= n_max (n_max sizes
 * the launches; n = 0 is a no-op on the device).  For callers whose range is produced by a kernel (the replay
 * shard's device-side ring cursor): no host round trip between that kernel and the update. */
int hb_st_update_range_indirect(hb_sumtree* t, const int64_t* d_rng, int64_t n_max, const float* d_fill,
                                void* stream) {
  HB_REQUIRE(d_rng != nullptr && d_fill != nullptr && n_max >= 0, "hb_st_update_range_indirect: bad arguments");
  HB_REQUIRE(t && n_max <= t->cap, "hb_st_update_range_indirect: n_max beyond the capacity");
  return st_update_range_impl(t, 0, n_max, nullptr, 0.f, d_fill, stream, d_rng);
}

static int st_update_range_impl(hb_sumtree* t, int64_t start, int64_t n, const float* d_val,
                                float fill, const float* d_fill, void* stream, const int64_t* d_rng) {
  HB_REQUIRE(t && n >= 0, "hb_st_update_range: bad arguments");
  if (n == 0) return HB_OK;
  if (d_rng == nullptr && (start < 0 || start + n > t->cap)) {
    hb::set_error("hb_st_update_range: [%lld,%lld) outside capacity %lld", (long long)start,
                  (long long)(start + n), (long long)t->cap);
    return HB_ERR_INDEX;
  }
  hb::DeviceGuard g(t->device);
  cudaStream_t s = (cudaStream_t)stream;
  if (t->scratch_n < n) {
    HB_CUDA(cudaStreamSynchronize(s));
    if (t->scratch) cudaFree(t->scratch);
    t->scratch = nullptr;
    t->scratch_n = 0;
    HB_CUDA(cudaMalloc(&t->scratch, sizeof(float) * n));
    t->scratch_n = n;
  }
  const int threads = 256;
  st_range_leaf_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, s>>>(
      t->heap, t->cap, start, n, d_val, fill, d_fill, t->scratch, d_rng);
  HB_CUDA(cudaGetLastError());
  if (t->depth > 0) {
    // levels whose nodes cover >= 2048 leaves: CTA per node, ordered sum as a scan;
    // >= 64 leaves: warp per node, ordered chain; deeper levels: thread per node
    const int scan_levels = t->depth >= kScanSpanLog2 ? t->depth - kScanSpanLog2 + 1 : 0;   // 0 .. depth-11
    const int warp_levels = (t->depth >= 6 ? t->depth - 5 : 0) - scan_levels;               // .. depth-6
    if (scan_levels > 0) {
      const int64_t widest = (n >> kScanSpanLog2) + 2;
      dim3 grid((unsigned)widest, (unsigned)scan_levels);
      st_range_internal_scan_kernel<<<grid, kScanThreads, 0, s>>>(t->heap, t->depth, t->cap, start, n,
                                                                   t->scratch, d_rng);
      HB_CUDA(cudaGetLastError());
    }
    if (warp_levels > 0) {
      const int64_t widest = n / 64 + 2;
      dim3 grid((unsigned)((widest + 3) / 4), (unsigned)warp_levels);
      st_range_internal_warp_kernel<<<grid, 128, 0, s>>>(t->heap, t->depth, t->cap, start, n, t->scratch,
                                                         scan_levels, d_rng);
      HB_CUDA(cudaGetLastError());
    }
    const int upper = scan_levels + warp_levels;
    const int64_t widest = n / 2 + 2;  // nodes covered at the deepest internal level
    const int th = 128;
    dim3 grid((unsigned)((widest + th - 1) / th), (unsigned)(t->depth - upper));
    st_range_internal_kernel<<<grid, th, 0, s>>>(t->heap, t->depth, t->cap, start, n, t->scratch, upper, d_rng);
    HB_CUDA(cudaGetLastError());
  }
  return HB_OK;
}

// ---- allocation-free range update ---------------------------------------------------------------
// hb_st_update_range keeps its per-element diffs in t->scratch, grown with cudaMalloc on demand.  Callers
// that must not allocate (stream capture, serving loops, harnesses that forbid cudaMalloc) get the same
// kernels over a fixed scratch array that lives in the library image: kNoAllocMaxRange floats (256 KB) on
// every GPU that loads the library, until the process exits.
constexpr int64_t kNoAllocMaxRange = 65536;
static __device__ __align__(256) float g_noalloc_range_scratch[kNoAllocMaxRange];

namespace {
// In host code `g_noalloc_range_scratch` is only a stand-in for the symbol, not a device address: the
// address is looked up once per device, with that device current, and handed to the kernels.
int noalloc_scratch_ptr(int device, float** out) {
  static float* ptr[64] = {nullptr};
  if (device < 0 || device >= 64) { hb::set_error("hb_st_update_range_noalloc: device %d out of range", device); return HB_ERR_INVALID; }
  if (ptr[device] == nullptr) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_range_scratch);   // the caller made `device` current
    if (e != cudaSuccess) return hb::cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_range_scratch)", __FILE__, __LINE__);
    if (p == nullptr) { hb::set_error("hb_st_update_range_noalloc: NULL scratch address"); return HB_ERR_CUDA; }
    ptr[device] = p;
  }
  *out = ptr[device];
  return HB_OK;
}
}  // namespace

int hb_st_update_range_noalloc(hb_sumtree* t, int64_t start, int64_t n, const float* d_val, float fill,
                               void* stream) {
  HB_REQUIRE(t && n >= 0, "hb_st_update_range_noalloc: bad arguments");
  if (n == 0) return HB_OK;
  if (n > kNoAllocMaxRange) {
    hb::set_error("hb_st_update_range_noalloc: %lld elements exceed the fixed scratch of %lld (use hb_st_update_range, "
                  "or split the range)", (long long)n, (long long)kNoAllocMaxRange);
    return HB_ERR_INVALID;
  }
  if (start < 0 || start + n > t->cap) {
    hb::set_error("hb_st_update_range_noalloc: [%lld,%lld) outside capacity %lld", (long long)start,
                  (long long)(start + n), (long long)t->cap);
    return HB_ERR_INDEX;
  }
  hb::DeviceGuard g(t->device);
  float* scratch = nullptr;
  int rc = noalloc_scratch_ptr(t->device, &scratch);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  // The array keeps the previous call's diffs; nothing has to start at zero: st_range_leaf_kernel writes
  // diff[0 .. n) before any kernel of this call reads it (stream order), and nothing beyond n is read.
  const int threads = 256;
  st_range_leaf_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, s>>>(
      t->heap, t->cap, start, n, d_val, fill, nullptr, scratch);
  HB_CUDA(cudaGetLastError());
  if (t->depth > 0) {   // the level split of hb_st_update_range: scan / warp chain / thread per node
    const int scan_levels = t->depth >= kScanSpanLog2 ? t->depth - kScanSpanLog2 + 1 : 0;
    const int warp_levels = (t->depth >= 6 ? t->depth - 5 : 0) - scan_levels;
    if (scan_levels > 0) {
      dim3 grid((unsigned)((n >> kScanSpanLog2) + 2), (unsigned)scan_levels);
      st_range_internal_scan_kernel<<<grid, kScanThreads, 0, s>>>(t->heap, t->depth, t->cap, start, n, scratch);
      HB_CUDA(cudaGetLastError());
    }
    if (warp_levels > 0) {
      dim3 grid((unsigned)((n / 64 + 2 + 3) / 4), (unsigned)warp_levels);
      st_range_internal_warp_kernel<<<grid, 128, 0, s>>>(t->heap, t->depth, t->cap, start, n, scratch, scan_levels);
      HB_CUDA(cudaGetLastError());
    }
    const int upper = scan_levels + warp_levels;
    const int th = 128;
    dim3 grid((unsigned)((n / 2 + 2 + th - 1) / th), (unsigned)(t->depth - upper));
    st_range_internal_kernel<<<grid, th, 0, s>>>(t->heap, t->depth, t->cap, start, n, scratch, upper);
    HB_CUDA(cudaGetLastError());
  }
  return HB_OK;
}

int hb_st_get_indices(hb_sumtree* t, const float* d_q, int64_t* d_out, int64_t n, void* stream) {
  HB_REQUIRE(t && n >= 0, "hb_st_get_indices: bad arguments");
  if (n == 0) return HB_OK;
  HB_REQUIRE(d_q && d_out, "hb_st_get_indices: NULL buffer");
  hb::DeviceGuard g(t->device);
  const unsigned blocks = (unsigned)((n + kDescendThreads - 1) / kDescendThreads);
  if (n <= kLatencyKeys)
    st_descend_kernel<false, true><<<blocks, kDescendThreads, 0, (cudaStream_t)stream>>>(
        t->heap, t->depth, t->cap, d_q, d_out, nullptr, nullptr, n);
  else
    st_descend_kernel<false, false><<<blocks, kDescendThreads, 0, (cudaStream_t)stream>>>(
        t->heap, t->depth, t->cap, d_q, d_out, nullptr, nullptr, n);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_st_sample(hb_sumtree* t, const float* d_q, int64_t* d_idx_out, float* d_val_out,
                 float* d_total, int64_t n, void* stream) {
  HB_REQUIRE(t && n >= 0, "hb_st_sample: bad arguments");
  if (n == 0) {
    if (d_total) return hb_st_total(t, d_total, stream);
    return HB_OK;
  }
  HB_REQUIRE(d_q && d_idx_out && d_val_out, "hb_st_sample: NULL buffer");
  hb::DeviceGuard g(t->device);
  const unsigned blocks = (unsigned)((n + kDescendThreads - 1) / kDescendThreads);
  if (n <= kLatencyKeys)
    st_descend_kernel<true, true><<<blocks, kDescendThreads, 0, (cudaStream_t)stream>>>(
        t->heap, t->depth, t->cap, d_q, d_idx_out, d_val_out, d_total, n);
  else
    st_descend_kernel<true, false><<<blocks, kDescendThreads, 0, (cudaStream_t)stream>>>(
        t->heap, t->depth, t->cap, d_q, d_idx_out, d_val_out, d_total, n);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_st_get_values(hb_sumtree* t, const int64_t* d_idx, float* d_out, int64_t n, void* stream) {
  HB_REQUIRE(t && n >= 0, "hb_st_get_values: bad arguments");
  if (n == 0) return HB_OK;
  HB_REQUIRE(d_idx && d_out, "hb_st_get_values: NULL buffer");
  hb::DeviceGuard g(t->device);
  st_gather_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      t->heap, t->cap, d_idx, d_out, n, t->err);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_st_total(hb_sumtree* t, float* d_out, void* stream) {
  HB_REQUIRE(t && d_out, "hb_st_total: bad arguments");
  hb::DeviceGuard g(t->device);
  st_total_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(t->heap, d_out);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_st_check(hb_sumtree* t, void* stream) {
  HB_REQUIRE(t, "hb_st_check: NULL tree");
  hb::DeviceGuard g(t->device);
  int flag = 0;
  HB_CUDA(cudaMemcpyAsync(&flag, t->err, sizeof(int), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  HB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  if (flag) {
    HB_CUDA(cudaMemsetAsync(t->err, 0, sizeof(int), (cudaStream_t)stream));
    hb::set_error("sum tree: index out of range");
    return HB_ERR_INDEX;
  }
  return HB_OK;
}

// ---- host-buffer forms (the pybind surface) --------------------------------
// Synchronous calls, as the reference's are (it holds the GIL for the whole call).  They run on the legacy
// default stream, which does not order against non-blocking streams (the learner's, a caller's own):
// every form therefore starts with a device-wide synchronisation, so whatever was enqueued on ANY stream
// before the call -- an overlapped priority update, say -- is complete before the heap is touched, and
// returns only when its own work is done.
int hb_st_update_host(hb_sumtree* t, const int64_t* h_idx, const float* h_val, int64_t n) {
  HB_REQUIRE(t && n >= 0, "hb_st_update_host: bad arguments");
  if (n == 0) return HB_OK;
  HB_REQUIRE(h_idx && h_val, "hb_st_update_host: NULL buffer");
  hb::DeviceGuard g(t->device);
  HB_CUDA(cudaDeviceSynchronize());
  int rc = ensure_stage(t, (size_t)n * 12);
  if (rc) return rc;
  int64_t* d_idx = (int64_t*)t->stage;
  float* d_val = (float*)((char*)t->stage + (size_t)n * 8);
  HB_CUDA(cudaMemcpyAsync(d_idx, h_idx, (size_t)n * 8, cudaMemcpyHostToDevice, 0));
  HB_CUDA(cudaMemcpyAsync(d_val, h_val, (size_t)n * 4, cudaMemcpyHostToDevice, 0));
  rc = hb_st_update(t, d_idx, d_val, n, nullptr);
  if (rc) return rc;
  return hb_st_check(t, nullptr);
}

int hb_st_get_indices_host(hb_sumtree* t, const float* h_q, int64_t* h_out, int64_t n) {
  HB_REQUIRE(t && n >= 0, "hb_st_get_indices_host: bad arguments");
  if (n == 0) return HB_OK;
  HB_REQUIRE(h_q && h_out, "hb_st_get_indices_host: NULL buffer");
  hb::DeviceGuard g(t->device);
  HB_CUDA(cudaDeviceSynchronize());
  int rc = ensure_stage(t, (size_t)n * 12);
  if (rc) return rc;
  int64_t* d_out = (int64_t*)t->stage;
  float* d_q = (float*)((char*)t->stage + (size_t)n * 8);
  HB_CUDA(cudaMemcpyAsync(d_q, h_q, (size_t)n * 4, cudaMemcpyHostToDevice, 0));
  rc = hb_st_get_indices(t, d_q, d_out, n, nullptr);
  if (rc) return rc;
  HB_CUDA(cudaMemcpy(h_out, d_out, (size_t)n * 8, cudaMemcpyDeviceToHost));
  return HB_OK;
}

int hb_st_get_values_host(hb_sumtree* t, const int64_t* h_idx, float* h_out, int64_t n) {
  HB_REQUIRE(t && n >= 0, "hb_st_get_values_host: bad arguments");
  if (n == 0) return HB_OK;
  HB_REQUIRE(h_idx && h_out, "hb_st_get_values_host: NULL buffer");
  hb::DeviceGuard g(t->device);
  HB_CUDA(cudaDeviceSynchronize());
  int rc = ensure_stage(t, (size_t)n * 12);
  if (rc) return rc;
  int64_t* d_idx = (int64_t*)t->stage;
  float* d_out = (float*)((char*)t->stage + (size_t)n * 8);
  HB_CUDA(cudaMemcpyAsync(d_idx, h_idx, (size_t)n * 8, cudaMemcpyHostToDevice, 0));
  rc = hb_st_get_values(t, d_idx, d_out, n, nullptr);
  if (rc) return rc;
  HB_CUDA(cudaMemcpy(h_out, d_out, (size_t)n * 4, cudaMemcpyDeviceToHost));
  return hb_st_check(t, nullptr);
}

int hb_st_total_host(hb_sumtree* t, float* h_out) {
  HB_REQUIRE(t && h_out, "hb_st_total_host: bad arguments");
  hb::DeviceGuard g(t->device);
  HB_CUDA(cudaDeviceSynchronize());
  HB_CUDA(cudaMemcpy(h_out, t->heap + 1, sizeof(float), cudaMemcpyDeviceToHost));
  return HB_OK;
}

int hb_st_heap_host(hb_sumtree* t, float* h_out) {
  HB_REQUIRE(t && h_out, "hb_st_heap_host: bad arguments");
  hb::DeviceGuard g(t->device);
  HB_CUDA(cudaDeviceSynchronize());
  HB_CUDA(cudaMemcpy(h_out, t->heap, sizeof(float) * 2 * t->cap, cudaMemcpyDeviceToHost));
  return HB_OK;
}

}  // extern "C"
