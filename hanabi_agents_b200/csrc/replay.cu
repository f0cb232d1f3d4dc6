// Device-resident replay shard: ring buffer of bit-packed transitions + prioritized sampling.
//
// Replaces (path:line under /root/reference/hanabi_agents/rlax_dqn)
//   ExperienceBuffer   experience_buffer.py:5-97   (ring of obs_tm1 / act / rew / obs_t / lms / term)
//   PriorityBuffer     priority_buffer.py:13-52    (ring + SumTree: insert at max priority,
//                                                   stratified sample, priority update)
// and the transition bookkeeping of DQNAgent.add_experience_first / add_experience
// (rlax_rainbow.py:292-308: rows with step_type != 0 are appended in order, last_obs is refreshed).
//
// Observations stay in the packed form the env kernels emit (Full-2p: 96 B per row instead of
// the reference's 658 int8); legal moves are 1-2 words per row.  Rows are appended with one warp
// per transition (coalesced word copies); the masked append first ranks the selected rows with a
// two-level ballot scan so that their order -- and therefore the ring and sum-tree contents --
// is the reference's.
#include "common.cuh"
#include "sum_tree_dev.cuh"
#include <new>

struct hb_replay {
  int device = 0;
  int obs_len = 0, stride = 0, num_moves = 0, legal_words = 0;
  int64_t cap = 0, size = 0, oldest = 0;
  int prioritized = 0;
  float alpha = 0.6f;
  uint32_t* obs_tm1 = nullptr;
  uint32_t* obs_t = nullptr;
  uint32_t* lms = nullptr;
  int32_t* act = nullptr;
  float* rew = nullptr;
  uint8_t* term = nullptr;
  hb_sumtree* tree = nullptr;
  float* prio_range = nullptr;        // device: [0] max_priority, [1] min_priority, [2] fixed insert priority
  bool fixed_insert = false;          // new rows take prio_range[2] instead of the running max
  // scratch for the masked append
  int64_t scan_n = 0;
  int32_t* block_sum = nullptr;       // per 1024-row block
  int32_t* total = nullptr;           // device count of selected rows
  int32_t* h_total = nullptr;         // pinned host copy
  float* vals = nullptr;              // scratch for priorities / sampled values
  int64_t vals_n = 0;
  // A learner's captured graphs hold `vals`: a larger request allocates a NEW buffer, the old ones
  // stay alive until hb_rb_destroy, and vals_gen tells the learner to capture again (hb_rb_info 6).
  int64_t vals_gen = 0;
  float* vals_retired[16] = {nullptr};
  int n_retired = 0;
  // Device-side ring cursor (hb_rb_add_masked_async): d_cur = {oldest, size, write base of the current
  // append, its row count, range 0 {start, n}, range 1 {start, n}}.  While cur_on_device the host fields
  // `oldest` / `size` are stale; pull_cursor() brings them back (one stream synchronisation).
  int64_t* d_cur = nullptr;
  int64_t* h_cur = nullptr;           // pinned
  bool cur_on_device = false;
  cudaStream_t cur_stream = nullptr;
  // a copy of the cursor follows every asynchronous append to the host WITHOUT being waited for; when the
  // previous append's copy has arrived by the time of the next one, the host knows where that one starts
  // and skips the launches of the wrapped-around range when it cannot wrap
  int64_t* h_peek = nullptr;          // pinned: {oldest, size} after the last asynchronous append
  cudaEvent_t ev_peek = nullptr;
  bool peek_valid = false;            // a copy is in flight or has arrived, and nothing was appended since
};

namespace {

constexpr int kScanBlock = 1024;

// ---- masked append: rank the rows with step_type != 0 ----------------------------------
__global__ void __launch_bounds__(kScanBlock)
mask_count_kernel(const uint8_t* __restrict__ step_type, int64_t n, int32_t* __restrict__ block_sum) {
  __shared__ int warp_cnt[32];
  const int64_t i = (int64_t)blockIdx.x * kScanBlock + threadIdx.x;
  const bool sel = i < n && step_type[i] != 0;
  const unsigned b = __ballot_sync(0xffffffffu, sel);
  if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = __popc(b);
  __syncthreads();
  if (threadIdx.x < 32) {
    int v = warp_cnt[threadIdx.x];
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) block_sum[blockIdx.x] = v;
  }
}

// exclusive scan of the block sums in place (one block; sequential over chunks of 1024)
__global__ void __launch_bounds__(kScanBlock)
mask_scan_kernel(int32_t* __restrict__ block_sum, int n_blocks, int32_t* __restrict__ total) {
  __shared__ int sh[kScanBlock];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n_blocks; base += kScanBlock) {
    const int i = base + threadIdx.x;
    const int v = i < n_blocks ? block_sum[i] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < kScanBlock; o <<= 1) {
      const int t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
      __syncthreads();
      sh[threadIdx.x] += t;
      __syncthreads();
    }
    if (i < n_blocks) block_sum[i] = carry + sh[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 0) carry += sh[kScanBlock - 1];
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}

// Ranks the selected rows of a 1024-row block (ballot + block offsets from the scan), then copies
// the selected transitions with 16-byte accesses, one (row, 16-byte chunk) item per thread and
// pass: every load is independent of the others (the first version walked a warp's rows one after
// the other, 24 of 32 lanes busy, one row's latency at a time: 33 us for 60,000 rows; this: ~6 us).
// Also refreshes last_obs for the selected rows (rlax_rainbow.py:308).
__global__ void __launch_bounds__(kScanBlock)
masked_append_kernel(hb_replay rb, uint32_t* __restrict__ last_obs, const uint32_t* __restrict__ obs,
                     const uint32_t* __restrict__ lms, const int32_t* __restrict__ act,
                     const float* __restrict__ rew, const uint8_t* __restrict__ step_type, int64_t n,
                     const int32_t* __restrict__ block_off, const int64_t* __restrict__ d_cur = nullptr) {
  __shared__ int warp_off[32];
  __shared__ int64_t dst_row[kScanBlock];          // ring position of each selected row, -1 otherwise
  const int64_t write_base = d_cur != nullptr ? d_cur[2] : rb.oldest;   // device-side cursor (async form)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t row0 = (int64_t)blockIdx.x * kScanBlock;
  const int64_t i = row0 + threadIdx.x;
  const uint8_t st = i < n ? step_type[i] : 0;
  const bool sel = st != 0;
  const unsigned b = __ballot_sync(0xffffffffu, sel);
  if (lane == 0) warp_off[warp] = __popc(b);
  __syncthreads();
  if (threadIdx.x < 32) {
    const int v = warp_off[threadIdx.x];
    int s = v;
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, s, o);
      if (threadIdx.x >= o) s += t;
    }
    warp_off[threadIdx.x] = s - v;
  }
  __syncthreads();
  int64_t pos = -1;
  if (sel) {
    pos = write_base + block_off[blockIdx.x] + warp_off[warp] + __popc(b & ((1u << lane) - 1));
    if (pos >= rb.cap) pos -= rb.cap;                       // ring wrap (experience_buffer.py:62-81)
    for (int w = 0; w < rb.legal_words; ++w) rb.lms[pos * rb.legal_words + w] = lms[i * rb.legal_words + w];
    rb.act[pos] = act[i];
    rb.rew[pos] = rew[i];
    rb.term[pos] = st == 2;
  }
  dst_row[threadIdx.x] = pos;
  __syncthreads();
  const int q = rb.stride >> 2;                             // 16-byte chunks per row (stride % 4 == 0)
  const uint4* src_prev = reinterpret_cast<const uint4*>(last_obs) + row0 * q;
  const uint4* src_cur = reinterpret_cast<const uint4*>(obs) + row0 * q;
  uint4* upd = reinterpret_cast<uint4*>(last_obs) + row0 * q;
  uint4* o1 = reinterpret_cast<uint4*>(rb.obs_tm1);
  uint4* o2 = reinterpret_cast<uint4*>(rb.obs_t);
  const int items = kScanBlock * q;
  for (int item = threadIdx.x; item < items; item += kScanBlock) {
    const int r = item / q, c = item - r * q;
    const int64_t d = dst_row[r];
    if (d < 0) continue;
    const uint4 prev = src_prev[item];
    const uint4 cur = src_cur[item];
    o1[d * q + c] = prev;
    o2[d * q + c] = cur;
    upd[item] = cur;
  }
}

// add_experience_first: last_obs[step_type == FIRST] = obs  (rlax_rainbow.py:292-295), 16-byte items
__global__ void store_first_kernel(uint32_t* __restrict__ last_obs, const uint32_t* __restrict__ obs,
                                   const uint8_t* __restrict__ step_type, int stride, int64_t n) {
  const int q = stride >> 2;
  const int64_t total = n * q;
  const uint4* src = reinterpret_cast<const uint4*>(obs);
  uint4* dst = reinterpret_cast<uint4*>(last_obs);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / q;
    if (step_type[r] == 0) dst[i] = src[i];
  }
}

// contiguous append of n given rows (ExperienceBuffer.add_transitions), warp per row
__global__ void append_kernel(hb_replay rb, const uint32_t* __restrict__ obs_tm1,
                              const int32_t* __restrict__ act, const float* __restrict__ rew,
                              const uint32_t* __restrict__ obs_t, const uint32_t* __restrict__ lms,
                              const uint8_t* __restrict__ term, int64_t n) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= n) return;
  int64_t pos = rb.oldest + row;
  if (pos >= rb.cap) pos -= rb.cap;
  for (int w = lane; w < rb.stride; w += 32) {
    rb.obs_tm1[pos * rb.stride + w] = obs_tm1[row * rb.stride + w];
    rb.obs_t[pos * rb.stride + w] = obs_t[row * rb.stride + w];
  }
  if (lane < rb.legal_words) rb.lms[pos * rb.legal_words + lane] = lms[row * rb.legal_words + lane];
  if (lane == 0) {
    rb.act[pos] = act[row];
    rb.rew[pos] = rew[row];
    rb.term[pos] = term[row] != 0;
  }
}

// __getitem__(indices): fancy-index gather (experience_buffer.py:83-87), warp per row
__global__ void gather_kernel(hb_replay rb, const int64_t* __restrict__ idx, int64_t n,
                              uint32_t* __restrict__ obs_tm1, int32_t* __restrict__ act,
                              float* __restrict__ rew, uint32_t* __restrict__ obs_t,
                              uint32_t* __restrict__ lms, uint8_t* __restrict__ term, int* err) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= n) return;
  int64_t pos = idx[row];
  if (pos < 0) pos += rb.cap;                              // numpy negative indexing
  if (pos < 0 || pos >= rb.cap) { if (lane == 0) *err = 1; return; }
  for (int w = lane; w < rb.stride; w += 32) {
    if (obs_tm1) obs_tm1[row * rb.stride + w] = rb.obs_tm1[pos * rb.stride + w];
    if (obs_t) obs_t[row * rb.stride + w] = rb.obs_t[pos * rb.stride + w];
  }
  if (lms && lane < rb.legal_words) lms[row * rb.legal_words + lane] = rb.lms[pos * rb.legal_words + lane];
  if (lane == 0) {
    if (act) act[row] = rb.act[pos];
    if (rew) rew[row] = rb.rew[pos];
    if (term) term[row] = rb.term[pos];
  }
}

// prios = (value + 1e-10) / total in float64  (priority_buffer.py:42)
__global__ void prios_kernel(const float* __restrict__ val, const float* __restrict__ total, int64_t n,
                             double* __restrict__ p64, float* __restrict__ p32) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double p = ((double)val[i] + 1e-10) / (double)*total;
  if (p64) p64[i] = p;
  if (p32) p32[i] = (float)p;
}

// The learner's front end in ONE launch (one CTA of 1024 threads, B <= 1024):
//   keys -> indices      SumTree::getIndices   (priority_buffer.py:41; sum_tree.h:46-59)
//   prios                (getValues(indices) + 1e-10) / getTotalVal() in float64      (:42)
//   weights_is           (1 / prios)^beta / max(.) in float32         (rlax_rainbow.py:188-189)
//   transitions          self[indices]         (priority_buffer.py:46; experience_buffer.py:83-87)
// It replaces four dependent launches (H2D copy of the keys, descent, prios, gather) and the
// is_weights kernel of the network on the update's critical path.
// `keys` and `dyn_in` may be pinned host memory (read in place over PCIe: 1 KB).  dyn_in[0..7] are the
// step's scalars (beta, discount, ...): copied to dyn_out (device) for the kernels that follow.
__global__ void __launch_bounds__(1024)
sample_gather_kernel(hb_replay rb, const float* __restrict__ heap, int depth, int64_t tree_cap,
                     const float* __restrict__ keys, int B, const float* __restrict__ dyn_in,
                     float* __restrict__ dyn_out, float beta_value, int64_t* __restrict__ idx_out,
                     float* __restrict__ prios32, double* __restrict__ prios64, float* __restrict__ w_is,
                     uint32_t* __restrict__ obs_tm1, int32_t* __restrict__ act, float* __restrict__ rew,
                     uint32_t* __restrict__ obs_t, uint32_t* __restrict__ lms, uint8_t* __restrict__ term,
                     int* err) {
  __shared__ float top[hbst::kTopNodes];
  __shared__ int64_t s_pos[1024];
  __shared__ float red[32];
  extern __shared__ float4 scratch[];                      // [8 chunks][nd threads], nd = B rounded up to 32
  const int nd = (B + 31) & ~31;
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int k = threadIdx.x;
  if (k == 0) *err = 0;                                   // ordered before any `*err = 1` by stage_top's barrier
  const float quant = k < B ? keys[k] : 0.f;              // in flight while the top levels are staged
  float beta = beta_value;
  if (dyn_in != nullptr) {
    beta = dyn_in[0];
    if (k < 8) dyn_out[k] = dyn_in[k];
  }
  const int staged = hbst::stage_top(heap, tree_cap, top);
  const float root = top[1];
  float w = 0.f;
  if (k < B) {
    const float q = __fmul_rn(quant, root);               // sum_tree.h:47
    float value;
    const int64_t node = hbst::descend(heap, top, staged, depth, q, &value,
                                       reinterpret_cast<float*>(scratch + k), nd);
    int64_t pos = node - tree_cap;
    idx_out[k] = pos;
    if (pos < 0) pos += rb.cap;                           // numpy negative indexing (never taken here)
    if (pos < 0 || pos >= rb.cap) { *err = 1; pos = 0; }  // a leaf beyond the ring: caller's IndexError
    s_pos[k] = pos;
    // the row's scalars: loads issued before the arithmetic below (they miss to HBM; the shard is far
    // larger than L2), stored after it
    const int32_t r_act = act ? rb.act[pos] : 0;
    const float r_rew = rew ? rb.rew[pos] : 0.f;
    const uint8_t r_term = term ? rb.term[pos] : (uint8_t)0;
    uint32_t r_lms[2] = {0u, 0u};
    if (lms)
      for (int j = 0; j < rb.legal_words; ++j) r_lms[j] = rb.lms[pos * rb.legal_words + j];
    const double p = ((double)value + 1e-10) / (double)root;
    const float p32 = (float)p;
    if (prios64 != nullptr) prios64[k] = p;
    if (prios32 != nullptr) prios32[k] = p32;
    w = powf(1.0f / p32, beta);
    if (act) act[k] = r_act;
    if (rew) rew[k] = r_rew;
    if (term) term[k] = r_term;
    if (lms)
      for (int j = 0; j < rb.legal_words; ++j) lms[(int64_t)k * rb.legal_words + j] = r_lms[j];
  }
  // max over the batch (weights are positive; lanes past B hold 0)
  float mx = w;
  for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((k & 31) == 0) red[k >> 5] = mx;
  __syncthreads();                                        // also publishes s_pos
  // gather: (row, 16-byte chunk) items over the whole CTA.  Up to four items per thread and observation
  // array, ALL their loads (HBM misses) issued before the first store and before the weight normalisation.
  const int q16 = rb.stride >> 2;
  const uint4* s1 = reinterpret_cast<const uint4*>(rb.obs_tm1);
  const uint4* s2 = reinterpret_cast<const uint4*>(rb.obs_t);
  uint4* d1 = reinterpret_cast<uint4*>(obs_tm1);
  uint4* d2 = reinterpret_cast<uint4*>(obs_t);
  const int items = B * q16;
  uint4 v1[4], v2[4];
  auto load4 = [&](int it0) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int it = it0 + u * 1024;
      if (it < items) {
        const int r = it / q16, c = it - r * q16;
        const int64_t src = s_pos[r] * q16 + c;
        if (d1) v1[u] = s1[src];
        if (d2) v2[u] = s2[src];
      }
    }
  };
  auto store4 = [&](int it0) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int it = it0 + u * 1024;
      if (it < items) {
        if (d1) d1[it] = v1[u];
        if (d2) d2[it] = v2[u];
      }
    }
  };
  load4(k);
  if (k < 32) {                                            // warp 0, all lanes: the block maximum while the loads fly
    mx = red[k];
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (k == 0) red[0] = mx;
  }
  store4(k);
  for (int it0 = k + 4 * 1024; it0 < items; it0 += 4 * 1024) {
    load4(it0);
    store4(it0);
  }
  __syncthreads();
  if (k < B && w_is != nullptr) w_is[k] = w / red[0];
}

// priorities = (|td| + 1e-10) ** alpha in float32, running max / min  (priority_buffer.py:49-51).
// The power is evaluated in double and rounded once (= a correctly rounded powf).
__global__ void priorities_kernel(const float* __restrict__ td, int64_t n, float alpha,
                                  float* __restrict__ out, float* __restrict__ range) {
  __shared__ float smax[32], smin[32];
  float mx = -INFINITY, mn = INFINITY;
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
    const float x = fabsf(td[i]) + 1e-10f;       // the agent passes onp.abs(tds) (rlax_rainbow.py:334)
    const float p = (float)pow((double)x, (double)alpha);
    out[i] = p;
    mx = fmaxf(mx, p);
    mn = fminf(mn, p);
  }
  for (int o = 16; o; o >>= 1) {
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  if ((threadIdx.x & 31) == 0) { smax[threadIdx.x >> 5] = mx; smin[threadIdx.x >> 5] = mn; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { mx = fmaxf(mx, smax[w]); mn = fminf(mn, smin[w]); }
    range[0] = fmaxf(range[0], mx);
    range[1] = fminf(range[1], mn);
  }
}

// one byte per bit <-> packed words (the reference's np.byte layout, experience_buffer.py:9-12)
__global__ void pack_bits_kernel(const int8_t* __restrict__ in, int bits, uint32_t* __restrict__ out,
                                 int stride, int64_t n) {
  const int64_t total = n * stride;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / stride;
    const int w = (int)(i - r * stride);
    uint32_t v = 0;
    for (int j = 0; j < 32; ++j) {
      const int b = w * 32 + j;
      if (b < bits && in[r * bits + b] != 0) v |= 1u << j;
    }
    out[i] = v;
  }
}

__global__ void unpack_bits_kernel2(const uint32_t* __restrict__ packed, int stride, int bits,
                                    int8_t* __restrict__ out, int64_t n) {
  const int64_t total = n * bits;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / bits;
    const int b = (int)(i - r * bits);
    out[i] = (int8_t)((packed[r * stride + (b >> 5)] >> (b & 31)) & 1u);
  }
}

unsigned blocks_for(int64_t total, int threads) {
  int64_t b = (total + threads - 1) / threads;
  if (b > 148 * 32) b = 148 * 32;
  if (b < 1) b = 1;
  return (unsigned)b;
}

int ensure_vals(hb_replay* rb, int64_t n, cudaStream_t s) {
  if (rb->vals_n >= n) return HB_OK;
  (void)s;
  // sizes double, so at most a handful of buffers are ever retired; none is freed while graphs
  // captured with it may still be replayed
  int64_t want = rb->vals_n > 0 ? rb->vals_n : 256;
  while (want < n) want *= 2;
  float* fresh = nullptr;
  HB_CUDA(cudaMalloc(&fresh, sizeof(float) * (want + 1)));
  if (rb->vals) {
    if (rb->n_retired < 16) rb->vals_retired[rb->n_retired++] = rb->vals;
    else { cudaDeviceSynchronize(); cudaFree(rb->vals); }
  }
  rb->vals = fresh;
  rb->vals_n = want;
  ++rb->vals_gen;
  return HB_OK;
}

// bookkeeping of ExperienceBuffer.add_transitions (experience_buffer.py:49-81)
void advance(hb_replay* rb, int64_t n) {
  if (rb->oldest + n <= rb->cap) {
    if (rb->oldest + n == rb->cap) rb->size = rb->cap;
    rb->oldest = (rb->oldest + n) % rb->cap;
    if (rb->oldest > rb->size) rb->size = rb->oldest;
  } else {
    rb->oldest = rb->oldest + n - rb->cap;
    rb->size = rb->cap;
  }
}

// The same bookkeeping on the device (one thread): cur = {oldest, size, write base, count, range 0, range 1}.
// The ranges are the sum-tree leaves of the rows about to be written, [oldest, cap) then [0, tail).
__global__ void cursor_kernel(int64_t* __restrict__ cur, const int32_t* __restrict__ total, int64_t cap) {
  const int64_t n = *total;
  int64_t oldest = cur[0], size = cur[1];
  cur[2] = oldest;
  cur[3] = n;
  const int64_t first = oldest + n <= cap ? n : cap - oldest;
  cur[4] = oldest; cur[5] = first;
  cur[6] = 0;      cur[7] = n - first;
  if (n > 0) {                                   // advance(), word for word
    if (oldest + n <= cap) {
      if (oldest + n == cap) size = cap;
      oldest = (oldest + n) % cap;
      if (oldest > size) size = oldest;
    } else {
      oldest = oldest + n - cap;
      size = cap;
    }
  }
  cur[0] = oldest;
  cur[1] = size;
}
__global__ void cursor_set_kernel(int64_t* __restrict__ cur, int64_t oldest, int64_t size) {
  cur[0] = oldest;
  cur[1] = size;
}

// host fields <- device cursor (after asynchronous appends); every host-side use of oldest / size starts here
int pull_cursor(hb_replay* rb) {
  if (!rb->cur_on_device) return HB_OK;
  hb::DeviceGuard g(rb->device);
  HB_CUDA(cudaMemcpyAsync(rb->h_cur, rb->d_cur, 16, cudaMemcpyDeviceToHost, rb->cur_stream));
  HB_CUDA(cudaStreamSynchronize(rb->cur_stream));
  rb->oldest = rb->h_cur[0];
  rb->size = rb->h_cur[1];
  rb->cur_on_device = false;
  rb->peek_valid = false;
  return HB_OK;
}

// PriorityBuffer.add_transitions: leaves of the rows about to be written take max_priority, in the
// order of get_update_indices (experience_buffer.py:21-24): [oldest, cap) then [0, tail).
int insert_priorities(hb_replay* rb, int64_t n, cudaStream_t s) {
  if (!rb->prioritized || n == 0) return HB_OK;
  const int64_t first = rb->oldest + n <= rb->cap ? n : rb->cap - rb->oldest;
  const float* fill = rb->fixed_insert ? rb->prio_range + 2 : rb->prio_range;
  int rc = hb_st_update_range_dev(rb->tree, rb->oldest, first, fill, s);
  if (rc) return rc;
  if (first < n) rc = hb_st_update_range_dev(rb->tree, 0, n - first, fill, s);
  return rc;
}

}  // namespace

extern "C" {

int hb_rb_destroy(hb_replay* rb);

int hb_rb_create(int obs_len, int obs_stride_words, int num_moves, int64_t capacity, int prioritized,
                 float alpha, int device, hb_replay** out) {
  HB_REQUIRE(out != nullptr, "hb_rb_create: out is NULL");
  HB_REQUIRE(obs_len >= 1 && obs_stride_words * 32 >= obs_len && num_moves >= 1 && num_moves <= 64 &&
                 capacity >= 1,
             "hb_rb_create: bad arguments");
  HB_REQUIRE(obs_stride_words % 4 == 0, "hb_rb_create: rows must be a multiple of 16 bytes (stride %d words)",
             obs_stride_words);
  hb::DeviceGuard g(device);
  if (!g.ok) { hb::set_error("hb_rb_create: cannot select CUDA device %d", device); return HB_ERR_CUDA; }
  hb_replay* rb = new (std::nothrow) hb_replay();
  HB_REQUIRE(rb != nullptr, "hb_rb_create: out of host memory");
  rb->device = device;
  rb->obs_len = obs_len;
  rb->stride = obs_stride_words;
  rb->num_moves = num_moves;
  rb->legal_words = (num_moves + 31) / 32;
  rb->cap = capacity;
  rb->prioritized = prioritized ? 1 : 0;
  rb->alpha = alpha;
  const size_t c = (size_t)capacity;
  cudaError_t e = cudaMalloc(&rb->obs_tm1, c * rb->stride * 4);
  if (e == cudaSuccess) e = cudaMalloc(&rb->obs_t, c * rb->stride * 4);
  if (e == cudaSuccess) e = cudaMalloc(&rb->lms, c * rb->legal_words * 4);
  if (e == cudaSuccess) e = cudaMalloc(&rb->act, c * 4);
  if (e == cudaSuccess) e = cudaMalloc(&rb->rew, c * 4);
  if (e == cudaSuccess) e = cudaMalloc(&rb->term, c);
  if (e == cudaSuccess) e = cudaMalloc(&rb->total, 4);
  if (e == cudaSuccess) e = cudaMallocHost(&rb->h_total, 4);
  if (e == cudaSuccess) e = cudaMalloc(&rb->d_cur, 8 * sizeof(int64_t));
  if (e == cudaSuccess) e = cudaMemset(rb->d_cur, 0, 8 * sizeof(int64_t));
  if (e == cudaSuccess) e = cudaMallocHost(&rb->h_cur, 2 * sizeof(int64_t));
  if (e == cudaSuccess) e = cudaMallocHost(&rb->h_peek, 2 * sizeof(int64_t));
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&rb->ev_peek, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaMalloc(&rb->prio_range, 12);
  if (e == cudaSuccess) {
    // np.empty in the reference; zero here so that unwritten rows are defined
    cudaMemset(rb->obs_tm1, 0, c * rb->stride * 4);
    cudaMemset(rb->obs_t, 0, c * rb->stride * 4);
    cudaMemset(rb->lms, 0, c * rb->legal_words * 4);
    cudaMemset(rb->act, 0, c * 4);
    cudaMemset(rb->rew, 0, c * 4);
    cudaMemset(rb->term, 0, c);
    const float init[2] = {alpha, alpha};      // max_priority = min_priority = alpha  (:18-19)
    e = cudaMemcpy(rb->prio_range, init, 8, cudaMemcpyHostToDevice);
  }
  if (e != cudaSuccess) {
    hb_rb_destroy(rb);
    return hb::cuda_fail(e, "hb_rb_create alloc", __FILE__, __LINE__);
  }
  if (prioritized) {
    int rc = hb_st_create(capacity, device, &rb->tree);     // SumTree(capacity)  :17
    if (rc) { hb_rb_destroy(rb); return rc; }
  }
  *out = rb;
  return HB_OK;
}

int hb_rb_destroy(hb_replay* rb) {
  if (!rb) return HB_OK;
  hb::DeviceGuard g(rb->device);
  cudaDeviceSynchronize();
  if (rb->obs_tm1) cudaFree(rb->obs_tm1);
  if (rb->obs_t) cudaFree(rb->obs_t);
  if (rb->lms) cudaFree(rb->lms);
  if (rb->act) cudaFree(rb->act);
  if (rb->rew) cudaFree(rb->rew);
  if (rb->term) cudaFree(rb->term);
  if (rb->total) cudaFree(rb->total);
  if (rb->h_total) cudaFreeHost(rb->h_total);
  if (rb->d_cur) cudaFree(rb->d_cur);
  if (rb->h_cur) cudaFreeHost(rb->h_cur);
  if (rb->h_peek) cudaFreeHost(rb->h_peek);
  if (rb->ev_peek) cudaEventDestroy(rb->ev_peek);
  if (rb->prio_range) cudaFree(rb->prio_range);
  if (rb->block_sum) cudaFree(rb->block_sum);
  if (rb->vals) cudaFree(rb->vals);
  for (int i = 0; i < rb->n_retired; ++i) cudaFree(rb->vals_retired[i]);
  if (rb->tree) hb_st_destroy(rb->tree);
  delete rb;
  return HB_OK;
}

int64_t hb_rb_info(const hb_replay* rb, int what) {
  if (!rb) return -1;
  if ((what == 1 || what == 2) && pull_cursor(const_cast<hb_replay*>(rb))) return -1;
  switch (what) {
    case 0: return rb->cap;
    case 1: return rb->size;
    case 2: return rb->oldest;
    case 3: return rb->stride;
    case 4: return rb->legal_words;
    case 5: return rb->prioritized;
    case 6: return rb->vals_gen;
    default: return -1;
  }
}

hb_sumtree* hb_rb_tree(hb_replay* rb) { return rb ? rb->tree : nullptr; }

int hb_rb_add(hb_replay* rb, const uint32_t* d_obs_tm1, const int32_t* d_act, const float* d_rew,
              const uint32_t* d_obs_t, const uint32_t* d_lms, const uint8_t* d_term, int64_t n,
              void* stream) {
  HB_REQUIRE(rb && n >= 0, "hb_rb_add: bad arguments");
  if (n == 0) return HB_OK;
  HB_REQUIRE(d_obs_tm1 && d_act && d_rew && d_obs_t && d_lms && d_term, "hb_rb_add: NULL buffer");
  HB_REQUIRE(n <= rb->cap, "hb_rb_add: %lld rows exceed the capacity %lld", (long long)n, (long long)rb->cap);
  hb::DeviceGuard g(rb->device);
  cudaStream_t s = (cudaStream_t)stream;
  int rc = pull_cursor(rb);
  if (rc) return rc;
  rc = insert_priorities(rb, n, s);
  if (rc) return rc;
  append_kernel<<<(unsigned)((n + 7) / 8), 256, 0, s>>>(*rb, d_obs_tm1, d_act, d_rew, d_obs_t, d_lms,
                                                       d_term, n);
  HB_CUDA(cudaGetLastError());
  advance(rb, n);
  return HB_OK;
}

int hb_rb_store_first(hb_replay* rb, uint32_t* d_last_obs, const uint32_t* d_obs,
                      const uint8_t* d_step_type, int64_t n, void* stream) {
  HB_REQUIRE(rb && d_last_obs && d_obs && d_step_type && n >= 0, "hb_rb_store_first: bad arguments");
  HB_REQUIRE((((uintptr_t)d_last_obs | (uintptr_t)d_obs) & 15) == 0, "hb_rb_store_first: rows must be 16-byte aligned");
  if (n == 0) return HB_OK;
  hb::DeviceGuard g(rb->device);
  store_first_kernel<<<blocks_for(n * (rb->stride / 4), 256), 256, 0, (cudaStream_t)stream>>>(
      d_last_obs, d_obs, d_step_type, rb->stride, n);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

/* hb_rb_add_masked in two halves, so that a caller with several shards (the self-play session: one per
 * learning seat) can start every shard's row count before it waits for the first one.
 * begin: counts the rows that will be appended (step types -> per-block counts -> scan) and starts the copy
 * of the total to the host.  finish (same arguments, same stream): waits for that total -- the ring cursor
 * lives on the host, as in the reference -- and appends. */
int hb_rb_add_masked_begin(hb_replay* rb, const uint8_t* d_step_type, int64_t n, void* stream) {
  HB_REQUIRE(rb && d_step_type && n >= 0, "hb_rb_add_masked_begin: bad arguments");
  if (n == 0) return HB_OK;
  hb::DeviceGuard g(rb->device);
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t n_blocks = (n + kScanBlock - 1) / kScanBlock;
  if (rb->scan_n < n_blocks) {
    HB_CUDA(cudaStreamSynchronize(s));
    if (rb->block_sum) cudaFree(rb->block_sum);
    rb->block_sum = nullptr;
    rb->scan_n = 0;
    HB_CUDA(cudaMalloc(&rb->block_sum, sizeof(int32_t) * n_blocks));
    rb->scan_n = n_blocks;
  }
  mask_count_kernel<<<(unsigned)n_blocks, kScanBlock, 0, s>>>(d_step_type, n, rb->block_sum);
  HB_CUDA(cudaGetLastError());
  mask_scan_kernel<<<1, kScanBlock, 0, s>>>(rb->block_sum, (int)n_blocks, rb->total);
  HB_CUDA(cudaGetLastError());
  HB_CUDA(cudaMemcpyAsync(rb->h_total, rb->total, 4, cudaMemcpyDeviceToHost, s));
  return HB_OK;
}

int hb_rb_add_masked_finish(hb_replay* rb, uint32_t* d_last_obs, const uint32_t* d_obs, const uint32_t* d_lms,
                            const int32_t* d_act, const float* d_rew, const uint8_t* d_step_type, int64_t n,
                            int64_t* n_added, void* stream) {
  HB_REQUIRE(rb && d_last_obs && d_obs && d_lms && d_act && d_rew && d_step_type && n >= 0,
             "hb_rb_add_masked: bad arguments");
  HB_REQUIRE((((uintptr_t)d_last_obs | (uintptr_t)d_obs) & 15) == 0, "hb_rb_add_masked: rows must be 16-byte aligned");
  if (n_added) *n_added = 0;
  if (n == 0) return HB_OK;
  hb::DeviceGuard g(rb->device);
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t n_blocks = (n + kScanBlock - 1) / kScanBlock;
  HB_REQUIRE(rb->scan_n >= n_blocks, "hb_rb_add_masked_finish: no matching hb_rb_add_masked_begin");
  HB_CUDA(cudaStreamSynchronize(s));       // the ring cursor lives on the host, as in the reference
  {
    const int prc = pull_cursor(rb);       // ... unless asynchronous appends moved it to the device
    if (prc) return prc;
  }
  const int64_t cnt = *rb->h_total;
  if (n_added) *n_added = cnt;
  if (cnt == 0) return HB_OK;
  HB_REQUIRE(cnt <= rb->cap, "hb_rb_add_masked: %lld rows exceed the capacity %lld", (long long)cnt,
             (long long)rb->cap);
  int rc = insert_priorities(rb, cnt, s);
  if (rc) return rc;
  masked_append_kernel<<<(unsigned)n_blocks, kScanBlock, 0, s>>>(*rb, d_last_obs, d_obs, d_lms, d_act,
                                                                d_rew, d_step_type, n, rb->block_sum);
  HB_CUDA(cudaGetLastError());
  advance(rb, cnt);
  return HB_OK;
}

int hb_rb_add_masked(hb_replay* rb, uint32_t* d_last_obs, const uint32_t* d_obs, const uint32_t* d_lms,
                     const int32_t* d_act, const float* d_rew, const uint8_t* d_step_type, int64_t n,
                     int64_t* n_added, void* stream) {
  HB_REQUIRE(rb && d_last_obs && d_obs && d_lms && d_act && d_rew && d_step_type && n >= 0,
             "hb_rb_add_masked: bad arguments");
  if (n_added) *n_added = 0;
  int rc = hb_rb_add_masked_begin(rb, d_step_type, n, stream);
  if (rc) return rc;
  return hb_rb_add_masked_finish(rb, d_last_obs, d_obs, d_lms, d_act, d_rew, d_step_type, n, n_added, stream);
}

/* hb_rb_add_masked without the host in the loop: the number of appended rows never travels to the host.  The
 * ring cursor moves to device memory (cursor_kernel), the sum-tree leaves of the new rows are set through
 * hb_st_update_range_indirect, the append reads its write position from the device.  The host's view of
 * size / oldest_entry (hb_rb_info, hb_rb_add, hb_rb_add_masked) is refreshed on demand, with one stream
 * synchronisation.  Same results as hb_rb_add_masked, bit for bit (tests/test_replay_gpu.py). */
int hb_rb_add_masked_async(hb_replay* rb, uint32_t* d_last_obs, const uint32_t* d_obs, const uint32_t* d_lms,
                           const int32_t* d_act, const float* d_rew, const uint8_t* d_step_type, int64_t n,
                           void* stream) {
  HB_REQUIRE(rb && d_last_obs && d_obs && d_lms && d_act && d_rew && d_step_type && n >= 0,
             "hb_rb_add_masked_async: bad arguments");
  HB_REQUIRE((((uintptr_t)d_last_obs | (uintptr_t)d_obs) & 15) == 0,
             "hb_rb_add_masked_async: rows must be 16-byte aligned");
  HB_REQUIRE(n <= rb->cap, "hb_rb_add_masked_async: %lld rows exceed the capacity %lld", (long long)n,
             (long long)rb->cap);
  if (n == 0) return HB_OK;
  hb::DeviceGuard g(rb->device);
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t n_blocks = (n + kScanBlock - 1) / kScanBlock;
  if (rb->scan_n < n_blocks) {
    HB_CUDA(cudaStreamSynchronize(s));
    if (rb->block_sum) cudaFree(rb->block_sum);
    rb->block_sum = nullptr;
    rb->scan_n = 0;
    HB_CUDA(cudaMalloc(&rb->block_sum, sizeof(int32_t) * n_blocks));
    rb->scan_n = n_blocks;
  }
  // can the rows of this call wrap around the end of the ring?  Known when the host's cursor is current, or
  // when the copy that followed the previous asynchronous append has arrived; assumed otherwise.
  bool may_wrap = true;
  if (!rb->cur_on_device) may_wrap = rb->oldest + n > rb->cap;
  else if (rb->peek_valid && cudaEventQuery(rb->ev_peek) == cudaSuccess) may_wrap = rb->h_peek[0] + n > rb->cap;
  else (void)cudaGetLastError();             // cudaErrorNotReady is not an error
  if (!rb->cur_on_device) {                  // the host's cursor is the current one: hand it over
    cursor_set_kernel<<<1, 1, 0, s>>>(rb->d_cur, rb->oldest, rb->size);
    HB_CUDA(cudaGetLastError());
    rb->cur_on_device = true;
  } else if (rb->cur_stream != s) {
    // appends of one shard are ordered by the caller; a change of stream is ordered here
    HB_CUDA(cudaStreamSynchronize(rb->cur_stream));
  }
  rb->cur_stream = s;
  mask_count_kernel<<<(unsigned)n_blocks, kScanBlock, 0, s>>>(d_step_type, n, rb->block_sum);
  HB_CUDA(cudaGetLastError());
  mask_scan_kernel<<<1, kScanBlock, 0, s>>>(rb->block_sum, (int)n_blocks, rb->total);
  HB_CUDA(cudaGetLastError());
  cursor_kernel<<<1, 1, 0, s>>>(rb->d_cur, rb->total, rb->cap);
  HB_CUDA(cudaGetLastError());
  if (rb->prioritized) {
    const float* fill = rb->fixed_insert ? rb->prio_range + 2 : rb->prio_range;
    int rc = hb_st_update_range_indirect(rb->tree, rb->d_cur + 4, n, fill, s);
    if (rc) return rc;
    if (may_wrap) {                          // [0, tail): four launches that do nothing when there is no tail
      rc = hb_st_update_range_indirect(rb->tree, rb->d_cur + 6, n, fill, s);
      if (rc) return rc;
    }
  }
  masked_append_kernel<<<(unsigned)n_blocks, kScanBlock, 0, s>>>(*rb, d_last_obs, d_obs, d_lms, d_act,
                                                                d_rew, d_step_type, n, rb->block_sum, rb->d_cur);
  HB_CUDA(cudaGetLastError());
  HB_CUDA(cudaMemcpyAsync(rb->h_peek, rb->d_cur, 16, cudaMemcpyDeviceToHost, s));   // not waited for
  HB_CUDA(cudaEventRecord(rb->ev_peek, s));
  rb->peek_valid = true;
  return HB_OK;
}

int hb_rb_gather(hb_replay* rb, const int64_t* d_idx, int64_t n, uint32_t* d_obs_tm1, int32_t* d_act,
                 float* d_rew, uint32_t* d_obs_t, uint32_t* d_lms, uint8_t* d_term, void* stream) {
  HB_REQUIRE(rb && d_idx && n >= 0, "hb_rb_gather: bad arguments");
  if (n == 0) return HB_OK;
  hb::DeviceGuard g(rb->device);
  cudaStream_t s = (cudaStream_t)stream;
  HB_CUDA(cudaMemsetAsync(rb->total, 0, 4, s));
  gather_kernel<<<(unsigned)((n + 7) / 8), 256, 0, s>>>(*rb, d_idx, n, d_obs_tm1, d_act, d_rew, d_obs_t,
                                                       d_lms, d_term, rb->total);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

/* Reads the out-of-range flag of the last hb_rb_gather (synchronises). */
int hb_rb_check(hb_replay* rb, void* stream) {
  HB_REQUIRE(rb, "hb_rb_check: NULL buffer");
  hb::DeviceGuard g(rb->device);
  cudaStream_t s = (cudaStream_t)stream;
  HB_CUDA(cudaMemcpyAsync(rb->h_total, rb->total, 4, cudaMemcpyDeviceToHost, s));
  HB_CUDA(cudaStreamSynchronize(s));
  if (*rb->h_total) { hb::set_error("replay: index out of range"); return HB_ERR_INDEX; }
  return HB_OK;
}

int hb_rb_sample(hb_replay* rb, const float* d_keys, int64_t n, int64_t* d_idx, double* d_prios64,
                 float* d_prios32, void* stream) {
  HB_REQUIRE(rb && rb->prioritized && d_keys && d_idx && n >= 1, "hb_rb_sample: bad arguments");
  hb::DeviceGuard g(rb->device);
  cudaStream_t s = (cudaStream_t)stream;
  int rc = ensure_vals(rb, n, s);
  if (rc) return rc;
  // get_indices + get_values + get_total_val  (priority_buffer.py:41-42)
  rc = hb_st_sample(rb->tree, d_keys, d_idx, rb->vals, rb->vals + n, n, s);
  if (rc) return rc;
  prios_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(rb->vals, rb->vals + n, n, d_prios64, d_prios32);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_rb_sample_gather(hb_replay* rb, const float* d_keys, int B, const float* d_dyn_in, float* d_dyn_out,
                        float beta, int64_t* d_idx, float* d_prios32, double* d_prios64, float* d_w_is,
                        uint32_t* d_obs_tm1, int32_t* d_act, float* d_rew, uint32_t* d_obs_t,
                        uint32_t* d_lms, uint8_t* d_term, int pdl, void* stream) {
  HB_REQUIRE(rb && rb->prioritized && d_keys && d_idx && B >= 1 && B <= 1024,
             "hb_rb_sample_gather: bad arguments (prioritized shard, 1 <= B <= 1024)");
  HB_REQUIRE((d_dyn_in == nullptr) == (d_dyn_out == nullptr), "hb_rb_sample_gather: dyn_in and dyn_out go together");
  HB_REQUIRE((((uintptr_t)d_obs_tm1 | (uintptr_t)d_obs_t) & 15) == 0, "hb_rb_sample_gather: rows must be 16-byte aligned");
  hb::DeviceGuard g(rb->device);
  cudaStream_t s = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl ? 1 : 0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(1);
  cfg.blockDim = dim3(1024);
  cfg.dynamicSmemBytes = (size_t)((B + 31) & ~31) * hbst::kScratchBytesPerThread;
  cfg.stream = s;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  {   // static (top levels, positions) + dynamic (descent scratch) shared memory exceeds 48 KB from B = 256 on
    static bool configured[64] = {false};
    if (rb->device < 0 || rb->device >= 64 || !configured[rb->device]) {
      HB_CUDA(cudaFuncSetAttribute(sample_gather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
      if (rb->device >= 0 && rb->device < 64) configured[rb->device] = true;
    }
  }
  const float* heap = hb_st_heap(rb->tree);
  const int depth = hb_st_depth(rb->tree);
  const int64_t tcap = hb_st_capacity(rb->tree);
  HB_CUDA(cudaLaunchKernelEx(&cfg, sample_gather_kernel, *rb, heap, depth, tcap, d_keys, B, d_dyn_in, d_dyn_out,
                             beta, d_idx, d_prios32, d_prios64, d_w_is, d_obs_tm1, d_act, d_rew, d_obs_t,
                             d_lms, d_term, (int*)rb->total));
  return HB_OK;
}

int hb_rb_update_priorities(hb_replay* rb, const int64_t* d_idx, const float* d_td_abs, int64_t n,
                            void* stream) {
  HB_REQUIRE(rb && rb->prioritized && d_idx && d_td_abs && n >= 0, "hb_rb_update_priorities: bad arguments");
  if (n == 0) return HB_OK;
  hb::DeviceGuard g(rb->device);
  cudaStream_t s = (cudaStream_t)stream;
  int rc = ensure_vals(rb, n, s);
  if (rc) return rc;
  priorities_kernel<<<1, 256, 0, s>>>(d_td_abs, n, rb->alpha, rb->vals, rb->prio_range);
  HB_CUDA(cudaGetLastError());
  return hb_st_update(rb->tree, d_idx, rb->vals, n, s);
}

int hb_rb_set_insert_priority(hb_replay* rb, float value) {
  HB_REQUIRE(rb && rb->prioritized, "hb_rb_set_insert_priority: not a prioritized buffer");
  hb::DeviceGuard g(rb->device);
  rb->fixed_insert = value > 0.f;
  if (rb->fixed_insert) HB_CUDA(cudaMemcpy(rb->prio_range + 2, &value, 4, cudaMemcpyHostToDevice));
  return HB_OK;
}

int hb_rb_priority_range(hb_replay* rb, float* h_max, float* h_min, void* stream) {
  HB_REQUIRE(rb && h_max && h_min, "hb_rb_priority_range: bad arguments");
  hb::DeviceGuard g(rb->device);
  float v[2];
  HB_CUDA(cudaMemcpyAsync(v, rb->prio_range, 8, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  HB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  *h_max = v[0];
  *h_min = v[1];
  return HB_OK;
}

int hb_pack_bits(const int8_t* d_in, int bits, uint32_t* d_out, int stride_words, int64_t n, void* stream) {
  HB_REQUIRE(d_in && d_out && bits >= 1 && stride_words * 32 >= bits && n >= 0, "hb_pack_bits: bad arguments");
  if (n == 0) return HB_OK;
  pack_bits_kernel<<<blocks_for(n * stride_words, 256), 256, 0, (cudaStream_t)stream>>>(d_in, bits, d_out,
                                                                                       stride_words, n);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_unpack_bits(const uint32_t* d_packed, int stride_words, int bits, int8_t* d_out, int64_t n,
                   void* stream) {
  HB_REQUIRE(d_packed && d_out && bits >= 1 && stride_words * 32 >= bits && n >= 0, "hb_unpack_bits: bad arguments");
  if (n == 0) return HB_OK;
  unpack_bits_kernel2<<<blocks_for(n * bits, 256), 256, 0, (cudaStream_t)stream>>>(d_packed, stride_words,
                                                                                  bits, d_out, n);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

}  // extern "C"
