// Q-network (NoisyMLP / MLP), policy and learner step on tcgen05 GEMM tiles, and their C ABI.
//
// Reference (path:line under /root/reference/hanabi_agents/rlax_dqn):
//   NoisyLinear / NoisyMLP            noisy_mlp.py:49-91, 154-185
//   DQNPolicy.policy / eval_policy    rlax_rainbow.py:34-150
//   DQNLearning.update_q (C51)        rlax_rainbow.py:154-217
//   DQNLearning.update_q (double-Q)   rlax_dqn.py:144-233
//   optix.adam                        rlax_rainbow.py:257
//
// A noisy layer  y = x.w + b + x.(w_mu + w_sigma*eps_w) + (b_mu + b_sigma*eps_b)  has noise drawn
// at trace time (constants), so it is ONE effective layer  W = w + w_mu + w_sigma*eps_w,
// bias likewise; gradients fan out as dw = dw_mu = dW, dw_sigma = dW*eps_w.
// Masters, Adam moments and gradients are fp32; the GEMM operands are the fp16 roundings of the
// effective weights ([N, K] and its transpose) and fp16 activations, accumulated in fp32.
#include "tc_host.cuh"
#include "tc_chain.cuh"
#include <math.h>
#include <new>
#include <vector>

namespace {

constexpr int kMaxLayers = 6;
constexpr int kActChunk = 131072;   // rows per acting pass (bounds the workspace)

struct LayerGeo {
  int in, out;          // logical sizes (haiku w: [in, out])
  int K, N;             // GEMM sizes: K = padded in, N = padded / re-laid-out out
  long long eoff;       // offset of [W (N*K) | b (N)] in the flat effective-parameter space
  long long woff;       // offset of W in the fp16 weight arrays
  long long boff;       // offset of b in the effective-bias array
};

struct LayerTable {      // kernel argument for the fused Adam / refresh kernel
  int n_layers;
  LayerGeo l[kMaxLayers];
};

}  // namespace

struct hb_qnet {
  int device = 0;
  int obs_len = 0, obs_stride = 0;
  int A = 0, n_atoms = 1, noisy = 1;
  float vmax = 25.f;
  int n_layers = 0;
  LayerTable tab{};
  long long n_eff = 0, n_w = 0, n_b = 0;
  int kinds = 1;                       // master tensors per effective parameter (3 noisy, 1 plain)
  float* master = nullptr;             // [kinds][n_eff]
  float* tmaster = nullptr;            // masters of the target snapshot (save_weights)
  float* eps = nullptr;                // [n_eff] noise constants (noisy)
  float* adam_m = nullptr;             // [kinds == 3 ? 2 : 1][n_eff]: kinds 0 and 1 share their moments
  float* adam_v = nullptr;
  float* grad = nullptr;               // [n_eff] flat gradient bucket
  __half* w16 = nullptr;               // online effective weights [N, K] per layer
  __half* w16t = nullptr;              // transposes [K, N]
  float* beff = nullptr;               // effective biases
  __half* t16 = nullptr;               // target snapshot
  float* tbeff = nullptr;
  long long adam_step = 0;
  // acting workspace
  long long act_rows = 0;
  __half* act_x = nullptr;
  __half* act_h[2] = {nullptr, nullptr};
  float* act_q = nullptr;
  // learner workspace (batch capacity lb)
  int lb = 0, lbp = 0;
  uint32_t* xb = nullptr;              // [2B, obs_stride] packed rows: obs_tm1 then obs_t
  cudaStream_t side = nullptr;         // second stream: target forward / weight gradients run beside
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_join2 = nullptr, ev_loss = nullptr, ev_wt = nullptr,
              ev_fwd = nullptr,
              ev_dz[kMaxLayers] = {nullptr};
  cudaStream_t side2 = nullptr;        // small kernels beside the GEMMs of `side`
  // one stream per layer for {dW_l GEMM -> Adam_l}: layer l's update starts as soon as dZ_l exists and
  // runs beside the rest of the backward pass (single-process learner, hb_net_grads_ex adam_per_layer)
  cudaStream_t lstream[kMaxLayers] = {nullptr};
  cudaEvent_t ev_bias[kMaxLayers] = {nullptr}, ev_layer[kMaxLayers] = {nullptr}, ev_tick = nullptr;
  __half* x = nullptr;                 // [2B, K0]
  __half* xt = nullptr;                // [K0, Bp]
  __half* h[kMaxLayers] = {nullptr};   // hidden activations, online [2B, N]
  __half* ht[kMaxLayers] = {nullptr};  // transposed [N, 2Bp]
  __half* th[kMaxLayers] = {nullptr};  // target hidden [B, N]
  __half* dz[kMaxLayers] = {nullptr};  // dL/d(pre-activation) [B, N]
  __half* dzt[kMaxLayers] = {nullptr}; // transposed [N, Bp]
  float* logits = nullptr;             // online head [2B, N_last]
  float* q_on = nullptr;               // [2B, ldq]
  float* q_tg = nullptr;               // [B, ldq] target-network q at obs_t (TF-Rainbow variant)
  float* tlogits = nullptr;            // target head [B, N_last]
  float* w_is = nullptr;               // [B]
  float* loss_b = nullptr;             // [B] per-sample loss / td
  float* scalars = nullptr;            // [4]: mean loss, max weight, ...
  int act_reserve_sms = 0;             // hb_net_set_act_reserve
  unsigned int* chain = nullptr;       // tc_chain.cuh: [0..1] launch state, [2..] row-block counters
  int chain_B = 0;                     // batch the counters were last used with
  unsigned int* chain2 = nullptr;      // the same for the backward tail (dZ_0^T -> dW_0)
  int chain2_B = 0;
  float* bias_part = nullptr;          // [ceil(B / 64)][N_0] row sums of dZ_0^T (EpiParams::row_part)
};

namespace {

int ceil_to(int x, int m) { return (x + m - 1) / m * m; }

// ----------------------------------------------------------------- kernels ----
// bit-packed rows -> fp16 0/1 rows (and, optionally, the transposed copy of the first rows_t rows)
__global__ void unpack_rows_kernel(const uint32_t* __restrict__ packed, int stride, long long rows,
                                   __half* __restrict__ out, int ld, __half* __restrict__ out_t,
                                   int ld_t, long long rows_t) {
  const long long total = rows * stride;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / stride;
    const int w = (int)(i - r * stride);
    const uint32_t bits = __ldg(packed + i);
    uint32_t p[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const uint32_t lo = (bits >> (2 * j)) & 1u, hi = (bits >> (2 * j + 1)) & 1u;
      p[j] = (lo ? 0x3c00u : 0u) | (hi ? 0x3c000000u : 0u);   // half(1.0) = 0x3c00
    }
    uint4* dst = reinterpret_cast<uint4*>(out + r * ld + w * 32);
#pragma unroll
    for (int j = 0; j < 4; ++j) dst[j] = make_uint4(p[4 * j], p[4 * j + 1], p[4 * j + 2], p[4 * j + 3]);
    if (out_t != nullptr && r < rows_t) {
#pragma unroll 4
      for (int j = 0; j < 32; ++j)
        out_t[(long long)(w * 32 + j) * ld_t + r] = __ushort_as_half(((bits >> j) & 1u) ? 0x3c00 : 0);
    }
  }
}

// xt[k, r] = bit k of packed row r as fp16 (the K-major operand of the first layer's weight
// gradient): one thread per (k, 8 consecutive rows) -> one 16-byte store, coalesced along r.
__global__ void unpack_t_kernel(const uint32_t* __restrict__ packed, int stride, int rows,
                                __half* __restrict__ out_t, int ld_t) {
  const int groups = (rows + 7) / 8;
  const long long total = (long long)stride * 32 * groups;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i / groups), r0 = (int)(i - (long long)k * groups) * 8;
    uint32_t h[4] = {0u, 0u, 0u, 0u};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int r = r0 + j;
      const uint32_t bit = r < rows ? (__ldg(packed + (long long)r * stride + (k >> 5)) >> (k & 31)) & 1u : 0u;
      h[j >> 1] |= (bit ? 0x3c00u : 0u) << (16 * (j & 1));
    }
    *reinterpret_cast<uint4*>(out_t + (long long)k * ld_t + r0) = make_uint4(h[0], h[1], h[2], h[3]);
  }
}

__device__ __forceinline__ float uniform_from(unsigned long long seed, unsigned long long ctr,
                                              unsigned long long row) {
  unsigned long long x = seed + 0x9E3779B97F4A7C15ull * (row + 1);
  x ^= ctr * 0xBF58476D1CE4E5B9ull;
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  // (0, 1): the centre of one of 2^23 cells.  Never exactly 0 -- with u == 0 the reference's
  // `argmin(rnds > cpi)` (rlax_rainbow.py:44) returns action 0 even when it is illegal (cpi[0] == 0);
  // jax.random.uniform hits that once in 2^23 draws, a 131,072-game session every ~64 ticks.
  // Explicit `uniform` inputs keep the reference's behaviour, quirk included.
  return ((float)(x >> 41) + 0.5f) * (1.0f / 8388608.0f);
}

// DQNPolicy.policy / eval_policy after the network (rlax_rainbow.py:34-59,119-123,146-150):
//   q = where(legal, q, -inf); greedy probs = ties of the row max split evenly;
//   explore: probs = (1-eps)*greedy + eps*legal/num_legal; cpi = cumsum(probs), cpi[-1] = 1;
//            action = first a with not (u > cpi[a]), u in [0, 0.999)
//   exploit: rlax.greedy().sample = uniform among the maxima
__global__ void policy_kernel(const float* __restrict__ q, int ldq, const uint32_t* __restrict__ legal,
                              int legal_words, int A, long long n, int greedy, float epsilon,
                              const float* __restrict__ uniform, unsigned long long seed,
                              unsigned long long ctr, long long row0, int32_t* __restrict__ action,
                              float* __restrict__ q_out) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const float* qr = q + r * ldq;
  unsigned long long lm = legal[r * legal_words];
  if (legal_words > 1) lm |= (unsigned long long)legal[r * legal_words + 1] << 32;
  float mx = -INFINITY;
  int num_legal = 0;
  for (int a = 0; a < A; ++a) {
    const bool ok = (lm >> a) & 1ull;
    const float v = ok ? qr[a] : -INFINITY;
    if (q_out != nullptr) q_out[r * A + a] = v;
    mx = fmaxf(mx, v);
    num_legal += ok ? 1 : 0;
  }
  int n_max = 0;
  for (int a = 0; a < A; ++a) {
    const float v = ((lm >> a) & 1ull) ? qr[a] : -INFINITY;
    n_max += (v == mx) ? 1 : 0;       // with no legal move every entry is -inf == max, as in jnp
  }
  float u = uniform != nullptr ? uniform[r] : uniform_from(seed, ctr, (unsigned long long)(row0 + r));
  int act = 0;
  if (greedy) {
    // uniform among the n_max maxima
    int pick = min((int)(u * (float)n_max), n_max - 1), seen = 0;
    for (int a = 0; a < A; ++a) {
      const float v = ((lm >> a) & 1ull) ? qr[a] : -INFINITY;
      if (v == mx) {
        if (seen == pick) { act = a; break; }
        ++seen;
      }
    }
  } else {
    if (uniform == nullptr) u *= 0.999f;   // jax.random.uniform(maxval=0.999)
    const float pg = 1.0f / (float)n_max;
    const float pu = 1.0f / (float)num_legal;   // inf when no legal move, as legal / 0 in jnp
    float c = 0.f;
    act = A - 1;                                 // cpi[:, -1] = 1 and u < 1
    for (int a = 0; a < A - 1; ++a) {
      const bool ok = (lm >> a) & 1ull;
      const float v = ok ? qr[a] : -INFINITY;
      // separate roundings as in the reference's elementwise ops (no fused multiply-add)
      const float p = __fadd_rn(__fmul_rn(1.0f - epsilon, (v == mx) ? pg : 0.f),
                                __fmul_rn(epsilon, ok ? pu : 0.f));
      c = __fadd_rn(c, p);
      if (!(u > c)) { act = a; break; }
    }
  }
  action[r] = act;
}

// weights_is = (1/prios)^beta / max(.)   rlax_rainbow.py:188-189 (fp32)
__global__ void is_weights_kernel(const float* __restrict__ prios, int B, float beta,
                                  float* __restrict__ w) {
  __shared__ float red[32];
  float mx = 0.f;
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    const float v = powf(1.0f / prios[i], beta);
    w[i] = v;
    mx = fmaxf(mx, v);
  }
  for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x < 32) {
    mx = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (threadIdx.x == 0) red[0] = mx;
  }
  __syncthreads();
  mx = red[0];
  for (int i = threadIdx.x; i < B; i += blockDim.x) w[i] = w[i] / mx;
}

// The head layer's backward product folded into the C51 loss (see c51_loss_kernel<true>).
struct HeadBackward {
  const __half* w;          // head's effective weights [N, K]
  int K;                    // width of the previous layer (multiple of 8)
  const __half* gate;       // previous layer's activations [B.., K] (ReLU backward)
  __half* dz_prev;          // [B, K]
  __half* dzt_prev;         // [K, ld_dzt]
  unsigned long long* trace;  // debugging (HB_LOSS_TRACE): 8 globaltimer stamps per sample, or nullptr
};
constexpr int kHeadBwSmemMax = 112 * 1024;     // rows of one action (n_atoms x K halves) staged per sample

// One warp, one sample.  rlax.categorical_double_q_learning (rlax_rainbow.py:172-184):
//   a* = argmax_a q_sel;  p = softmax(target_logits[a*]);  target_z = r + gamma * z;
//   m = categorical_l2_project(target_z, p, z);  loss = -sum_i m_i log_softmax(logits_tm1[a])_i
// Writes loss_b and dL/dlogits * B / w-normalisation folded later:  dz[b, a*64+i] = w_b (p_i S - m_i);
// returns the lane's two values (atoms lane and lane + 32) as the fp16 numbers that were stored.
__device__ __forceinline__ void c51_row(int b, int lane, const float* __restrict__ logits_tm1,
                                        const float* __restrict__ q_sel, int ldq,
                                        const uint32_t* __restrict__ legal_t, int legal_words,
                                        const float* __restrict__ tlogits, int ldl, int a,
                                        const float* __restrict__ r_t, const uint8_t* __restrict__ term_t,
                                        int mask_terminal, const float* __restrict__ w_is, int A, int n_atoms,
                                        float vmax, float gamma, __half* __restrict__ dz,
                                        __half* __restrict__ dzt, int ld_dzt, float* __restrict__ loss_b,
                                        __half& h0, __half& h1) {
  const unsigned full = 0xffffffffu;
  const float vmin = -vmax, dzv = 2.0f * vmax / (float)(n_atoms - 1);
  const int i0 = lane, i1 = lane + 32;
  // everything that does not depend on a* is requested first (these loads used to sit behind the
  // argmax -> target logits -> projection chain, one L2 round trip each)
  const float* ol = logits_tm1 + (long long)b * ldl + a * 64;
  const float l0 = i0 < n_atoms ? ol[i0] : -INFINITY, l1 = i1 < n_atoms ? ol[i1] : -INFINITY;
  const float r = r_t[b];
  const float w = w_is[b];
  const bool terminal = mask_terminal && term_t != nullptr && term_t[b];
  // a* = first maximum of q_sel[b, :]  (jnp.argmax)
  float best = -INFINITY;
  int besta = 0x7fffffff;
  for (int aa = lane; aa < A; aa += 32) {
    float v = q_sel[(long long)b * ldq + aa];
    // TF-Rainbow: argmax(next_qt + next_legal_actions), illegal = -inf  (rainbow_agent.py:203-204)
    if (legal_t != nullptr && !((legal_t[(long long)b * legal_words + (aa >> 5)] >> (aa & 31)) & 1u)) v = -INFINITY;
    if (v > best) { best = v; besta = aa; }
  }
  for (int o = 16; o; o >>= 1) {
    const float ov = __shfl_xor_sync(full, best, o);
    const int oa = __shfl_xor_sync(full, besta, o);
    if (ov > best || (ov == best && oa < besta)) { best = ov; besta = oa; }
  }
  const int astar = besta == 0x7fffffff ? 0 : besta;
  // target distribution p_j
  const float* tl = tlogits + (long long)b * ldl + astar * 64;
  const float t0 = i0 < n_atoms ? tl[i0] : -INFINITY, t1 = i1 < n_atoms ? tl[i1] : -INFINITY;
  float mx = fmaxf(t0, t1);
  for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(full, mx, o));
  float e0 = i0 < n_atoms ? expf(t0 - mx) : 0.f, e1 = i1 < n_atoms ? expf(t1 - mx) : 0.f;
  float den = e0 + e1;
  for (int o = 16; o; o >>= 1) den += __shfl_xor_sync(full, den, o);
  const float p0 = e0 / den, p1 = e1 / den;
  // projection onto the fixed support: m_i = sum_j clip(1 - |clip(tz_j) - z_i| / dz, 0, 1) p_j
  const float g = terminal ? 0.f : gamma;
  const float zi0 = vmin + dzv * (float)i0, zi1 = vmin + dzv * (float)i1;
  float m0 = 0.f, m1 = 0.f;
  const float inv_dz = 1.0f / dzv;
#pragma unroll 4
  for (int j = 0; j < n_atoms; ++j) {
    const float pj = __shfl_sync(full, j < 32 ? p0 : p1, j & 31);
    float tz = r + g * (vmin + dzv * (float)j);
    tz = fminf(fmaxf(tz, vmin), vmax);
    // rlax.categorical_l2_project multiplies by the reciprocal spacing (d_pos = 1 / d_pos)
    m0 += fminf(fmaxf(1.0f - fabsf(tz - zi0) * inv_dz, 0.f), 1.f) * pj;
    m1 += fminf(fmaxf(1.0f - fabsf(tz - zi1) * inv_dz, 0.f), 1.f) * pj;
  }
  if (i0 >= n_atoms) m0 = 0.f;
  if (i1 >= n_atoms) m1 = 0.f;
  // log-softmax of the online logits of the action taken
  float lmx = fmaxf(l0, l1);
  for (int o = 16; o; o >>= 1) lmx = fmaxf(lmx, __shfl_xor_sync(full, lmx, o));
  const float f0 = i0 < n_atoms ? expf(l0 - lmx) : 0.f, f1 = i1 < n_atoms ? expf(l1 - lmx) : 0.f;
  float lden = f0 + f1;
  for (int o = 16; o; o >>= 1) lden += __shfl_xor_sync(full, lden, o);
  const float lse = lmx + logf(lden);
  float part = (i0 < n_atoms ? m0 * (l0 - lse) : 0.f) + (i1 < n_atoms ? m1 * (l1 - lse) : 0.f);
  float S = m0 + m1;
  for (int o = 16; o; o >>= 1) {
    part += __shfl_xor_sync(full, part, o);
    S += __shfl_xor_sync(full, S, o);
  }
  if (lane == 0) loss_b[b] = -part;
  const float d0 = w * (f0 / lden * S - m0), d1 = w * (f1 / lden * S - m1);
  h0 = __float2half_rn(i0 < n_atoms ? d0 : 0.f);
  h1 = __float2half_rn(i1 < n_atoms ? d1 : 0.f);
  const long long col = (long long)a * 64;
  if (i0 < n_atoms) {
    dz[(long long)b * ldl + col + i0] = h0;
    dzt[(col + i0) * ld_dzt + b] = h0;
  }
  if (i1 < n_atoms) {
    dz[(long long)b * ldl + col + i1] = h1;
    dzt[(col + i1) * ld_dzt + b] = h1;
  }
}

// The same row program spread over the four warps of a CTA (HEAD_BW): as one warp it was a 7.5 us chain of
// dependent loads, shuffle reductions and a 51-step projection loop (profiles/r02_loss_trace_*.txt).
//   warp 0: argmax -> target logits -> softmax p          warp 1: log-softmax of the online logits
//   all:    a quarter of the projection's 51 terms each   warp 0: the sums, loss and dL/dlogits
// The arithmetic and the order of every sum over atoms are c51_row's, except that m_i is accumulated in
// four partial sums (j = 0..12, 13..25, 26..38, 39..50) added in that order.
struct C51Shared {
  float p[64], soft[64], logp[64], pm[4][64];
};
__device__ __forceinline__ void c51_row_cta(C51Shared& sh, int b, int warp, int lane,
                                            const float* __restrict__ logits_tm1,
                                            const float* __restrict__ q_sel, int ldq,
                                            const uint32_t* __restrict__ legal_t, int legal_words,
                                            const float* __restrict__ tlogits, int ldl, int a,
                                            const float* __restrict__ r_t, const uint8_t* __restrict__ term_t,
                                            int mask_terminal, const float* __restrict__ w_is, int A, int n_atoms,
                                            float vmax, float gamma, __half* __restrict__ dz,
                                            __half* __restrict__ dzt, int ld_dzt, float* __restrict__ loss_b,
                                            float* __restrict__ gsm) {
  const unsigned full = 0xffffffffu;
  const float vmin = -vmax, dzv = 2.0f * vmax / (float)(n_atoms - 1);
  const int i0 = lane, i1 = lane + 32;
  const float r = r_t[b];
  const bool terminal = mask_terminal && term_t != nullptr && term_t[b];
  if (warp == 0) {
    float best = -INFINITY;
    int besta = 0x7fffffff;
    for (int aa = lane; aa < A; aa += 32) {
      float v = q_sel[(long long)b * ldq + aa];
      if (legal_t != nullptr && !((legal_t[(long long)b * legal_words + (aa >> 5)] >> (aa & 31)) & 1u)) v = -INFINITY;
      if (v > best) { best = v; besta = aa; }
    }
    for (int o = 16; o; o >>= 1) {
      const float ov = __shfl_xor_sync(full, best, o);
      const int oa = __shfl_xor_sync(full, besta, o);
      if (ov > best || (ov == best && oa < besta)) { best = ov; besta = oa; }
    }
    const int astar = besta == 0x7fffffff ? 0 : besta;
    const float* tl = tlogits + (long long)b * ldl + astar * 64;
    const float t0 = i0 < n_atoms ? tl[i0] : -INFINITY, t1 = i1 < n_atoms ? tl[i1] : -INFINITY;
    float mx = fmaxf(t0, t1);
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(full, mx, o));
    const float e0 = i0 < n_atoms ? expf(t0 - mx) : 0.f, e1 = i1 < n_atoms ? expf(t1 - mx) : 0.f;
    float den = e0 + e1;
    for (int o = 16; o; o >>= 1) den += __shfl_xor_sync(full, den, o);
    sh.p[i0] = e0 / den;
    sh.p[i1] = e1 / den;
  } else if (warp == 1) {
    const float* ol = logits_tm1 + (long long)b * ldl + a * 64;
    const float l0 = i0 < n_atoms ? ol[i0] : -INFINITY, l1 = i1 < n_atoms ? ol[i1] : -INFINITY;
    float lmx = fmaxf(l0, l1);
    for (int o = 16; o; o >>= 1) lmx = fmaxf(lmx, __shfl_xor_sync(full, lmx, o));
    const float f0 = i0 < n_atoms ? expf(l0 - lmx) : 0.f, f1 = i1 < n_atoms ? expf(l1 - lmx) : 0.f;
    float lden = f0 + f1;
    for (int o = 16; o; o >>= 1) lden += __shfl_xor_sync(full, lden, o);
    const float lse = lmx + logf(lden);
    sh.soft[i0] = f0 / lden;
    sh.soft[i1] = f1 / lden;
    sh.logp[i0] = i0 < n_atoms ? l0 - lse : 0.f;
    sh.logp[i1] = i1 < n_atoms ? l1 - lse : 0.f;
  }
  __syncthreads();
  {
    // projection onto the fixed support: m_i = sum_j clip(1 - |clip(tz_j) - z_i| / dz, 0, 1) p_j
    const float g = terminal ? 0.f : gamma;
    const float zi0 = vmin + dzv * (float)i0, zi1 = vmin + dzv * (float)i1;
    const float inv_dz = 1.0f / dzv;
    const int q4 = (n_atoms + 3) >> 2;
    const int j0 = warp * q4, j1 = (j0 + q4 < n_atoms) ? j0 + q4 : n_atoms;
    float m0 = 0.f, m1 = 0.f;
#pragma unroll 4
    for (int j = j0; j < j1; ++j) {
      const float pj = sh.p[j];
      float tz = r + g * (vmin + dzv * (float)j);
      tz = fminf(fmaxf(tz, vmin), vmax);
      m0 += fminf(fmaxf(1.0f - fabsf(tz - zi0) * inv_dz, 0.f), 1.f) * pj;
      m1 += fminf(fmaxf(1.0f - fabsf(tz - zi1) * inv_dz, 0.f), 1.f) * pj;
    }
    sh.pm[warp][i0] = m0;
    sh.pm[warp][i1] = m1;
  }
  __syncthreads();
  if (warp == 0) {
    float m0 = ((sh.pm[0][i0] + sh.pm[1][i0]) + sh.pm[2][i0]) + sh.pm[3][i0];
    float m1 = ((sh.pm[0][i1] + sh.pm[1][i1]) + sh.pm[2][i1]) + sh.pm[3][i1];
    if (i0 >= n_atoms) m0 = 0.f;
    if (i1 >= n_atoms) m1 = 0.f;
    float part = m0 * sh.logp[i0] + m1 * sh.logp[i1];
    float S = m0 + m1;
    for (int o = 16; o; o >>= 1) {
      part += __shfl_xor_sync(full, part, o);
      S += __shfl_xor_sync(full, S, o);
    }
    if (lane == 0) loss_b[b] = -part;
    const float w = w_is[b];
    const float d0 = w * (sh.soft[i0] * S - m0), d1 = w * (sh.soft[i1] * S - m1);
    const __half h0 = __float2half_rn(i0 < n_atoms ? d0 : 0.f), h1 = __float2half_rn(i1 < n_atoms ? d1 : 0.f);
    const long long col = (long long)a * 64;
    if (i0 < n_atoms) {
      dz[(long long)b * ldl + col + i0] = h0;
      dzt[(col + i0) * ld_dzt + b] = h0;
    }
    if (i1 < n_atoms) {
      dz[(long long)b * ldl + col + i1] = h1;
      dzt[(col + i1) * ld_dzt + b] = h1;
    }
    gsm[i0] = __half2float(h0);
    gsm[i1] = __half2float(h1);
  }
}

// HEAD_BW == false: one warp per sample (grid = ceil(B / warps per CTA)).
// HEAD_BW == true: one CTA per sample that also computes the previous layer's dZ row.  dL/dlogits of a
// sample is n_atoms values of ONE action, so that row is n_atoms consecutive rows of the head's effective
// weights [N, K] combined -- not a dense [B, N] x [N, K] product (that GEMM was 15 us of every update's
// critical path: 20 k-blocks for 51 non-zero columns per row).  The rows depend only on the action taken:
// one bulk copy (cp.async.bulk, 51 KB) brings them to shared memory while warp 0 computes the loss.
// The fp16-rounded dL/dlogits are used, as the GEMM would read them; fp32 accumulation.
template <bool HEAD_BW>
__global__ void c51_loss_kernel(const float* __restrict__ logits_tm1, const float* __restrict__ q_sel,
                                int ldq, const uint32_t* __restrict__ legal_t, int legal_words,
                                const float* __restrict__ tlogits, int ldl,
                                const int32_t* __restrict__ a_tm1, const float* __restrict__ r_t,
                                const uint8_t* __restrict__ term_t, int mask_terminal,
                                const float* __restrict__ w_is, int B, int A, int n_atoms, float vmax,
                                float gamma, const float* __restrict__ dyn, __half* __restrict__ dz,
                                __half* __restrict__ dzt, int ld_dzt, float* __restrict__ loss_b,
                                HeadBackward hb, int early) {
  extern __shared__ __align__(128) unsigned char hb_smem[];
  tc::pdl_launch_dependents();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = HEAD_BW ? (int)blockIdx.x : (int)(blockIdx.x * (blockDim.x >> 5)) + warp;
  // early != 0: the launch in front of this one released its dependents only after ITS dependencies were
  // complete (tc_chain.cuh), so the sampled actions are visible here, in front of the wait, and the weight
  // rows (written by the previous update) can be on their way while the forward pass still runs
  int a = 0;
  if (HEAD_BW && early && b < B) {
    a = a_tm1[b];
    if (threadIdx.x == 32) {
      const uint32_t bytes = (uint32_t)n_atoms * (uint32_t)hb.K * 2u;
      uint64_t* bar = reinterpret_cast<uint64_t*>(hb_smem + ((bytes + 127u) & ~127u) + 64 * 4);
      tc::mbar_init(bar, 1);
      tc::fence_barrier_init();
      tc::mbar_expect_tx(bar, bytes);
      const __half* src = hb.w + (size_t)a * 64 * hb.K;
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   :: "r"(tc::smem_u32(hb_smem)), "l"(src), "r"(bytes), "r"(tc::smem_u32(bar)) : "memory");
    }
  }
#define HB_LOSS_STAMP(k) do { if (HEAD_BW && hb.trace != nullptr && (threadIdx.x & 31) == 0) \
    hb.trace[(size_t)blockIdx.x * 8 + (k)] = tc::globaltimer_ns(); } while (0)
  tc::pdl_wait();
  if (threadIdx.x == 0) HB_LOSS_STAMP(0);
  if (dyn != nullptr) gamma = dyn[1];            // the step's discount, staged by the sample kernel
  if (b >= B) return;
  if (!(HEAD_BW && early)) a = a_tm1[b];
  __half h0, h1;
  if constexpr (!HEAD_BW) {
    c51_row(b, lane, logits_tm1, q_sel, ldq, legal_t, legal_words, tlogits, ldl, a, r_t, term_t, mask_terminal,
            w_is, A, n_atoms, vmax, gamma, dz, dzt, ld_dzt, loss_b, h0, h1);
  } else {
    const uint32_t bytes = (uint32_t)n_atoms * (uint32_t)hb.K * 2u;
    const __half* wsm = reinterpret_cast<const __half*>(hb_smem);
    float* gsm = reinterpret_cast<float*>(hb_smem + ((bytes + 127u) & ~127u));
    uint64_t* bar = reinterpret_cast<uint64_t*>(gsm + 64);
    if (threadIdx.x == 32 && !early) {           // warp 1: warp 0 goes straight to the loss
      tc::mbar_init(bar, 1);
      tc::fence_barrier_init();
      tc::mbar_expect_tx(bar, bytes);
      const __half* src = hb.w + (size_t)a * 64 * hb.K;
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   :: "r"(tc::smem_u32(hb_smem)), "l"(src), "r"(bytes), "r"(tc::smem_u32(bar)) : "memory");
    }
    __shared__ C51Shared c51sh;
    c51_row_cta(c51sh, b, warp, lane, logits_tm1, q_sel, ldq, legal_t, legal_words, tlogits, ldl, a, r_t, term_t,
                mask_terminal, w_is, A, n_atoms, vmax, gamma, dz, dzt, ld_dzt, loss_b, gsm);
    if (warp == 0) HB_LOSS_STAMP(1);
    __syncthreads();
    tc::mbar_wait(bar, 0);
    if (threadIdx.x == 0) HB_LOSS_STAMP(2);
    for (int k4 = threadIdx.x; k4 < (hb.K >> 2); k4 += blockDim.x) {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      const __half* col = wsm + 4 * k4;
#pragma unroll 4
      for (int i = 0; i < n_atoms; ++i) {
        const uint2 v = *reinterpret_cast<const uint2*>(col + (size_t)i * hb.K);
        const float gi = gsm[i];
        const float2 f0 = __half22float2(*reinterpret_cast<const __half2*>(&v.x));
        const float2 f1 = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
        acc[0] = fmaf(gi, f0.x, acc[0]);
        acc[1] = fmaf(gi, f0.y, acc[1]);
        acc[2] = fmaf(gi, f1.x, acc[2]);
        acc[3] = fmaf(gi, f1.y, acc[3]);
      }
      const uint2 gv = *reinterpret_cast<const uint2*>(hb.gate + (size_t)b * hb.K + 4 * k4);   // ReLU backward
      const float2 g0 = __half22float2(*reinterpret_cast<const __half2*>(&gv.x));
      const float2 g1 = __half22float2(*reinterpret_cast<const __half2*>(&gv.y));
      if (!(g0.x > 0.f)) acc[0] = 0.f;
      if (!(g0.y > 0.f)) acc[1] = 0.f;
      if (!(g1.x > 0.f)) acc[2] = 0.f;
      if (!(g1.y > 0.f)) acc[3] = 0.f;
      const __half2 o0 = __floats2half2_rn(acc[0], acc[1]), o1 = __floats2half2_rn(acc[2], acc[3]);
      uint2 o;
      o.x = *reinterpret_cast<const uint32_t*>(&o0);
      o.y = *reinterpret_cast<const uint32_t*>(&o1);
      *reinterpret_cast<uint2*>(hb.dz_prev + (size_t)b * hb.K + 4 * k4) = o;
      hb.dzt_prev[(size_t)(4 * k4) * ld_dzt + b] = __low2half(o0);
      hb.dzt_prev[(size_t)(4 * k4 + 1) * ld_dzt + b] = __high2half(o0);
      hb.dzt_prev[(size_t)(4 * k4 + 2) * ld_dzt + b] = __low2half(o1);
      hb.dzt_prev[(size_t)(4 * k4 + 3) * ld_dzt + b] = __high2half(o1);
    }
    if (threadIdx.x == 0) HB_LOSS_STAMP(3);
#undef HB_LOSS_STAMP
  }
}

// rlax.double_q_learning + l2 loss (rlax_dqn.py:170-205): td = r + gamma * q_t[argmax q_sel] - q_tm1[a],
// q_t zeroed on terminal; L = mean(w * 0.5 td^2)  ->  dz[b, a] = -w_b td_b.  One thread per sample.
__global__ void dqn_loss_kernel(const float* __restrict__ q_on, int ldq, const float* __restrict__ q_t,
                                const int32_t* __restrict__ a_tm1, const float* __restrict__ r_t,
                                const uint8_t* __restrict__ term_t, const float* __restrict__ w_is,
                                int B, int A, float gamma, const float* __restrict__ dyn,
                                __half* __restrict__ dz, int ld_dz,
                                __half* __restrict__ dzt, int ld_dzt, float* __restrict__ td_out) {
  tc::pdl_launch_dependents();
  tc::pdl_wait();
  if (dyn != nullptr) gamma = dyn[1];
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const float* qs = q_on + (long long)(B + b) * ldq;     // online(obs_t)
  int astar = 0;
  float best = qs[0];
  for (int a = 1; a < A; ++a)
    if (qs[a] > best) { best = qs[a]; astar = a; }
  const bool term = term_t != nullptr && term_t[b];
  const float qt = term ? 0.f : q_t[(long long)b * ldq + astar];
  const int a = a_tm1[b];
  const float td = r_t[b] + gamma * qt - q_on[(long long)b * ldq + a];
  td_out[b] = td;
  const __half d = __float2half_rn(-w_is[b] * td);
  dz[(long long)b * ld_dz + a] = d;
  dzt[(long long)a * ld_dzt + b] = d;
}

// mean(loss_b * w) in a fixed order (one block, tree over fixed strides)
__global__ void mean_loss_kernel(const float* __restrict__ loss_b, const float* __restrict__ w, int B,
                                 int squared_half, float* __restrict__ out) {
  __shared__ float red[256];
  float s = 0.f;
  for (int i = threadIdx.x; i < B; i += blockDim.x) {
    const float l = squared_half ? 0.5f * loss_b[i] * loss_b[i] : loss_b[i];
    s += l * w[i];
  }
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = blockDim.x >> 1; o; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = red[0] / (float)B;
}

// gb[n] = alpha * sum_b zt[n, b]   (one warp per output unit, fixed order)
__global__ void bias_grad_kernel(const __half* __restrict__ zt, int ld, int N, int B, float alpha,
                                 float* __restrict__ gb) {
  tc::pdl_wait();              // launched programmatically behind the dZ GEMM (first layer): resident early
  const int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (n >= N) return;
  float s = 0.f;
  for (int b = lane; b < B; b += 32) s += __half2float(zt[(long long)n * ld + b]);
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) gb[n] = alpha * s;
}

// Fused Adam (optix.adam: m, v, bias correction, theta -= lr * mhat / (sqrt(vhat) + eps)) over the
// flat effective-parameter space + refresh of the fp16 effective weights in both layouts.
// do_update == 0: refresh only (after parameters were loaded).
// [i4_begin, i4_end): the range of 4-element groups to update (everything, or one layer).
template <int KINDS>
__global__ void adam_refresh_kernel(LayerTable tab, long long n_eff, long long i4_begin, long long i4_end,
                                    float* __restrict__ master,
                                    const float* __restrict__ eps_noise, float* __restrict__ am,
                                    float* __restrict__ av, const float* __restrict__ grad,
                                    float grad_scale, int do_update, float lr, float b1, float b2,
                                    float eps, const int* __restrict__ step_ptr, __half* __restrict__ w16,
                                    __half* __restrict__ /*w16t*/, float* __restrict__ beff, int tf_style) {
  // bias corrections of this step, computed once by adam_tick_kernel from the device-side step counter (so
  // that the launch can sit in a CUDA graph): [1] 1 - b1^t, [2] 1 - b2^t, [3] [4] their reciprocals,
  // [5] lr sqrt(1 - b2^t) / (1 - b1^t).  (Two powf and 24 IEEE divisions per thread were most of this kernel.)
  float r1 = 1.f, r2 = 1.f, lr_tf = 0.f;
  if (do_update) {
    // the step's hyper-parameters as the tick kernel left them ([6..10]: they may have come from the
    // learner's staging slot rather than from this launch's arguments, which a replayed graph freezes)
    const float* sc = reinterpret_cast<const float*>(step_ptr);
    r1 = sc[3]; r2 = sc[4]; lr_tf = sc[5];
    lr = sc[6]; b1 = sc[7]; b2 = sc[8]; eps = sc[9]; grad_scale = sc[10];
  }
  // four consecutive parameters per thread (every block boundary is a multiple of 8 elements)
  for (long long i4 = i4_begin + (long long)blockIdx.x * blockDim.x + threadIdx.x; i4 < i4_end;
       i4 += (long long)gridDim.x * blockDim.x) {
    const long long i = i4 << 2;
    int l = 0;
#pragma unroll
    for (int k = 1; k < kMaxLayers; ++k)
      if (k < tab.n_layers && i >= tab.l[k].eoff) l = k;
    const LayerGeo g = tab.l[l];
    const long long loc = i - g.eoff;
    float noise[4] = {0.f, 0.f, 0.f, 0.f};
    if (KINDS == 3) {
      const float4 t = *reinterpret_cast<const float4*>(eps_noise + i);
      noise[0] = t.x; noise[1] = t.y; noise[2] = t.z; noise[3] = t.w;
    }
    float gr4[4] = {0.f, 0.f, 0.f, 0.f};
    if (do_update) {
      const float4 t = *reinterpret_cast<const float4*>(grad + i);
      gr4[0] = t.x; gr4[1] = t.y; gr4[2] = t.z; gr4[3] = t.w;
    }
    // NoisyLinear's plain weight and its w_mu (kinds 0 and 1) enter the output as a sum (noisy_mlp.py:62,80):
    // their gradients are the same number, and from optix.adam's zero state so are their moments and their
    // steps.  One moment pair serves both (state slot 0; w_sigma's is slot 1): 16 of the 82 bytes this
    // kernel moved per parameter.
    constexpr int STATES = KINDS == 3 ? 2 : 1;
    float p[KINDS][4];
#pragma unroll
    for (int k = 0; k < KINDS; ++k) {
      const float4 pv = *reinterpret_cast<const float4*>(master + k * n_eff + i);
      p[k][0] = pv.x; p[k][1] = pv.y; p[k][2] = pv.z; p[k][3] = pv.w;
    }
    if (do_update) {
#pragma unroll
      for (int st = 0; st < STATES; ++st) {
        const float4 mv = *reinterpret_cast<const float4*>(am + st * n_eff + i);
        const float4 vv = *reinterpret_cast<const float4*>(av + st * n_eff + i);
        float m[4] = {mv.x, mv.y, mv.z, mv.w}, v[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float gr = gr4[j] * grad_scale * (st == 1 ? noise[j] : 1.0f);
          m[j] = b1 * m[j] + (1.0f - b1) * gr;
          v[j] = b2 * v[j] + (1.0f - b2) * gr * gr;
          float step;
          if (tf_style) {
            // tf.train.AdamOptimizer (rainbow_agent.py:228-230): lr_t = lr sqrt(1-b2^t)/(1-b1^t),
            // theta -= lr_t m / (sqrt(v) + eps) -- eps is NOT bias corrected
            step = lr_tf * m[j] / (sqrtf(v[j]) + eps);
          } else {
            const float mhat = m[j] * r1, vhat = v[j] * r2;
            step = lr * mhat / (sqrtf(vhat) + eps);
          }
          if (st == 0) {
            p[0][j] = p[0][j] - step;
            if (KINDS == 3) p[1][j] = p[1][j] - step;
          } else {
            p[KINDS - 1][j] = p[KINDS - 1][j] - step;
          }
        }
        *reinterpret_cast<float4*>(am + st * n_eff + i) = make_float4(m[0], m[1], m[2], m[3]);
        *reinterpret_cast<float4*>(av + st * n_eff + i) = make_float4(v[0], v[1], v[2], v[3]);
      }
#pragma unroll
      for (int k = 0; k < KINDS; ++k)
        *reinterpret_cast<float4*>(master + k * n_eff + i) = make_float4(p[k][0], p[k][1], p[k][2], p[k][3]);
    }
    float eff[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < KINDS; ++k)
#pragma unroll
      for (int j = 0; j < 4; ++j) eff[j] += (k == 2) ? p[k][j] * noise[j] : p[k][j];
    if (loc < (long long)g.N * g.K) {   // transposes: transpose16_kernel (coalesced)
      const __half2 h0 = __floats2half2_rn(eff[0], eff[1]), h1 = __floats2half2_rn(eff[2], eff[3]);
      uint2 o;
      o.x = *reinterpret_cast<const uint32_t*>(&h0);
      o.y = *reinterpret_cast<const uint32_t*>(&h1);
      *reinterpret_cast<uint2*>(w16 + g.woff + loc) = o;
    } else {
      *reinterpret_cast<float4*>(beff + g.boff + (loc - (long long)g.N * g.K)) =
          make_float4(eff[0], eff[1], eff[2], eff[3]);
    }
  }
}

// out[c, r] = in[r, c] for r < rows, c < cols (activations for the weight-gradient GEMMs): 32 x 32 tiles
// through shared memory, grid (ceil(cols / 32), ceil(rows / 32)), 256 threads
// (all hidden layers in one launch: blockIdx.z = layer -- as one launch per layer the second one ended 2 us
// after the loss and the backward tail waited for it)
struct TransposeActs {
  const __half* in[kMaxLayers];
  __half* out[kMaxLayers];
  int cols[kMaxLayers];
};
__global__ void transpose_act_kernel(TransposeActs ta, int rows, int ld_out) {
  __shared__ __half tile[32][33];
  const __half* __restrict__ in = ta.in[blockIdx.z];
  __half* __restrict__ out = ta.out[blockIdx.z];
  const int cols = ta.cols[blockIdx.z], ld_in = cols;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  if (c0 >= cols) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int i = ty; i < 32; i += 8) {
    const int r = r0 + i, c = c0 + tx;
    tile[i][tx] = (r < rows && c < cols) ? in[(size_t)r * ld_in + c] : __float2half(0.f);
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    const int c = c0 + i, r = r0 + tx;
    if (r < rows && c < cols) out[(size_t)c * ld_out + r] = tile[tx][i];
  }
}

// w16t[k, n] = w16[n, k] for every layer: 32 x 32 tiles through shared memory so that both the
// reads and the writes are coalesced (the v1 Adam kernel scattered 2-byte stores: 38 us per step).
__global__ void transpose16_kernel(LayerTable tab, const __half* __restrict__ w16, __half* __restrict__ w16t) {
  __shared__ __half tile[32][33];
  int t = blockIdx.x, l = 0;
  for (; l < tab.n_layers; ++l) {
    const int nt = ((tab.l[l].N + 31) / 32) * ((tab.l[l].K + 31) / 32);
    if (t < nt) break;
    t -= nt;
  }
  if (l >= tab.n_layers) return;
  const LayerGeo g = tab.l[l];
  const int tiles_k = (g.K + 31) / 32;
  const int n0 = (t / tiles_k) * 32, k0 = (t % tiles_k) * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;       // 32 x 8 threads
  for (int r = ty; r < 32; r += 8) {
    const int n = n0 + r, k = k0 + tx;
    tile[r][tx] = (n < g.N && k < g.K) ? w16[g.woff + (long long)n * g.K + k] : __float2half(0.f);
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int k = k0 + r, n = n0 + tx;
    if (n < g.N && k < g.K) w16t[g.woff + (long long)k * g.N + n] = tile[tx][r];
  }
}

// step counter and this step's bias corrections: step[0] = t, then as floats [1] 1 - b1^t, [2] 1 - b2^t,
// [3] [4] their reciprocals, [5] lr sqrt(1 - b2^t) / (1 - b1^t)  (tf.train.AdamOptimizer's lr_t)
// [6..10] lr, b1, b2, eps, grad_scale of this step.  dyn (optional; device or pinned host memory):
// [2] lr [3] b1 [4] b2 [5] eps [6] grad_scale replace the by-value arguments -- a replayed CUDA graph
// freezes its kernel arguments, the learner's schedules (learning rate, ...) must not force a re-capture.
__global__ void adam_tick_kernel(int* step, float b1, float b2, float lr, float eps, float grad_scale,
                                 const float* __restrict__ dyn) {
  if (dyn != nullptr) { lr = dyn[2]; b1 = dyn[3]; b2 = dyn[4]; eps = dyn[5]; grad_scale = dyn[6]; }
  const int t = *step + 1;
  *step = t;
  float* sc = reinterpret_cast<float*>(step);
  const float bc1 = 1.0f - powf(b1, (float)t), bc2 = 1.0f - powf(b2, (float)t);
  sc[1] = bc1; sc[2] = bc2; sc[3] = 1.0f / bc1; sc[4] = 1.0f / bc2; sc[5] = lr * sqrtf(bc2) / bc1;
  sc[6] = lr; sc[7] = b1; sc[8] = b2; sc[9] = eps; sc[10] = grad_scale;
}
__global__ void fill_kernel(float* p, float v, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// ---------------------------------------------------------------- host helpers ----
template <class T>
int dev_alloc(T** p, size_t n) {
  *p = nullptr;
  HB_CUDA(cudaMalloc((void**)p, n * sizeof(T)));
  HB_CUDA(cudaMemset(*p, 0, n * sizeof(T)));
  return HB_OK;
}
template <class T>
void dev_free(T*& p) {
  if (p) cudaFree(p);
  p = nullptr;
}

void free_learner_ws(hb_qnet* q) {
  dev_free(q->x); dev_free(q->xt); dev_free(q->xb);
  for (int l = 0; l < kMaxLayers; ++l) {
    dev_free(q->h[l]); dev_free(q->ht[l]); dev_free(q->th[l]); dev_free(q->dz[l]); dev_free(q->dzt[l]);
  }
  dev_free(q->logits); dev_free(q->q_on); dev_free(q->q_tg); dev_free(q->tlogits); dev_free(q->w_is);
  dev_free(q->loss_b);
  dev_free(q->bias_part);
  q->lb = q->lbp = 0;
}

int ensure_learner_ws(hb_qnet* q, int B) {
  if (q->lb >= B) return HB_OK;
  HB_CUDA(cudaDeviceSynchronize());
  free_learner_ws(q);
  const int Bp = ceil_to(B, 8);
  const int K0 = q->tab.l[0].K;
  const int nl = q->n_layers;
  const int NL = q->tab.l[nl - 1].N;
  int rc;
  if ((rc = dev_alloc(&q->x, (size_t)2 * B * K0))) return rc;
  if ((rc = dev_alloc(&q->xb, (size_t)2 * B * q->obs_stride))) return rc;
  if (q->side == nullptr) {
    // Priorities: the target network's forward pass (side) and the first layer's gradient + update
    // (lstream[0]) are on the critical path of an update; the other layers' optix.adam kernels are
    // bandwidth-bound grids of hundreds of CTAs that would otherwise hold the SMs' registers while a
    // critical-path GEMM waits to be placed (seen in the CUPTI timeline: 13 us).  Captured kernel nodes
    // keep their stream's priority.
    int prio_lo = 0, prio_hi = 0;
    HB_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    HB_CUDA(cudaStreamCreateWithPriority(&q->side, cudaStreamNonBlocking, prio_hi));
    HB_CUDA(cudaStreamCreateWithPriority(&q->side2, cudaStreamNonBlocking, prio_lo));
    HB_CUDA(cudaEventCreateWithFlags(&q->ev_join2, cudaEventDisableTiming));
    HB_CUDA(cudaEventCreateWithFlags(&q->ev_fork, cudaEventDisableTiming));
    HB_CUDA(cudaEventCreateWithFlags(&q->ev_join, cudaEventDisableTiming));
    HB_CUDA(cudaEventCreateWithFlags(&q->ev_loss, cudaEventDisableTiming));
    HB_CUDA(cudaEventCreateWithFlags(&q->ev_wt, cudaEventDisableTiming));
    HB_CUDA(cudaEventCreateWithFlags(&q->ev_fwd, cudaEventDisableTiming));
    for (int l = 0; l < kMaxLayers; ++l) HB_CUDA(cudaEventCreateWithFlags(&q->ev_dz[l], cudaEventDisableTiming));
    for (int l = 0; l < kMaxLayers; ++l) {
      HB_CUDA(cudaStreamCreateWithPriority(&q->lstream[l], cudaStreamNonBlocking, l == 0 ? prio_hi : prio_lo));
      HB_CUDA(cudaEventCreateWithFlags(&q->ev_bias[l], cudaEventDisableTiming));
      HB_CUDA(cudaEventCreateWithFlags(&q->ev_layer[l], cudaEventDisableTiming));
    }
    HB_CUDA(cudaEventCreateWithFlags(&q->ev_tick, cudaEventDisableTiming));
  }
  if ((rc = dev_alloc(&q->xt, (size_t)K0 * Bp))) return rc;
  for (int l = 0; l < nl; ++l) {
    const int N = q->tab.l[l].N;
    if (l < nl - 1) {
      if ((rc = dev_alloc(&q->h[l], (size_t)2 * B * N))) return rc;
      if ((rc = dev_alloc(&q->ht[l], (size_t)N * 2 * Bp))) return rc;
      if ((rc = dev_alloc(&q->th[l], (size_t)B * N))) return rc;
    }
    if ((rc = dev_alloc(&q->dz[l], (size_t)B * N))) return rc;
    if ((rc = dev_alloc(&q->dzt[l], (size_t)N * Bp))) return rc;
  }
  if ((rc = dev_alloc(&q->logits, (size_t)2 * B * NL))) return rc;
  if ((rc = dev_alloc(&q->q_on, (size_t)2 * B * NL))) return rc;
  if ((rc = dev_alloc(&q->tlogits, (size_t)B * NL))) return rc;
  if ((rc = dev_alloc(&q->q_tg, (size_t)B * NL))) return rc;
  if ((rc = dev_alloc(&q->w_is, (size_t)B))) return rc;
  if ((rc = dev_alloc(&q->loss_b, (size_t)B))) return rc;
  if ((rc = dev_alloc(&q->bias_part, (size_t)((B + 63) / 64) * q->tab.l[0].N))) return rc;
  q->lb = B;
  q->lbp = Bp;
  return HB_OK;
}

int ensure_act_ws(hb_qnet* q, long long rows) {
  if (q->act_rows >= rows) return HB_OK;
  HB_CUDA(cudaDeviceSynchronize());
  dev_free(q->act_x); dev_free(q->act_h[0]); dev_free(q->act_h[1]); dev_free(q->act_q);
  q->act_rows = 0;
  int nmax = 8;
  for (int l = 0; l + 1 < q->n_layers; ++l) nmax = nmax > q->tab.l[l].N ? nmax : q->tab.l[l].N;
  const int NL = q->tab.l[q->n_layers - 1].N;
  int rc;
  if ((rc = dev_alloc(&q->act_x, (size_t)rows * q->tab.l[0].K))) return rc;
  if ((rc = dev_alloc(&q->act_h[0], (size_t)rows * nmax))) return rc;
  if ((rc = dev_alloc(&q->act_h[1], (size_t)rows * nmax))) return rc;
  if ((rc = dev_alloc(&q->act_q, (size_t)rows * (q->n_atoms > 1 ? q->A : NL)))) return rc;
  q->act_rows = rows;
  return HB_OK;
}

int launch_unpack(const uint32_t* packed, int stride, long long rows, __half* out, int ld,
                  __half* out_t, int ld_t, long long rows_t, cudaStream_t s) {
  const long long total = rows * stride;
  const unsigned blocks = (unsigned)((total + 255) / 256 < 148 * 32 ? (total + 255) / 256 : 148 * 32);
  unpack_rows_kernel<<<blocks ? blocks : 1, 256, 0, s>>>(packed, stride, rows, out, ld, out_t, ld_t, rows_t);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

// Forward pass of `rows` rows already unpacked in x0 [rows, K0].  weights: w16 (online) or t16.
// hidden activations go to hbuf[l] (ld = N_l) with optional transposed copies; the head writes
// q (and logits when given).
int forward(hb_qnet* net, const __half* wts, const float* bias, const __half* x0, long long rows,
            __half* const* hbuf, __half* const* hbuf_t, int ld_t, float* logits, float* q_out, int ldq,
            cudaStream_t s, const uint32_t* x_bits = nullptr, int pdl = 0, int max_ctas = 0) {
  const __half* in = x0;
  int ld_in = net->tab.l[0].K;
  const int nl = net->n_layers;
  for (int l = 0; l < nl; ++l) {
    const LayerGeo& g = net->tab.l[l];
    tc::EpiParams ep{};
    ep.alpha = 1.f;
    ep.bias = bias + g.boff;
    ep.pdl = pdl;
    ep.max_ctas = max_ctas;
    int rc;
    if (l < nl - 1) {
      ep.relu = 1;
      ep.out16 = hbuf[l];
      ep.ld_out = g.N;
      ep.out16_t = hbuf_t ? hbuf_t[l] : nullptr;
      ep.ld_out_t = ld_t;
      if (l == 0 && x_bits != nullptr)   // bits expanded inside the kernel: no fp16 copy of the observations
        rc = tc::gemm_bits<tc::EPI_F16>(x_bits, net->obs_stride, wts + g.woff, g.K, (int)rows, g.N, g.K, ep, s);
      else
        rc = tc::gemm<tc::EPI_F16>(in, ld_in, wts + g.woff, g.K, (int)rows, g.N, g.K, ep, s);
      in = hbuf[l];
      ld_in = g.N;
    } else if (net->n_atoms > 1) {
      ep.out32 = logits;
      ep.ld_out = g.N;
      ep.q = q_out;
      ep.ld_q = ldq;
      ep.n_atoms = net->n_atoms;
      ep.z0 = -net->vmax;
      ep.dz = 2.f * net->vmax / (float)(net->n_atoms - 1);
      ep.q_scale = 1.0f / (float)net->n_atoms;     // jnp.mean(probs * atoms)  rlax_rainbow.py:118
      rc = tc::gemm<tc::EPI_C51>(in, ld_in, wts + g.woff, g.K, (int)rows, g.N, g.K, ep, s);
    } else {
      ep.out32 = q_out;
      ep.ld_out = ldq;
      rc = tc::gemm<tc::EPI_F32>(in, ld_in, wts + g.woff, g.K, (int)rows, g.N, g.K, ep, s);
    }
    if (rc) return rc;
  }
  return HB_OK;
}

// Launch of tc::mlp_chain_kernel (+ HB_CHAIN_TRACE: per-job phase times of an eager launch on stderr).
int launch_chain(hb_qnet* q, tc::ChainParams& P, int tiles, cudaStream_t s, int pdl, const char* tag) {
  static const int compact = getenv("HB_CHAIN_COMPACT") ? atoi(getenv("HB_CHAIN_COMPACT")) : 1;
  P.compact = compact;
  static bool configured[64] = {false};
  const int dev = q->device;
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    HB_CUDA(cudaFuncSetAttribute(tc::mlp_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 tc::smem_bytes_for(tc::kChainBN)));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl ? 1 : 0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)tiles);
  cfg.blockDim = dim3(tc::kThreads);
  cfg.dynamicSmemBytes = tc::smem_bytes_for(tc::kChainBN);
  cfg.stream = s;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  // HB_CHAIN_TRACE=1 (debugging, eager launches only): per-job phase times of the launch on stderr
  static const bool trace = getenv("HB_CHAIN_TRACE") && atoi(getenv("HB_CHAIN_TRACE")) != 0;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (trace) cudaStreamIsCapturing(s, &cap);
  static unsigned long long* d_trace = nullptr;
  if (trace && cap == cudaStreamCaptureStatusNone) {
    if (d_trace == nullptr) HB_CUDA(cudaMalloc(&d_trace, 256 * 16 * 8));
    HB_CUDA(cudaMemsetAsync(d_trace, 0, 256 * 16 * 8, s));
    P.trace = d_trace;
  }
  HB_CUDA(cudaLaunchKernelEx(&cfg, tc::mlp_chain_kernel, P));
  if (P.trace != nullptr) {
    std::vector<unsigned long long> h(256 * 16);
    HB_CUDA(cudaStreamSynchronize(s));
    HB_CUDA(cudaMemcpy(h.data(), d_trace, 256 * 16 * 8, cudaMemcpyDeviceToHost));
    unsigned long long t0 = ~0ull;
    for (int c = 0; c < tiles; ++c) if (h[c * 16] && h[c * 16] < t0) t0 = h[c * 16];
    for (int jn = 0; jn < P.n_jobs; ++jn) {
      const int c0 = P.job[jn].tile_base, c1 = jn + 1 < P.n_jobs ? P.job[jn + 1].tile_base : tiles;
      double mx[6] = {0, 0, 0, 0, 0, 0}, mn[6] = {1e30, 1e30, 1e30, 1e30, 1e30, 1e30};
      for (int c = c0; c < c1; ++c)
        for (int k = 0; k < 6; ++k) {
          const double v = ((double)h[c * 16 + k] - (double)t0) * 1e-3;
          if (v > mx[k]) mx[k] = v;
          if (v < mn[k]) mn[k] = v;
        }
      for (int c = c0; c < c1; c += (c1 - c0 > 20 ? 20 : c1 - c0)) {   // first CTA (and every 20th) of the job
        fprintf(stderr, "  [cta %d: %.0f cycles in %.0f ns between 'acc' and 'stores done']\n", c,
                (double)(h[c * 16 + 7] - h[c * 16 + 6]), (double)(h[c * 16 + 5] - h[c * 16 + 3]));
        if (h[c * 16 + 8] && P.job[jn].abits)
          fprintf(stderr, "  [expand kb 0 cycles: wait slot %lld | expand+sts %lld | proxy fence %lld | arrive %lld]\n",
                  (long long)(h[c * 16 + 9] - h[c * 16 + 8]), (long long)(h[c * 16 + 10] - h[c * 16 + 9]),
                  (long long)(h[c * 16 + 11] - h[c * 16 + 10]), (long long)(h[c * 16 + 12] - h[c * 16 + 11]));
        else if (h[c * 16 + 8])
          fprintf(stderr, "  [c51 epilogue cycles: tmem %lld | bias+add %lld | stage %lld | row stores %lld | q %lld]\n",
                  (long long)(h[c * 16 + 9] - h[c * 16 + 8]), (long long)(h[c * 16 + 10] - h[c * 16 + 9]),
                  (long long)(h[c * 16 + 11] - h[c * 16 + 10]), (long long)(h[c * 16 + 12] - h[c * 16 + 11]),
                  (long long)(h[c * 16 + 13] - h[c * 16 + 12]));
      }
      if (atoi(getenv("HB_CHAIN_TRACE")) > 1)
        for (int c = c0; c < c1; ++c)
          fprintf(stderr, "    cta %d (m-block %d, n-tile %d): rows ready %.1f acc %.1f stores issued %.1f done %.1f\n", c,
                  (c - c0) / P.job[jn].tiles_n, (c - c0) % P.job[jn].tiles_n, ((double)h[c * 16 + 1] - (double)t0) * 1e-3,
                  ((double)h[c * 16 + 3] - (double)t0) * 1e-3, ((double)h[c * 16 + 4] - (double)t0) * 1e-3,
                  ((double)h[c * 16 + 5] - (double)t0) * 1e-3);
      fprintf(stderr, "%s job %d (%d tiles): start %.1f-%.1f | rows ready %.1f-%.1f | mma issued %.1f-%.1f | "
              "acc %.1f-%.1f | stores issued %.1f-%.1f | stores done %.1f-%.1f us\n", tag, jn, c1 - c0, mn[0], mx[0],
              mn[1], mx[1], mn[2], mx[2], mn[3], mx[3], mn[4], mx[4], mn[5], mx[5]);
    }
  }
  return HB_OK;
}

// The learner's two forward passes (online network on 2B rows, target network on B rows) as one launch
// of tc::mlp_chain_kernel.  Same operands and epilogues as forward() gives each layer.  Returns
// HB_OK with *done = 0 when the shapes do not qualify (the caller then runs forward() twice).
int forward_chain(hb_qnet* q, int B, const int32_t* d_a_tm1, int tf_variant, int ldq, cudaStream_t s, int pdl,
                  int* done) {
  *done = 0;
  const int nl = q->n_layers;
  static const bool off = getenv("HB_FWD_CHAIN") && atoi(getenv("HB_FWD_CHAIN")) == 0;
  if (off || nl < 2 || 2 * nl > tc::kChainMaxJobs) return HB_OK;
  tc::ChainParams P{};
  // Tile widths.  The target network's head writes every logit (the loss picks the action once the online
  // network's argmax is known): 64 KB per 128-wide tile at the ~45 GB/s one SM stores, so its tiles are 64
  // wide when all tiles of the launch still fit on the device.
  int bn_of[kMaxLayers][2];
  {
    static const bool narrow_off = getenv("HB_CHAIN_NARROW") && atoi(getenv("HB_CHAIN_NARROW")) == 0;
    long long total = 0;
    const long long tm_on = (2 * B + tc::BM - 1) / tc::BM, tm_tg = (B + tc::BM - 1) / tc::BM;
    for (int l = 0; l < nl; ++l) {
      bn_of[l][0] = bn_of[l][1] = 128;
      total += (tm_on + tm_tg) * ((q->tab.l[l].N + 127) / 128);
    }
    const int NH = q->tab.l[nl - 1].N;
    if (!narrow_off && q->n_atoms > 1 && total + tm_tg * ((NH + 63) / 64 - (NH + 127) / 128) <= tc::sm_count())
      bn_of[nl - 1][1] = 64;
  }
  int tiles = 0, flags = 0;
  const size_t row_words = (size_t)q->obs_stride;
  for (int l = 0; l < nl; ++l) {
    const LayerGeo& g = q->tab.l[l];
    for (int net = 0; net < 2; ++net) {
      const int rows = net == 0 ? 2 * B : B;
      const __half* wts = net == 0 ? q->w16 : q->t16;
      const float* bias = net == 0 ? q->beff : q->tbeff;
      __half* const* hbuf = net == 0 ? q->h : q->th;
      tc::ChainJob& J = P.job[2 * l + net];
      tc::EpiParams& ep = J.ep;
      ep = tc::EpiParams{};
      ep.M = rows;
      ep.N = g.N;
      ep.alpha = 1.f;
      ep.bias = bias + g.boff;
      J.K = g.K;
      J.bn = bn_of[l][net];
      J.tiles_n = (g.N + J.bn - 1) / J.bn;
      const int tiles_m = (rows + tc::BM - 1) / tc::BM;
      J.tile_base = tiles;
      tiles += tiles_m * J.tiles_n;
      J.flag_base = 2 + flags;
      flags += tiles_m;
      J.dep = l == 0 ? -1 : 2 * (l - 1) + net;
      J.dep_arrivals = l == 0 ? 0 : P.job[J.dep].tiles_n * (P.job[J.dep].bn == 64 ? tc::kEpiWarps / 2 : tc::kEpiWarps);
      J.has_consumer = l < nl - 1;
      int rc;
      if ((rc = tc::make_tmap(&J.tb, wts + g.woff, g.N, g.K, g.K, J.bn))) return rc;
      if (l == 0) {
        J.abits = 1;
        J.ta = J.tb;
        ep.a_bits = q->xb + (net == 0 ? 0 : (size_t)B * row_words);
        ep.a_stride = q->obs_stride;
        if (q->obs_stride % 2 != 0 || g.K > 32 * q->obs_stride) return HB_OK;
      } else {
        const int Kin = q->tab.l[l - 1].N;
        if ((rc = tc::make_tmap(&J.ta, hbuf[l - 1], rows, g.K, Kin, tc::BM))) return rc;
      }
      J.tout = J.tb;
      if (l < nl - 1) {
        J.epi = tc::EPI_F16;
        ep.relu = 1;
        ep.out16 = hbuf[l];
        ep.ld_out = g.N;
        // (no transposed copy here: the weight-gradient GEMMs want it much later, the caller makes it
        // beside the loss -- in the epilogue it was 1.5-2 us of every hidden layer)
        if ((rc = tc::out_map<tc::EPI_F16>(&J.tout, J.tb, ep))) return rc;
      } else if (q->n_atoms > 1) {
        J.epi = tc::EPI_C51;
        ep.out32 = net == 0 ? q->logits : q->tlogits;
        ep.ld_out = g.N;
        ep.q = net == 0 ? q->q_on : (tf_variant ? q->q_tg : nullptr);
        ep.ld_q = ldq;
        ep.n_atoms = q->n_atoms;
        ep.z0 = -q->vmax;
        ep.dz = 2.f * q->vmax / (float)(q->n_atoms - 1);
        ep.q_scale = 1.0f / (float)q->n_atoms;
        if (net == 0 && !tf_variant) {       // online head: logits of the action taken, q of the next state
          J.sel_action = d_a_tm1;
          J.sel_rows = B;
        }
      } else {
        J.epi = tc::EPI_F32;
        ep.out32 = net == 0 ? q->q_on : q->tlogits;
        ep.ld_out = ldq;
      }
    }
  }
  if (tiles > tc::sm_count() || 2 + flags > 128) return HB_OK;
  // The row-block counters only ever grow, by the same amount per launch -- as long as every launch has the
  // same row blocks.  Another batch size starts them over (stream ordered; a captured update never gets here
  // with a new batch without being captured again).
  if (q->chain_B != B) {
    HB_CUDA(cudaMemsetAsync(q->chain, 0, 128 * sizeof(unsigned int), s));
    q->chain_B = B;
  }
  P.n_jobs = 2 * nl;
  P.total = tiles;
  P.state = q->chain;
  P.flags = q->chain;                      // flag_base already counts from the start of the buffer
  int rc0;
  if ((rc0 = launch_chain(q, P, tiles, s, pdl, "chain"))) return rc0;
  *done = 1;
  return HB_OK;
}

// The tail of the backward pass as one launch of tc::mlp_chain_kernel: dZ_0^T [N_0, B] = W_1^T . dZ_1^T (ReLU
// gate read transposed, row sums left for the bias gradient) and, released row block by row block,
// dW_0 [N_0, K_0] = dZ_0^T . X^T / B, whose tiles of column block 0 finish db_0.  Replaces two GEMM launches
// and bias_grad_kernel on the critical path.  *done = 0: shapes do not qualify, nothing was launched.
int backward_tail_chain(hb_qnet* q, int B, int Bp, cudaStream_t s, int pdl, int* done) {
  *done = 0;
  static const bool off = getenv("HB_BWD_CHAIN") && atoi(getenv("HB_BWD_CHAIN")) == 0;
  if (off || q->n_layers < 2) return HB_OK;
  const LayerGeo& L0 = q->tab.l[0];
  const LayerGeo& L1 = q->tab.l[1];
  tc::ChainParams P{};
  int rc;
  tc::ChainJob& Z = P.job[0];
  Z.ep = tc::EpiParams{};
  Z.ep.M = L1.K;                       // = N_0
  Z.ep.N = B;
  Z.ep.alpha = 1.f;
  Z.ep.gate = q->ht[0];
  Z.ep.ld_gate = 2 * Bp;
  Z.ep.out16 = q->dzt[0];
  Z.ep.ld_out = Bp;
  Z.ep.row_part = q->bias_part;
  Z.ep.ld_part = L0.N;
  Z.K = L1.N;
  Z.epi = tc::EPI_F16;
  // 64-wide tiles for both jobs: one SM stores ~20 B / clock, and the epilogues (32 KB of fp16 + the gate per
  // dZ tile, 64 KB of fp32 per dW tile at 128 columns) were 4 us each of a 16 us launch
  static const int bn_env = getenv("HB_BWD_CHAIN_BN") ? atoi(getenv("HB_BWD_CHAIN_BN")) : 64;
  const int bn = bn_env == 128 ? 128 : 64;
  Z.bn = bn;
  Z.tiles_n = (B + bn - 1) / bn;
  Z.tile_base = 0;
  Z.dep = -1;
  Z.flag_base = 2;
  Z.has_consumer = 1;
  Z.b_late = 1;
  // Both operands that could be requested ahead of a job's wait (W_1^T, X^T) are written by THIS update's side
  // stream.  They are complete before this launch can start only if that dependency is an ordinary (full)
  // graph edge; the timelines say it is (the launch always starts when the side stream's last kernel ends),
  // the documentation does not promise it for a launch with the programmatic attribute -- and requesting
  // them early bought nothing measurable (72.0 vs 72.1 us).  HB_BWD_CHAIN_EARLY=1 requests them early.
  static const bool early = getenv("HB_BWD_CHAIN_EARLY") && atoi(getenv("HB_BWD_CHAIN_EARLY")) != 0;
  Z.no_early = early ? 0 : 1;
  if ((rc = tc::make_tmap(&Z.ta, q->w16t + L1.woff, L1.K, L1.N, L1.N, tc::BM))) return rc;
  if ((rc = tc::make_tmap(&Z.tb, q->dz[1], B, L1.N, L1.N, bn))) return rc;
  if ((rc = tc::out_map<tc::EPI_F16>(&Z.tout, Z.tb, Z.ep))) return rc;
  if (!Z.ep.tma_out) return HB_OK;     // the row sums are made on the tensor-store path only
  const int tiles_m = (L1.K + tc::BM - 1) / tc::BM;
  tc::ChainJob& W = P.job[1];
  W.ep = tc::EpiParams{};
  W.ep.M = L0.N;
  W.ep.N = L0.K;
  W.ep.alpha = 1.0f / (float)B;
  W.ep.out32 = q->grad + L0.eoff;
  W.ep.ld_out = L0.K;
  W.K = B;
  W.epi = tc::EPI_F32;
  W.bn = bn;
  W.tiles_n = (L0.K + bn - 1) / bn;
  W.tile_base = tiles_m * Z.tiles_n;
  W.dep = 0;
  W.dep_arrivals = Z.tiles_n * (bn == 64 ? tc::kEpiWarps / 2 : tc::kEpiWarps);
  W.flag_base = 2 + tiles_m;
  W.has_consumer = 0;
  W.no_early = Z.no_early;
  W.bias_in = q->bias_part;
  W.bias_parts = (B + 63) / 64;
  W.bias_ld = L0.N;
  W.bias_out = q->grad + L0.eoff + (long long)L0.N * L0.K;
  W.bias_alpha = 1.0f / (float)B;
  if ((rc = tc::make_tmap(&W.ta, q->dzt[0], L0.N, B, Bp, tc::BM))) return rc;
  if ((rc = tc::make_tmap(&W.tb, q->xt, L0.K, B, Bp, bn))) return rc;
  W.tout = W.tb;
  const int tiles = W.tile_base + tiles_m * W.tiles_n;
  if (tiles > tc::sm_count() || 2 + 2 * tiles_m > 128) return HB_OK;
  if (q->chain2_B != B) {
    HB_CUDA(cudaMemsetAsync(q->chain2, 0, 128 * sizeof(unsigned int), s));
    q->chain2_B = B;
  }
  P.n_jobs = 2;
  P.total = tiles;
  P.state = q->chain2;
  P.flags = q->chain2;
  if ((rc = launch_chain(q, P, tiles, s, pdl, "bwd chain"))) return rc;
  *done = 1;
  return HB_OK;
}

int launch_transposes(hb_qnet* net, cudaStream_t s) {
  int tiles = 0;
  for (int l = 0; l < net->n_layers; ++l)
    tiles += ((net->tab.l[l].N + 31) / 32) * ((net->tab.l[l].K + 31) / 32);
  transpose16_kernel<<<tiles, 256, 0, s>>>(net->tab, net->w16, net->w16t);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int adam_tick(hb_qnet* net, float lr, float b1, float b2, float eps, float grad_scale, const float* dyn,
              cudaStream_t s) {
  int* step_ptr = reinterpret_cast<int*>(net->scalars + 2);
  net->adam_step += 1;                       // host mirror (hb_net_info); the kernels use the device copy
  adam_tick_kernel<<<1, 1, 0, s>>>(step_ptr, b1, b2, lr, eps, grad_scale, dyn);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

// optix.adam / refresh over the 4-element groups [i4_begin, i4_end) of the flat parameter space
// ctas_per_sm > 0: a layer's update beside the backward pass is held to that many CTAs per SM (by asking
// for a matching share of the shared memory it never touches).  Left alone, the bandwidth-bound grid fills
// every SM's register file for its whole 10-15 us and a critical-path kernel that becomes ready in the
// meantime (the first layer's bias gradient, a gradient GEMM) waits for CTAs to retire whatever its priority.
int adam_range(hb_qnet* net, int do_update, long long i4_begin, long long i4_end, cudaStream_t s, int tf_style,
               int ctas_per_sm = 0) {
  int* step_ptr = reinterpret_cast<int*>(net->scalars + 2);
  const long long n4 = i4_end - i4_begin;
  if (n4 <= 0) return HB_OK;
  const unsigned blocks = (unsigned)((n4 + 255) / 256 < 148 * 16 ? (n4 + 255) / 256 : 148 * 16);
  size_t smem = 0;
  if (ctas_per_sm > 0) {
    smem = (size_t)(227 * 1024) / ctas_per_sm - 2048;
    static bool configured[64] = {false};
    const int dev = net->device;
    if (dev < 0 || dev >= 64 || !configured[dev]) {
      HB_CUDA(cudaFuncSetAttribute(adam_refresh_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
      HB_CUDA(cudaFuncSetAttribute(adam_refresh_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
      if (dev >= 0 && dev < 64) configured[dev] = true;
    }
  }
  if (net->kinds == 3)
    adam_refresh_kernel<3><<<blocks, 256, smem, s>>>(net->tab, net->n_eff, i4_begin, i4_end, net->master, net->eps,
                                                     net->adam_m, net->adam_v, net->grad, 1.f, do_update,
                                                     0.f, 0.f, 0.f, 0.f, step_ptr, net->w16, net->w16t, net->beff, tf_style);
  else
    adam_refresh_kernel<1><<<blocks, 256, smem, s>>>(net->tab, net->n_eff, i4_begin, i4_end, net->master, net->eps,
                                                     net->adam_m, net->adam_v, net->grad, 1.f, do_update,
                                                     0.f, 0.f, 0.f, 0.f, step_ptr, net->w16, net->w16t, net->beff, tf_style);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int run_adam(hb_qnet* net, int do_update, float lr, float b1, float b2, float eps, float grad_scale,
             cudaStream_t s, int tf_style = 0, const float* dyn = nullptr) {
  if (do_update) {
    int rc = adam_tick(net, lr, b1, b2, eps, grad_scale, dyn, s);
    if (rc) return rc;
  }
  return adam_range(net, do_update, 0, net->n_eff / 4, s, tf_style);   // w16t: refreshed by the next hb_net_grads
}

// maps (layer, kind, element of the reference-layout tensor) <-> flat effective index
// weights: reference [in, out] row-major, element (i, o); ours row n(o), column i.
inline int out_to_n(const hb_qnet* net, int layer, int o) {
  if (layer == net->n_layers - 1 && net->n_atoms > 1) return (o / net->n_atoms) * 64 + (o % net->n_atoms);
  return o;
}

}  // namespace

extern "C" {

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

int hb_net_create(int obs_len, int obs_stride_words, const int32_t* hidden, int n_hidden,
                  int num_actions, int n_atoms, float atom_vmax, int noisy, int device,
                  hb_qnet** out) {
  HB_REQUIRE(out != nullptr, "hb_net_create: out is NULL");
  HB_REQUIRE(obs_len >= 1 && obs_stride_words * 32 >= obs_len && obs_stride_words % 4 == 0,
             "hb_net_create: bad observation geometry (%d bits, %d words)", obs_len, obs_stride_words);
  HB_REQUIRE(n_hidden >= 0 && n_hidden + 1 <= kMaxLayers, "hb_net_create: 0..%d hidden layers", kMaxLayers - 1);
  HB_REQUIRE(num_actions >= 1 && num_actions <= 64, "hb_net_create: 1..64 actions");
  HB_REQUIRE(n_atoms >= 1 && n_atoms <= 64, "hb_net_create: 1..64 atoms");
  for (int l = 0; l < n_hidden; ++l)
    HB_REQUIRE(hidden[l] >= 8 && hidden[l] % 8 == 0, "hb_net_create: hidden sizes must be multiples of 8");
  hb::DeviceGuard g(device);
  if (!g.ok) { hb::set_error("hb_net_create: cannot select CUDA device %d", device); return HB_ERR_CUDA; }
  hb_qnet* q = new (std::nothrow) hb_qnet();
  HB_REQUIRE(q != nullptr, "hb_net_create: out of host memory");
  q->device = device;
  q->obs_len = obs_len;
  q->obs_stride = obs_stride_words;
  q->A = num_actions;
  q->n_atoms = n_atoms;
  q->noisy = noisy ? 1 : 0;
  q->kinds = noisy ? 3 : 1;
  q->vmax = atom_vmax;
  q->n_layers = n_hidden + 1;
  q->tab.n_layers = q->n_layers;
  long long eoff = 0, woff = 0, boff = 0;
  int in = obs_len, K = obs_stride_words * 32;
  for (int l = 0; l < q->n_layers; ++l) {
    LayerGeo& L = q->tab.l[l];
    const bool last = l == q->n_layers - 1;
    L.in = in;
    L.out = last ? num_actions * n_atoms : hidden[l];
    L.K = K;
    L.N = last ? (n_atoms > 1 ? num_actions * 64 : ceil_to(num_actions, 8)) : hidden[l];
    L.eoff = eoff; L.woff = woff; L.boff = boff;
    eoff += (long long)L.N * L.K + L.N;
    woff += (long long)L.N * L.K;
    boff += L.N;
    in = L.out;
    K = L.N;
  }
  q->n_eff = eoff; q->n_w = woff; q->n_b = boff;
  int rc = HB_OK;
  if (!rc) rc = dev_alloc(&q->master, (size_t)q->kinds * q->n_eff);
  if (!rc) rc = dev_alloc(&q->tmaster, (size_t)q->kinds * q->n_eff);
  if (!rc && noisy) rc = dev_alloc(&q->eps, (size_t)q->n_eff);
  if (!rc) rc = dev_alloc(&q->adam_m, (size_t)(q->kinds == 3 ? 2 : 1) * q->n_eff);
  if (!rc) rc = dev_alloc(&q->adam_v, (size_t)(q->kinds == 3 ? 2 : 1) * q->n_eff);
  if (!rc) rc = dev_alloc(&q->grad, (size_t)q->n_eff);
  if (!rc) rc = dev_alloc(&q->w16, (size_t)q->n_w);
  if (!rc) rc = dev_alloc(&q->w16t, (size_t)q->n_w);
  if (!rc) rc = dev_alloc(&q->t16, (size_t)q->n_w);
  if (!rc) rc = dev_alloc(&q->beff, (size_t)q->n_b);
  if (!rc) rc = dev_alloc(&q->tbeff, (size_t)q->n_b);
  if (!rc) rc = dev_alloc(&q->chain, 128);
  if (!rc) rc = dev_alloc(&q->chain2, 128);
  if (!rc) rc = dev_alloc(&q->scalars, 32);      // [0] mean loss [1] max weight [2] Adam step [3..7] its corrections
  if (rc) { hb_net_destroy(q); return rc; }
  *out = q;
  return HB_OK;
}

int hb_net_destroy(hb_qnet* q) {
  if (!q) return HB_OK;
  hb::DeviceGuard g(q->device);
  cudaDeviceSynchronize();
  dev_free(q->master); dev_free(q->tmaster); dev_free(q->eps); dev_free(q->adam_m); dev_free(q->adam_v); dev_free(q->grad);
  dev_free(q->w16); dev_free(q->w16t); dev_free(q->t16); dev_free(q->beff); dev_free(q->tbeff);
  dev_free(q->scalars);
  dev_free(q->chain);
  dev_free(q->chain2);
  if (q->side) {
    cudaStreamDestroy(q->side);
    if (q->side2) cudaStreamDestroy(q->side2);
    if (q->ev_join2) cudaEventDestroy(q->ev_join2);
    cudaEventDestroy(q->ev_fork);
    cudaEventDestroy(q->ev_join);
    if (q->ev_loss) cudaEventDestroy(q->ev_loss);
    if (q->ev_wt) cudaEventDestroy(q->ev_wt);
    if (q->ev_fwd) cudaEventDestroy(q->ev_fwd);
    for (int l = 0; l < kMaxLayers; ++l) cudaEventDestroy(q->ev_dz[l]);
    for (int l = 0; l < kMaxLayers; ++l) {
      if (q->lstream[l]) cudaStreamDestroy(q->lstream[l]);
      if (q->ev_bias[l]) cudaEventDestroy(q->ev_bias[l]);
      if (q->ev_layer[l]) cudaEventDestroy(q->ev_layer[l]);
    }
    if (q->ev_tick) cudaEventDestroy(q->ev_tick);
  }
  dev_free(q->act_x); dev_free(q->act_h[0]); dev_free(q->act_h[1]); dev_free(q->act_q);
  free_learner_ws(q);
  delete q;
  return HB_OK;
}

/* what: 0 n_layers, 1 n_eff (flat gradient length), 2 num_actions, 3 n_atoms, 4 noisy,
 *       5 adam step count, 10+l: in of layer l, 20+l: out of layer l, 30+l: K, 40+l: N */
int64_t hb_net_info(const hb_qnet* q, int what) {
  if (!q) return -1;
  if (what >= 10 && what < 50) {
    const int l = what % 10;
    if (l >= q->n_layers) return -1;
    const LayerGeo& L = q->tab.l[l];
    switch (what / 10) { case 1: return L.in; case 2: return L.out; case 3: return L.K; default: return L.N; }
  }
  switch (what) {
    case 0: return q->n_layers;
    case 1: return q->n_eff;
    case 2: return q->A;
    case 3: return q->n_atoms;
    case 4: return q->noisy;
    case 5: {   // the step counter lives on the device (graph replays advance it)
      int v = 0;
      hb::DeviceGuard g(q->device);
      if (cudaMemcpy(&v, q->scalars + 2, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
      return v;
    }
    default: return -1;
  }
}

float* hb_net_grad(hb_qnet* q) { return q ? q->grad : nullptr; }

/* kind: 0 w, 1 b, 2 w_mu, 3 b_mu, 4 w_sigma, 5 b_sigma, 6 eps_w, 7 eps_b, 8 dW (gradient of the
 * effective weight), 9 db.  HOST arrays in the reference's layout: weights [in, out] row-major
 * (noisy_mlp.py:61,73,78), biases [out]. */
static int net_param_io(hb_qnet* q, int layer, int kind, float* h_data, int64_t n, bool set,
                        bool target = false) {
  HB_REQUIRE(q && h_data && layer >= 0 && layer < q->n_layers && kind >= 0 && kind <= 9,
             "hb_net_param: bad arguments");
  HB_REQUIRE(set ? kind <= 7 : true, "hb_net_set_param: gradients are read-only");
  HB_REQUIRE(q->noisy || kind <= 1 || kind >= 8, "hb_net_param: plain MLP has only w and b");
  const LayerGeo& L = q->tab.l[layer];
  const bool is_w = (kind % 2) == 0;
  const int64_t want = is_w ? (int64_t)L.in * L.out : L.out;
  HB_REQUIRE(n == want, "hb_net_param: layer %d kind %d expects %lld values, got %lld", layer, kind,
             (long long)want, (long long)n);
  hb::DeviceGuard g(q->device);
  float* base;
  if (kind <= 5) base = (target ? q->tmaster : q->master) + (size_t)(kind / 2) * q->n_eff;
  else if (kind <= 7) base = q->eps;
  else base = q->grad;
  base += L.eoff + (is_w ? 0 : (long long)L.N * L.K);
  const size_t cnt = is_w ? (size_t)L.N * L.K : (size_t)L.N;
  std::vector<float> tmp(cnt, 0.f);
  HB_CUDA(cudaDeviceSynchronize());
  if (!set) HB_CUDA(cudaMemcpy(tmp.data(), base, cnt * 4, cudaMemcpyDeviceToHost));
  if (is_w) {
    for (int i = 0; i < L.in; ++i)
      for (int o = 0; o < L.out; ++o) {
        float& d = tmp[(size_t)out_to_n(q, layer, o) * L.K + i];
        float& h = h_data[(size_t)i * L.out + o];
        if (set) d = h; else h = d;
      }
  } else {
    for (int o = 0; o < L.out; ++o) {
      float& d = tmp[(size_t)out_to_n(q, layer, o)];
      if (set) d = h_data[o]; else h_data[o] = d;
    }
  }
  if (set) HB_CUDA(cudaMemcpy(base, tmp.data(), cnt * 4, cudaMemcpyHostToDevice));
  return HB_OK;
}

int hb_net_set_param(hb_qnet* q, int layer, int kind, const float* h_data, int64_t n) {
  return net_param_io(q, layer, kind, const_cast<float*>(h_data), n, true);
}
int hb_net_get_param(hb_qnet* q, int layer, int kind, float* h_out, int64_t n) {
  return net_param_io(q, layer, kind, h_out, n, false);
}
int hb_net_get_target_param(hb_qnet* q, int layer, int kind, float* h_out, int64_t n) {
  HB_REQUIRE(kind >= 0 && kind <= 5, "hb_net_get_target_param: kind 0..5");
  return net_param_io(q, layer, kind, h_out, n, false, true);
}

/* Recompute the fp16 effective weights from the masters (after hb_net_set_param). */
int hb_net_refresh(hb_qnet* q, void* stream) {
  HB_REQUIRE(q, "hb_net_refresh: NULL net");
  hb::DeviceGuard g(q->device);
  return run_adam(q, 0, 0.f, 0.f, 0.f, 0.f, 1.f, (cudaStream_t)stream);
}

/* trg_params = online_params  (rlax_rainbow.py:336-337) */
int hb_net_sync_target(hb_qnet* q, void* stream) {
  HB_REQUIRE(q, "hb_net_sync_target: NULL net");
  hb::DeviceGuard g(q->device);
  cudaStream_t s = (cudaStream_t)stream;
  HB_CUDA(cudaMemcpyAsync(q->t16, q->w16, (size_t)q->n_w * 2, cudaMemcpyDeviceToDevice, s));
  HB_CUDA(cudaMemcpyAsync(q->tbeff, q->beff, (size_t)q->n_b * 4, cudaMemcpyDeviceToDevice, s));
  HB_CUDA(cudaMemcpyAsync(q->tmaster, q->master, (size_t)q->kinds * q->n_eff * 4,
                          cudaMemcpyDeviceToDevice, s));
  return HB_OK;
}

/* The acting pass's persistent GEMMs leave `sms` SMs unoccupied (0: none), for work a caller overlaps with
 * them on another stream: their CTAs take a whole SM's registers and shared memory, so nothing else runs
 * beside them otherwise. */
int hb_net_set_act_reserve(hb_qnet* q, int sms) {
  HB_REQUIRE(q && sms >= 0 && sms < tc::sm_count(), "hb_net_set_act_reserve: bad arguments");
  q->act_reserve_sms = sms;
  return HB_OK;
}

/* DQNAgent.explore / exploit (rlax_rainbow.py:277-290) for n rows of bit-packed observations and
 * legal-move masks.  greedy != 0: eval_policy.  d_uniform: one U[0,1) (explore: U[0,0.999)) draw per
 * row, or NULL for the built-in counter-based stream keyed (seed, counter, row0 + row).
 * d_q (optional): masked q-values [n, num_actions]. */
int hb_net_act(hb_qnet* q, const uint32_t* d_obs, const uint32_t* d_legal, int legal_words, int64_t n,
               int greedy, float epsilon, const float* d_uniform, uint64_t seed, uint64_t counter,
               int64_t row0, int use_target, int32_t* d_action, float* d_q, void* stream) {
  HB_REQUIRE(q && d_obs && d_legal && d_action && n >= 0, "hb_net_act: bad arguments");
  HB_REQUIRE(legal_words == (q->A + 31) / 32, "hb_net_act: legal_words does not match %d actions", q->A);
  if (n == 0) return HB_OK;
  hb::DeviceGuard g(q->device);
  cudaStream_t s = (cudaStream_t)stream;
  const long long chunk = n < kActChunk ? n : kActChunk;
  int rc = ensure_act_ws(q, chunk);
  if (rc) return rc;
  const int ldq = q->n_atoms > 1 ? q->A : q->tab.l[q->n_layers - 1].N;
  for (long long r0 = 0; r0 < n; r0 += chunk) {
    const long long rows = n - r0 < chunk ? n - r0 : chunk;
    const bool fused = q->n_layers > 1;           // layer 0 reads the packed rows directly
    if (!fused) {
      rc = launch_unpack(d_obs + r0 * q->obs_stride, q->obs_stride, rows, q->act_x, q->tab.l[0].K,
                         nullptr, 0, 0, s);
      if (rc) return rc;
    }
    __half* hb[kMaxLayers];
    for (int l = 0; l < kMaxLayers; ++l) hb[l] = q->act_h[l & 1];
    rc = forward(q, use_target ? q->t16 : q->w16, use_target ? q->tbeff : q->beff, q->act_x, rows, hb,
                 nullptr, 0, nullptr, q->act_q, ldq, s, fused ? d_obs + r0 * q->obs_stride : nullptr, 0,
                 q->act_reserve_sms > 0 ? tc::sm_count() - q->act_reserve_sms : 0);
    if (rc) return rc;
    policy_kernel<<<(unsigned)((rows + 127) / 128), 128, 0, s>>>(
        q->act_q, ldq, d_legal + r0 * legal_words, legal_words, q->A, rows, greedy, epsilon,
        d_uniform ? d_uniform + r0 : nullptr, seed, counter, row0 + r0, d_action + r0,
        d_q ? d_q + r0 * q->A : nullptr);
    HB_CUDA(cudaGetLastError());
  }
  return HB_OK;
}

/* Forward + loss + backward of DQNLearning.update_q; leaves dL/dW_eff, dL/db_eff in the flat
 * gradient bucket (hb_net_grad) -- the buffer a data-parallel job all-reduces -- and
 *   d_loss_out[B]  per-sample loss (C51: cross-entropy; the reference's new_prios = |loss_b|,
 *                  rlax_rainbow.py:198) or TD error (double-Q, rlax_dqn.py:233)
 *   d_mean_loss[1] mean(loss_b * w_is)
 * d_prios: sampled probabilities (C51 variant: w = (1/prios)^beta / max, rlax_rainbow.py:188-189);
 * d_weights: importance weights given directly (rlax_dqn.py:144 signature); exactly one is non-NULL
 * (both NULL = all ones).  mask_terminal: gamma*(1-terminal) as rainbow_agent.py:191-197 (C51 only;
 * the rlax C51 learner ignores term_t).  Observations are bit-packed rows. */
static int net_grads_impl(hb_qnet* q, const uint32_t* d_obs_tm1, const int32_t* d_a_tm1, const float* d_r_t,
                          const uint32_t* d_obs_t, const uint8_t* d_term_t, const float* d_prios,
                          const float* d_weights, int B, float beta, float discount, int mask_terminal,
                          float* d_loss_out, float* d_mean_loss, void* stream, int tf_variant,
                          const uint32_t* d_legal_t, const hb_net_step_opts* opts = nullptr);

int hb_net_learner_buffers(hb_qnet* q, int B, uint32_t** obs_staging, float** loss, float** mean_loss) {
  HB_REQUIRE(q && B >= 1, "hb_net_learner_buffers: bad arguments");
  hb::DeviceGuard g(q->device);
  int rc = ensure_learner_ws(q, B);
  if (rc) return rc;
  if (obs_staging) *obs_staging = q->xb;
  if (loss) *loss = q->loss_b;
  if (mean_loss) *mean_loss = q->scalars;
  return HB_OK;
}

void* hb_net_loss_event(hb_qnet* q) { return q ? (void*)q->ev_loss : nullptr; }

// (internal, learner.cu) the host-side handles of the update's two PDL kernels: 0 = tc::mlp_chain_kernel,
// 1 = c51_loss_kernel<true>, 2 = c51_loss_kernel<false>
const void* hb_net_kernel_handle(int which) {
  switch (which) {
    case 0: return (const void*)tc::mlp_chain_kernel;
    case 1: return (const void*)c51_loss_kernel<true>;
    case 2: return (const void*)c51_loss_kernel<false>;
    default: return nullptr;
  }
}

int hb_net_grads(hb_qnet* q, const uint32_t* d_obs_tm1, const int32_t* d_a_tm1, const float* d_r_t,
                 const uint32_t* d_obs_t, const uint8_t* d_term_t, const float* d_prios,
                 const float* d_weights, int B, float beta, float discount, int mask_terminal,
                 float* d_loss_out, float* d_mean_loss, void* stream) {
  return net_grads_impl(q, d_obs_tm1, d_a_tm1, d_r_t, d_obs_t, d_term_t, d_prios, d_weights, B, beta, discount,
                        mask_terminal, d_loss_out, d_mean_loss, stream, 0, nullptr);
}

/* The learner's form of hb_net_grads (hb_net_step_opts in the header): importance weights already on the
 * device, the step's scalars read from device memory, kernels chained with programmatic dependent launch,
 * and optionally optix.adam applied layer by layer beside the backward pass. */
int hb_net_grads_ex(hb_qnet* q, const uint32_t* d_obs_tm1, const int32_t* d_a_tm1, const float* d_r_t,
                    const uint32_t* d_obs_t, const uint8_t* d_term_t, const float* d_prios,
                    const float* d_weights, int B, float beta, float discount, int mask_terminal,
                    float* d_loss_out, float* d_mean_loss, const hb_net_step_opts* opts, void* stream) {
  return net_grads_impl(q, d_obs_tm1, d_a_tm1, d_r_t, d_obs_t, d_term_t, d_prios, d_weights, B, beta, discount,
                        mask_terminal, d_loss_out, d_mean_loss, stream, 0, nullptr, opts);
}

float* hb_net_is_weights(hb_qnet* q, int B) {
  if (!q || B < 1) return nullptr;
  hb::DeviceGuard g(q->device);
  if (ensure_learner_ws(q, B)) return nullptr;
  return q->w_is;
}

/* TF-Rainbow target (rainbow/rainbow_agent.py:183-236): the greedy next action comes from the TARGET
 * network's q = sum(p z) over the LEGAL next moves, gamma is masked by the terminal flag, importance
 * weights are given (1/sqrt(priority), normalised by the caller's max).  Same kernels otherwise. */
int hb_net_grads_rainbow(hb_qnet* q, const uint32_t* d_obs_tm1, const int32_t* d_a_tm1, const float* d_r_t,
                         const uint32_t* d_obs_t, const uint32_t* d_legal_t, const uint8_t* d_term_t,
                         const float* d_weights, int B, float cumulative_gamma, float* d_loss_out,
                         float* d_mean_loss, void* stream) {
  HB_REQUIRE(q && q->n_atoms > 1 && d_legal_t && d_term_t, "hb_net_grads_rainbow: categorical net, legal moves "
             "and terminals are required");
  return net_grads_impl(q, d_obs_tm1, d_a_tm1, d_r_t, d_obs_t, d_term_t, nullptr, d_weights, B, 0.f,
                        cumulative_gamma, 1, d_loss_out, d_mean_loss, stream, 1, d_legal_t);
}

static int net_grads_impl(hb_qnet* q, const uint32_t* d_obs_tm1, const int32_t* d_a_tm1, const float* d_r_t,
                          const uint32_t* d_obs_t, const uint8_t* d_term_t, const float* d_prios,
                          const float* d_weights, int B, float beta, float discount, int mask_terminal,
                          float* d_loss_out, float* d_mean_loss, void* stream, int tf_variant,
                          const uint32_t* d_legal_t, const hb_net_step_opts* opts) {
  HB_REQUIRE(q && d_obs_tm1 && d_a_tm1 && d_r_t && d_obs_t && B >= 1, "hb_net_grads: bad arguments");
  HB_REQUIRE(!(d_prios && d_weights), "hb_net_grads: give prios or weights, not both");
  hb::DeviceGuard g(q->device);
  cudaStream_t s = (cudaStream_t)stream;
  int rc = ensure_learner_ws(q, B);
  if (rc) return rc;
  const int nl = q->n_layers, K0 = q->tab.l[0].K, Bp = q->lbp;
  const LayerGeo& LL = q->tab.l[nl - 1];
  const int ldq = q->n_atoms > 1 ? q->A : LL.N;
  const int pdl = opts ? opts->pdl : 0;
  const float* dyn = opts ? opts->d_dyn : nullptr;
  const bool per_layer = opts && opts->adam_per_layer;
  // C51: the head's backward product is part of the loss kernel (HB_DENSE_HEAD_BW=1: the dense GEMM)
  static const bool l_main = !(getenv("HB_DW0_OWN_STREAM") && atoi(getenv("HB_DW0_OWN_STREAM")) != 0);
  static const int adam_ctas = getenv("HB_ADAM_CTAS_PER_SM") ? atoi(getenv("HB_ADAM_CTAS_PER_SM")) : 0;
  static const bool bias0_main = getenv("HB_BIAS0_MAIN") && atoi(getenv("HB_BIAS0_MAIN")) != 0;
  static const bool dz0_plain = getenv("HB_DZ0_PLAIN") && atoi(getenv("HB_DZ0_PLAIN")) != 0;
  static const bool act_t_side = !(getenv("HB_ACT_T_SIDE") && atoi(getenv("HB_ACT_T_SIDE")) == 0);
  static const bool dense_head_bw = getenv("HB_DENSE_HEAD_BW") && atoi(getenv("HB_DENSE_HEAD_BW")) != 0;
  const bool head_bw = q->n_atoms > 1 && q->n_atoms <= 64 && nl >= 2 && LL.K % 8 == 0 &&
                       (size_t)q->n_atoms * LL.K * 2 <= (size_t)kHeadBwSmemMax && !dense_head_bw;
  // 1. packed rows side by side: [0, B) = obs_tm1, [B, 2B) = obs_t
  const size_t row_words = (size_t)q->obs_stride;
  // (a caller that gathered its rows straight into hb_net_obs_staging saves the two copies: each is a
  // node of a few microseconds on the critical path of the captured update)
  if (d_obs_tm1 != q->xb)
    HB_CUDA(cudaMemcpyAsync(q->xb, d_obs_tm1, (size_t)B * row_words * 4, cudaMemcpyDeviceToDevice, s));
  if (d_obs_t != q->xb + (size_t)B * row_words)
    HB_CUDA(cudaMemcpyAsync(q->xb + (size_t)B * row_words, d_obs_t, (size_t)B * row_words * 4,
                            cudaMemcpyDeviceToDevice, s));
  const bool fused = nl > 1;                 // layer 0 expands the bits inside the GEMM
  if (!fused && (rc = launch_unpack(q->xb, q->obs_stride, 2 * B, q->x, K0, nullptr, 0, 0, s))) return rc;
  // fork: the side stream runs the target network, the transposed observations and the importance
  // weights while the main stream runs the online network (both join before the loss)
  HB_CUDA(cudaEventRecord(q->ev_fork, s));
  // both networks, every layer, one launch (tc_chain.cuh) when every tile fits on the device at once ...
  int chained = 0;
  if (fused && (rc = forward_chain(q, B, d_a_tm1, tf_variant, ldq, s, pdl, &chained))) return rc;
  if (chained) HB_CUDA(cudaEventRecord(q->ev_fwd, s));
  if (!chained) {   // ... else the target network layer by layer on the side stream
    HB_CUDA(cudaStreamWaitEvent(q->side, q->ev_fork, 0));
    if ((rc = forward(q, q->t16, q->tbeff, q->x + (size_t)B * K0, B, q->th, nullptr, 0, q->tlogits,
                      q->n_atoms > 1 ? (tf_variant ? q->q_tg : nullptr) : q->tlogits, ldq, q->side,
                      fused ? q->xb + (size_t)B * row_words : nullptr, pdl))) return rc;
    HB_CUDA(cudaEventRecord(q->ev_join, q->side));
  }
  // a second side stream for the small kernels, so that the target network's GEMMs are all `side` runs
  HB_CUDA(cudaStreamWaitEvent(q->side2, q->ev_fork, 0));
  // importance weights (a caller that computed them with the sample -- hb_rb_sample_gather into
  // hb_net_is_weights -- passes opts->d_w_is == that buffer: nothing to do)
  if (opts && opts->d_w_is) {
    if (opts->d_w_is != q->w_is)
      HB_CUDA(cudaMemcpyAsync(q->w_is, opts->d_w_is, (size_t)B * 4, cudaMemcpyDeviceToDevice, q->side2));
  } else if (d_prios) {
    is_weights_kernel<<<1, 256, 0, q->side2>>>(d_prios, B, beta, q->w_is);
    HB_CUDA(cudaGetLastError());
  } else if (d_weights) {
    HB_CUDA(cudaMemcpyAsync(q->w_is, d_weights, (size_t)B * 4, cudaMemcpyDeviceToDevice, q->side2));
  } else {
    fill_kernel<<<(B + 255) / 256, 256, 0, q->side2>>>(q->w_is, 1.f, B);
    HB_CUDA(cudaGetLastError());
  }
  // dL/dlogits is written for the action taken only: the rest of both layouts is cleared here, beside
  // the forward passes (the [B, N] one used to sit on the main stream between the head and the loss)
  HB_CUDA(cudaMemsetAsync(q->dzt[nl - 1], 0, (size_t)LL.N * Bp * 2, q->side2));
  HB_CUDA(cudaMemsetAsync(q->dz[nl - 1], 0, (size_t)B * LL.N * 2, q->side2));
  HB_CUDA(cudaEventRecord(q->ev_join2, q->side2));
  {
    const long long tot = (long long)K0 * ((B + 7) / 8);
    unpack_t_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, q->side2>>>(q->xb, q->obs_stride, B, q->xt, Bp);
    HB_CUDA(cudaGetLastError());
  }
  // transposed effective weights for the dZ chain of the backward pass: refreshed beside the forward
  // passes instead of behind Adam where they were 8 us of every update's critical path (ev_wt: waited for
  // before the first dZ GEMM)
  if ((rc = launch_transposes(q, q->side2))) return rc;
  if (per_layer) {   // this step's Adam bias corrections and hyper-parameters, once, before any layer's update
    if ((rc = adam_tick(q, opts->lr, opts->b1, opts->b2, opts->eps, opts->grad_scale, dyn, q->side2))) return rc;
    HB_CUDA(cudaEventRecord(q->ev_tick, q->side2));
  }
  if (chained) {
    // transposed hidden activations of the sampled rows (B operand of the weight-gradient GEMMs), beside
    // the loss, on the high-priority side stream (idle when the forward passes are chained): behind the
    // weight transposes on side2 they started 5 us after the forward pass had ended and the backward tail
    // waited for them.  ev_join orders them before the backward pass.
    cudaStream_t ts = act_t_side ? q->side : q->side2;
    HB_CUDA(cudaStreamWaitEvent(ts, q->ev_fwd, 0));
    TransposeActs ta{};
    int widest = 0;
    for (int l = 0; l + 1 < nl; ++l) {
      ta.in[l] = q->h[l];
      ta.out[l] = q->ht[l];
      ta.cols[l] = q->tab.l[l].N;
      widest = ta.cols[l] > widest ? ta.cols[l] : widest;
    }
    static const bool per_layer_t = getenv("HB_MERGED_TRANSPOSE") && atoi(getenv("HB_MERGED_TRANSPOSE")) == 0;
    if (!per_layer_t) {
      transpose_act_kernel<<<dim3((widest + 31) / 32, (B + 31) / 32, nl - 1), 256, 0, ts>>>(ta, B, 2 * Bp);
      HB_CUDA(cudaGetLastError());
    } else {
      for (int l = 0; l + 1 < nl; ++l) {
        TransposeActs one{};
        one.in[0] = ta.in[l]; one.out[0] = ta.out[l]; one.cols[0] = ta.cols[l];
        transpose_act_kernel<<<dim3((ta.cols[l] + 31) / 32, (B + 31) / 32, 1), 256, 0, ts>>>(one, B, 2 * Bp);
        HB_CUDA(cudaGetLastError());
      }
    }
  }
  if (chained && act_t_side) HB_CUDA(cudaEventRecord(q->ev_join, q->side));
  HB_CUDA(cudaEventRecord(q->ev_wt, q->side2));
  // 2. online network on both observation sets (main stream)
  if (!chained) {
    if ((rc = forward(q, q->w16, q->beff, q->x, 2 * B, q->h, q->ht, 2 * Bp, q->logits, q->q_on, ldq, s,
                      fused ? q->xb : nullptr, pdl))) return rc;
    HB_CUDA(cudaStreamWaitEvent(s, q->ev_join, 0));
  }
  HB_CUDA(cudaStreamWaitEvent(s, q->ev_join2, 0));
  // 4. loss and dL/dlogits (scaled by B; 1/B is folded into the dW / db epilogues)
  {
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl ? 1 : 0;
    cudaLaunchConfig_t cfg = {};
    cfg.stream = s;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (q->n_atoms > 1) {
      cfg.gridDim = dim3((B + 3) / 4);
      cfg.blockDim = dim3(128);
      const float* q_sel = tf_variant ? q->q_tg : q->q_on + (size_t)B * ldq;
      const uint32_t* legal_t = tf_variant ? d_legal_t : nullptr;
      const float* logits = q->logits;
      const float* tlogits = q->tlogits;
      const float* w_is = q->w_is;
      HeadBackward hbw{};
      const float* qs = q_sel;
      if (head_bw) {
        hbw.w = q->w16 + LL.woff;
        hbw.K = LL.K;
        hbw.gate = q->h[nl - 2];
        hbw.dz_prev = q->dz[nl - 2];
        hbw.dzt_prev = q->dzt[nl - 2];
        cfg.gridDim = dim3(B);                   // one CTA per sample
        cfg.dynamicSmemBytes = (((size_t)q->n_atoms * LL.K * 2 + 127) & ~(size_t)127) + 64 * 4 + 16;
        static bool attr_set[64] = {};
        if (!attr_set[q->device & 63]) {
          HB_CUDA(cudaFuncSetAttribute(c51_loss_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       kHeadBwSmemMax + 1024));
          attr_set[q->device & 63] = true;
        }
        static const bool ltrace = getenv("HB_LOSS_TRACE") && atoi(getenv("HB_LOSS_TRACE")) != 0;
        static unsigned long long* d_ltrace = nullptr;
        cudaStreamCaptureStatus lcap = cudaStreamCaptureStatusNone;
        if (ltrace) cudaStreamIsCapturing(s, &lcap);
        if (ltrace && lcap == cudaStreamCaptureStatusNone) {
          if (d_ltrace == nullptr) HB_CUDA(cudaMalloc(&d_ltrace, 4096 * 8 * 8));
          HB_CUDA(cudaMemsetAsync(d_ltrace, 0, 4096 * 8 * 8, s));
          if (B <= 4096) hbw.trace = d_ltrace;
        }
        HB_CUDA(cudaLaunchKernelEx(&cfg, c51_loss_kernel<true>, logits, qs, ldq, legal_t, (q->A + 31) / 32,
                                   tlogits, LL.N, d_a_tm1, d_r_t, d_term_t, mask_terminal, w_is, B, q->A,
                                   q->n_atoms, q->vmax, discount, dyn, q->dz[nl - 1], q->dzt[nl - 1], Bp,
                                   q->loss_b, hbw, (chained && pdl) ? 1 : 0));
        if (hbw.trace != nullptr) {
          std::vector<unsigned long long> h((size_t)B * 8);
          HB_CUDA(cudaStreamSynchronize(s));
          HB_CUDA(cudaMemcpy(h.data(), d_ltrace, (size_t)B * 8 * 8, cudaMemcpyDeviceToHost));
          unsigned long long t0 = ~0ull;
          for (int c = 0; c < B; ++c) if (h[c * 8] && h[c * 8] < t0) t0 = h[c * 8];
          double mx[4] = {0, 0, 0, 0}, mn[4] = {1e30, 1e30, 1e30, 1e30};
          for (int c = 0; c < B; ++c)
            for (int k = 0; k < 4; ++k) {
              const double v = ((double)h[c * 8 + k] - (double)t0) * 1e-3;
              if (v > mx[k]) mx[k] = v;
              if (v < mn[k]) mn[k] = v;
            }
          fprintf(stderr, "c51 loss (%d samples): wait over %.1f-%.1f | row done %.1f-%.1f | weights in %.1f-%.1f | "
                  "dZ row stored %.1f-%.1f us\n", B, mn[0], mx[0], mn[1], mx[1], mn[2], mx[2], mn[3], mx[3]);
        }
      } else {
        HB_CUDA(cudaLaunchKernelEx(&cfg, c51_loss_kernel<false>, logits, qs, ldq, legal_t, (q->A + 31) / 32,
                                   tlogits, LL.N, d_a_tm1, d_r_t, d_term_t, mask_terminal, w_is, B, q->A,
                                   q->n_atoms, q->vmax, discount, dyn, q->dz[nl - 1], q->dzt[nl - 1], Bp,
                                   q->loss_b, hbw, 0));
      }
    } else {
      cfg.gridDim = dim3((B + 127) / 128);
      cfg.blockDim = dim3(128);
      const float* q_on = q->q_on;
      const float* q_t = q->tlogits;
      const float* w_is = q->w_is;
      HB_CUDA(cudaLaunchKernelEx(&cfg, dqn_loss_kernel, q_on, ldq, q_t, d_a_tm1, d_r_t, d_term_t, w_is, B, q->A,
                                 discount, dyn, q->dz[nl - 1], LL.N, q->dzt[nl - 1], Bp, q->loss_b));
    }
  }
  HB_CUDA(cudaEventRecord(q->ev_loss, s));   // the per-sample losses are complete (hb_net_loss_event)
  // mean(loss * w) is only reported: beside the backward pass, not in front of it
  HB_CUDA(cudaStreamWaitEvent(q->side2, q->ev_loss, 0));
  mean_loss_kernel<<<1, 256, 0, q->side2>>>(q->loss_b, q->w_is, B, q->n_atoms > 1 ? 0 : 1, q->scalars);
  HB_CUDA(cudaGetLastError());
  if (d_loss_out && d_loss_out != q->loss_b)
    HB_CUDA(cudaMemcpyAsync(d_loss_out, q->loss_b, (size_t)B * 4, cudaMemcpyDeviceToDevice, q->side2));
  if (d_mean_loss && d_mean_loss != q->scalars)
    HB_CUDA(cudaMemcpyAsync(d_mean_loss, q->scalars, 4, cudaMemcpyDeviceToDevice, q->side2));
  // 5. backward: the dZ chain (critical path) on the main stream with programmatic dependent launch;
  //    every layer's weight gradient on that layer's own stream as soon as its dZ exists, the bias
  //    gradients on side2, and (adam_per_layer) that layer's optix.adam right behind its gradient, so
  //    that only the first layer's update is left when the chain ends
  const float invB = 1.0f / (float)B;
  HB_CUDA(cudaStreamWaitEvent(s, q->ev_wt, 0));
  if (chained && act_t_side) HB_CUDA(cudaStreamWaitEvent(s, q->ev_join, 0));
  bool tail_chained = false;          // dZ_0^T, dW_0 and db_0 came out of one launch (backward_tail_chain)
  for (int l = nl - 1; l >= 0; --l) {
    const LayerGeo& L = q->tab.l[l];
    // the first layer's gradient and update are the tail of the critical path: they stay on the main
    // stream (nothing else is left for it), right behind the last dZ GEMM with a programmatic launch --
    // handed to another stream the gradient GEMM started 4 to 10 us after its input was complete
    const bool on_main = l == 0 && per_layer && l_main;
    cudaStream_t ls = on_main ? s : q->lstream[l];
    HB_CUDA(cudaEventRecord(q->ev_dz[l], s));              // dz[l] / dzt[l] are complete here
    if (!on_main) HB_CUDA(cudaStreamWaitEvent(ls, q->ev_dz[l], 0));
    // dW_l [N, K] = dZ_l^T . X_l  : A = dzt[l] [N, B], B = (layer input)^T [K, B]
    const __half* in_t = l == 0 ? q->xt : q->ht[l - 1];
    const int ld_in_t = l == 0 ? Bp : 2 * Bp;
    tc::EpiParams ep{};
    ep.alpha = invB;
    ep.out32 = q->grad + L.eoff;
    ep.ld_out = L.K;
    ep.pdl = on_main ? pdl : 0;
    const bool have_grads = l == 0 && tail_chained;
    if (!have_grads && (rc = tc::gemm<tc::EPI_F32>(q->dzt[l], Bp, in_t, ld_in_t, L.N, L.K, B, ep, ls))) return rc;
    // (the first layer's bias gradient on the high-priority side stream, idle since the target network's
    // forward pass: on side2 it queued behind the other layers' Adam CTAs and held the last update up)
    cudaStream_t bs = on_main ? (bias0_main ? s : q->side) : q->side2;
    if (bs != s && !have_grads) HB_CUDA(cudaStreamWaitEvent(bs, q->ev_dz[l], 0));
    if (!have_grads) {
      // programmatic launch for the first layer: its CTAs are placed while the SMs still have room and
      // wait inside the kernel, instead of queueing behind the other layers' Adam CTAs when dZ is ready
      cudaLaunchAttribute battr[1];
      battr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      battr[0].val.programmaticStreamSerializationAllowed = (on_main && pdl) ? 1 : 0;
      cudaLaunchConfig_t bcfg = {};
      bcfg.gridDim = dim3((L.N + 7) / 8);
      bcfg.blockDim = dim3(256);
      bcfg.stream = bs;
      bcfg.attrs = battr;
      bcfg.numAttrs = 1;
      const __half* zt = q->dzt[l];
      HB_CUDA(cudaLaunchKernelEx(&bcfg, bias_grad_kernel, zt, Bp, L.N, B, invB,
                                 q->grad + L.eoff + (long long)L.N * L.K));
    }
    if (per_layer) {
      if (!have_grads) {
        HB_CUDA(cudaEventRecord(q->ev_bias[l], bs));
        HB_CUDA(cudaStreamWaitEvent(ls, q->ev_bias[l], 0));
      }
      HB_CUDA(cudaStreamWaitEvent(ls, q->ev_tick, 0));
      const long long e0 = L.eoff, e1 = l + 1 < nl ? q->tab.l[l + 1].eoff : q->n_eff;
      // data parallel: this layer's slice of the bucket is summed over the ranks on the layer's own
      // channel while the dZ chain and the other layers' gradients are still being computed
      if (opts->comm != nullptr && (rc = hb_comm_allreduce_range(opts->comm, l, e0, e1 - e0, ls))) return rc;
      if ((rc = adam_range(q, 1, e0 / 4, e1 / 4, ls, 0, on_main ? 0 : adam_ctas))) return rc;
    }
    HB_CUDA(cudaEventRecord(q->ev_layer[l], ls));
    if (l == 1 && !(l == nl - 1 && head_bw) && per_layer && l_main && !dz0_plain) {
      // The last product of the chain, dZ_0, is only ever read transposed (first layer's weight and bias
      // gradients): computed as dZ_0^T [K, B] = W_1^T [K, N] . dZ_1^T, operands swapped, it leaves through
      // the tile's bulk tensor store instead of 2-byte scattered stores (2 us of the critical path).
      // The ReLU gate is read transposed too: ht[0] [K, 2Bp] (complete before ev_wt).
      // ... and, when every tile of it and of the first layer's weight gradient fits on the device at once,
      // the two are one launch that also finishes the bias gradient (backward_tail_chain)
      int tail = 0;
      if ((rc = backward_tail_chain(q, B, Bp, s, pdl, &tail))) return rc;
      tail_chained = tail != 0;
      if (!tail_chained) {
        tc::EpiParams eb{};
        eb.alpha = 1.f;
        eb.gate = q->ht[0];
        eb.ld_gate = 2 * Bp;
        eb.out16 = q->dzt[0];
        eb.ld_out = Bp;
        eb.pdl = pdl;
        if ((rc = tc::gemm<tc::EPI_F16>(q->w16t + L.woff, L.N, q->dz[1], L.N, L.K, B, L.N, eb, s))) return rc;
      }
    } else if (l > 0 && !(l == nl - 1 && head_bw)) {
      // dZ_{l-1} [B, K] = (dZ_l . W_l) * relu'(H_{l-1}) : A = dz[l] [B, N], B = W^T [K, N]
      tc::EpiParams eb{};
      eb.alpha = 1.f;
      eb.gate = q->h[l - 1];
      eb.ld_gate = L.K;
      eb.out16 = q->dz[l - 1];
      eb.ld_out = L.K;
      eb.out16_t = q->dzt[l - 1];
      eb.ld_out_t = Bp;
      eb.pdl = pdl;
      if ((rc = tc::gemm<tc::EPI_F16>(q->dz[l], L.N, q->w16t + L.woff, L.N, B, L.K, L.N, eb, s))) return rc;
    }
  }
  for (int l = 0; l < nl; ++l) HB_CUDA(cudaStreamWaitEvent(s, q->ev_layer[l], 0));
  HB_CUDA(cudaEventRecord(q->ev_join2, q->side2));
  HB_CUDA(cudaStreamWaitEvent(s, q->ev_join2, 0));
  return HB_OK;
}

/* optix.adam update of every master tensor from the flat gradient bucket (after the optional
 * all-reduce; grad_scale = 1/world_size) and refresh of the fp16 effective weights. */
int hb_net_adam_tf(hb_qnet* q, float lr, float b1, float b2, float eps, float grad_scale, void* stream) {
  HB_REQUIRE(q, "hb_net_adam_tf: NULL net");
  hb::DeviceGuard g(q->device);
  return run_adam(q, 1, lr, b1, b2, eps, grad_scale, (cudaStream_t)stream, 1);
}

int hb_net_adam(hb_qnet* q, float lr, float b1, float b2, float eps, float grad_scale, void* stream) {
  HB_REQUIRE(q, "hb_net_adam: NULL net");
  hb::DeviceGuard g(q->device);
  return run_adam(q, 1, lr, b1, b2, eps, grad_scale, (cudaStream_t)stream);
}

int hb_net_adam_dyn(hb_qnet* q, const float* d_dyn, void* stream) {
  HB_REQUIRE(q && d_dyn, "hb_net_adam_dyn: bad arguments");
  hb::DeviceGuard g(q->device);
  return run_adam(q, 1, 0.f, 0.f, 0.f, 0.f, 1.f, (cudaStream_t)stream, 0, d_dyn);
}

/* Q-values / logits of n packed observations (tests, evaluation): d_q [n, num_actions] unmasked;
 * d_logits [n, N_last] (categorical: column a*64 + j) or NULL. */
int hb_net_forward(hb_qnet* q, const uint32_t* d_obs, int64_t n, int use_target, float* d_q,
                   float* d_logits, void* stream) {
  HB_REQUIRE(q && d_obs && d_q && n >= 1 && n <= kActChunk, "hb_net_forward: bad arguments (n <= %d)", kActChunk);
  hb::DeviceGuard g(q->device);
  cudaStream_t s = (cudaStream_t)stream;
  int rc = ensure_act_ws(q, n);
  if (rc) return rc;
  const bool fused = q->n_layers > 1;
  if (!fused && (rc = launch_unpack(d_obs, q->obs_stride, n, q->act_x, q->tab.l[0].K, nullptr, 0, 0, s))) return rc;
  __half* hb[kMaxLayers];
  for (int l = 0; l < kMaxLayers; ++l) hb[l] = q->act_h[l & 1];
  const int ldq = q->n_atoms > 1 ? q->A : q->tab.l[q->n_layers - 1].N;
  rc = forward(q, use_target ? q->t16 : q->w16, use_target ? q->tbeff : q->beff, q->act_x, n, hb,
               nullptr, 0, d_logits, q->act_q, ldq, s, fused ? d_obs : nullptr);
  if (rc) return rc;
  HB_CUDA(cudaMemcpy2DAsync(d_q, (size_t)q->A * 4, q->act_q, (size_t)ldq * 4, (size_t)q->A * 4, (size_t)n,
                            cudaMemcpyDeviceToDevice, s));
  return HB_OK;
}

}  // extern "C"
