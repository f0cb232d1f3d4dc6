// Self-play session bookkeeping on the device: who gets which step type, the action each seat
// last took and the reward it has collected since (the per-player accounting of
// hanabi_agents/rainbow/run_experiment.py:280-336, batched over N lockstep games).
//
// Seats, not players: all games advance one move per tick and seat k = tick mod P moves in every
// game, so a seat is the agent object that acts at that tick.  A finished game spends one tick on
// its reset (the env's auto-reset step), which keeps every game on the same rotation.
#include "common.cuh"
#include <new>

struct hb_session {
  int device = 0;
  int P = 0;
  long long n = 0;
  float* acc = nullptr;          // [P, N] reward since that seat's last action   (:299,310)
  int32_t* pact = nullptr;       // [P, N] that seat's last action
  uint8_t* pending = nullptr;    // [P, N] seat has acted in this episode         (has_played :296)
  uint8_t* agent_type = nullptr; // [P, N] step type each seat is shown this tick
  uint8_t* prev_type = nullptr;  // [N] game-level step type before the env step of this tick
  float* ep_return = nullptr;    // [N] undiscounted return of the running episode (:307)
  int32_t* ep_len = nullptr;     // [N] moves of the running episode               (:312)
  int32_t* ep_count = nullptr;   // [N] finished episodes
  float* first_return = nullptr; // [N] return / length of the first finished episode (evaluation)
  int32_t* first_len = nullptr;
  double* stats = nullptr;       // [4] sum of returns, episodes, sum of lengths, sum of squared returns
  int turn_agent = -1;
};

namespace {

constexpr int kThreads = 256;
enum { FIRST = 0, MID = 1, LAST = 2 };

__global__ void begin_turn_kernel(hb_session s, int agent, const uint8_t* __restrict__ game_type) {
  const long long g = (long long)blockIdx.x * kThreads + threadIdx.x;
  if (g >= s.n) return;
  const int t = game_type[g];
  s.prev_type[g] = (uint8_t)t;
  for (int j = 0; j < s.P; ++j) {
    const bool pend = s.pending[(long long)j * s.n + g] != 0;
    int a;
    if (t == LAST) a = pend ? LAST : FIRST;            // end_episode: every seat that has played (:335)
    else if (j == agent) a = pend ? MID : FIRST;       // step / begin_episode                    (:319-327)
    else a = FIRST;                                    // not this seat's turn: nothing to record
    s.agent_type[(long long)j * s.n + g] = (uint8_t)a;
  }
}

__global__ void end_turn_kernel(hb_session s, int agent, const int32_t* __restrict__ action,
                                const float* __restrict__ reward, int lenient) {
  const long long g = (long long)blockIdx.x * kThreads + threadIdx.x;
  double ret = 0.0, len = 0.0, cnt = 0.0, sq = 0.0;
  if (g < s.n) {
    if (s.prev_type[g] == LAST) {
      // the env spent this tick on the reset: close the episode's books
      const float r = s.ep_return[g];
      const int l = s.ep_len[g];
      ret = r; len = l; cnt = 1.0; sq = (double)r * r;
      if (s.ep_count[g] == 0) { s.first_return[g] = r; s.first_len[g] = l; }
      s.ep_count[g] += 1;
      s.ep_return[g] = 0.0f;
      s.ep_len[g] = 0;
      for (int j = 0; j < s.P; ++j) {
        s.pending[(long long)j * s.n + g] = 0;
        s.acc[(long long)j * s.n + g] = 0.0f;
      }
    } else {
      float r = reward[g];
      if (lenient && r < 0.0f) r = 0.0f;                 // LENIENT_SCORE  (:306)
      s.pact[(long long)agent * s.n + g] = action[g];
      s.pending[(long long)agent * s.n + g] = 1;
      for (int j = 0; j < s.P; ++j) {
        const long long o = (long long)j * s.n + g;
        s.acc[o] = (j == agent ? 0.0f : s.acc[o]) + r;   // accumulator cleared when the seat acts (:331)
      }
      s.ep_return[g] += r;
      s.ep_len[g] += 1;
    }
  }
  // episode statistics: warp sums -> block sums -> four atomics per block that closed an episode
  // (one set per warp serialised 8,000 double atomics on one address: 16 us of a 65,536-game tick)
  __shared__ double red[4][kThreads / 32];
  for (int o = 16; o > 0; o >>= 1) {
    ret += __shfl_xor_sync(0xffffffffu, ret, o);
    len += __shfl_xor_sync(0xffffffffu, len, o);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    sq += __shfl_xor_sync(0xffffffffu, sq, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { red[0][warp] = ret; red[1][warp] = cnt; red[2][warp] = len; red[3][warp] = sq; }
  __syncthreads();
  if (warp == 0) {
    double v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      v[k] = lane < kThreads / 32 ? red[k][lane] : 0.0;
      for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    }
    if (lane == 0 && v[1] > 0.0) {
      atomicAdd(&s.stats[0], v[0]);
      atomicAdd(&s.stats[1], v[1]);
      atomicAdd(&s.stats[2], v[2]);
      atomicAdd(&s.stats[3], v[3]);
    }
  }
}

template <class T>
int dev_alloc(T** p, long long count) {
  HB_CUDA(cudaMalloc((void**)p, (size_t)count * sizeof(T)));
  return HB_OK;
}

int zero_all(hb_session* s, cudaStream_t st, bool with_stats) {
  const long long pn = (long long)s->P * s->n;
  HB_CUDA(cudaMemsetAsync(s->acc, 0, pn * sizeof(float), st));
  HB_CUDA(cudaMemsetAsync(s->pact, 0, pn * sizeof(int32_t), st));
  HB_CUDA(cudaMemsetAsync(s->pending, 0, pn, st));
  HB_CUDA(cudaMemsetAsync(s->agent_type, 0, pn, st));
  HB_CUDA(cudaMemsetAsync(s->prev_type, 0, s->n, st));
  HB_CUDA(cudaMemsetAsync(s->ep_return, 0, s->n * sizeof(float), st));
  HB_CUDA(cudaMemsetAsync(s->ep_len, 0, s->n * sizeof(int32_t), st));
  if (with_stats) {
    HB_CUDA(cudaMemsetAsync(s->ep_count, 0, s->n * sizeof(int32_t), st));
    HB_CUDA(cudaMemsetAsync(s->first_return, 0, s->n * sizeof(float), st));
    HB_CUDA(cudaMemsetAsync(s->first_len, 0, s->n * sizeof(int32_t), st));
    HB_CUDA(cudaMemsetAsync(s->stats, 0, 4 * sizeof(double), st));
  }
  return HB_OK;
}

}  // namespace

extern "C" {

int hb_session_create(int players, int64_t n_games, int device, hb_session** out) {
  HB_REQUIRE(out != nullptr, "hb_session_create: out is NULL");
  HB_REQUIRE(players >= 2 && players <= 5 && n_games > 0, "hb_session_create: players %d, games %lld",
             players, (long long)n_games);
  hb::DeviceGuard guard(device);
  HB_REQUIRE(guard.ok, "hb_session_create: cannot select device %d", device);
  hb_session* s = new (std::nothrow) hb_session();
  HB_REQUIRE(s != nullptr, "hb_session_create: out of host memory");
  s->device = device;
  s->P = players;
  s->n = n_games;
  const long long pn = (long long)players * n_games;
  int rc = dev_alloc(&s->acc, pn);
  if (!rc) rc = dev_alloc(&s->pact, pn);
  if (!rc) rc = dev_alloc(&s->pending, pn);
  if (!rc) rc = dev_alloc(&s->agent_type, pn);
  if (!rc) rc = dev_alloc(&s->prev_type, n_games);
  if (!rc) rc = dev_alloc(&s->ep_return, n_games);
  if (!rc) rc = dev_alloc(&s->ep_len, n_games);
  if (!rc) rc = dev_alloc(&s->ep_count, n_games);
  if (!rc) rc = dev_alloc(&s->first_return, n_games);
  if (!rc) rc = dev_alloc(&s->first_len, n_games);
  if (!rc) rc = dev_alloc(&s->stats, 4);
  if (!rc) rc = zero_all(s, nullptr, true);
  if (!rc && cudaStreamSynchronize(nullptr) != cudaSuccess) rc = HB_ERR_CUDA;
  if (rc) { hb_session_destroy(s); return rc; }
  *out = s;
  return HB_OK;
}

int hb_session_destroy(hb_session* s) {
  if (s == nullptr) return HB_OK;
  hb::DeviceGuard guard(s->device);
  cudaFree(s->acc); cudaFree(s->pact); cudaFree(s->pending); cudaFree(s->agent_type);
  cudaFree(s->prev_type); cudaFree(s->ep_return); cudaFree(s->ep_len); cudaFree(s->ep_count);
  cudaFree(s->first_return); cudaFree(s->first_len); cudaFree(s->stats);
  delete s;
  return HB_OK;
}

int hb_session_reset(hb_session* s, int clear_stats, void* stream) {
  HB_REQUIRE(s != nullptr, "hb_session_reset: NULL session");
  hb::DeviceGuard guard(s->device);
  s->turn_agent = -1;
  return zero_all(s, (cudaStream_t)stream, clear_stats != 0);
}

int hb_session_begin_turn(hb_session* s, int agent, const uint8_t* d_game_step_type, void* stream) {
  HB_REQUIRE(s != nullptr && d_game_step_type != nullptr, "hb_session_begin_turn: NULL argument");
  HB_REQUIRE(agent >= 0 && agent < s->P, "hb_session_begin_turn: seat %d of %d", agent, s->P);
  hb::DeviceGuard guard(s->device);
  const unsigned grid = (unsigned)((s->n + kThreads - 1) / kThreads);
  begin_turn_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(*s, agent, d_game_step_type);
  HB_CUDA(cudaGetLastError());
  s->turn_agent = agent;
  return HB_OK;
}

int hb_session_end_turn(hb_session* s, int agent, const int32_t* d_action, const float* d_reward,
                        int lenient, void* stream) {
  HB_REQUIRE(s != nullptr && d_action != nullptr && d_reward != nullptr, "hb_session_end_turn: NULL argument");
  if (agent != s->turn_agent) {
    hb::set_error("hb_session_end_turn: seat %d ends a turn that seat %d began", agent, s->turn_agent);
    return HB_ERR_STATE;
  }
  hb::DeviceGuard guard(s->device);
  const unsigned grid = (unsigned)((s->n + kThreads - 1) / kThreads);
  end_turn_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(*s, agent, d_action, d_reward, lenient);
  HB_CUDA(cudaGetLastError());
  s->turn_agent = -1;
  return HB_OK;
}

void* hb_session_buffer(hb_session* s, int which, int agent) {
  if (s == nullptr) return nullptr;
  const long long o = (agent >= 0 && agent < s->P) ? (long long)agent * s->n : 0;
  switch (which) {
    case 0: return s->acc + o;
    case 1: return s->pact + o;
    case 2: return s->agent_type + o;
    case 3: return s->pending + o;
    case 4: return s->ep_return;
    case 5: return s->ep_len;
    case 6: return s->ep_count;
    case 7: return s->first_return;
    case 8: return s->first_len;
    case 9: return s->stats;
    case 10: return s->prev_type;
    default: return nullptr;
  }
}

int hb_session_stats(hb_session* s, double* h_out4, void* stream) {
  HB_REQUIRE(s != nullptr && h_out4 != nullptr, "hb_session_stats: NULL argument");
  hb::DeviceGuard guard(s->device);
  HB_CUDA(cudaMemcpyAsync(h_out4, s->stats, 4 * sizeof(double), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  HB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return HB_OK;
}

}  // extern "C"
