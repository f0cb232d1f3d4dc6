// Batched Hanabi environment + rule-based policy kernels and their C ABI.
// Device code: hanabi_game.cuh (state, legality, transition), hanabi_encode.cuh (canonical
// observation), hanabi_rules.cuh (rule-based agents).  One thread per game, 64 games per CTA.
#include "common.cuh"
#include "hanabi_encode.cuh"
#include "hanabi_rules.cuh"
#include <algorithm>
#include <cstring>
#include <new>
#include <vector>

using namespace hbg;

struct hb_env {
  int device = 0;
  int P = 0, H = 0, C = 0, R = 0, shape = -1;
  int deck_size = 0, num_moves = 0, legal_words = 0, state_words = 0, obs_len = 0, obs_stride = 0;
  Rt rt{};
  uint32_t* state = nullptr;
  uint8_t* decks = nullptr;
  unsigned long long* illegal = nullptr;  // device counter
  // work queue of the fused self-play kernel (SpQueue below): [head, tail] + group[G] + state[G]
  unsigned long long* spq = nullptr;
  int spq_groups = 0;
  unsigned long long spq_seq = 0;   // self-play launches so far
  // A self-play launch may overlap the previous kernel of its stream only when that kernel is this env's
  // previous self-play launch (it hands the games over through the ring; anything else would have to be
  // waited for with griddepcontrol.wait to see its memory).  Every other entry point clears the flag.
  bool spq_chain = false;
  void* spq_chain_stream = nullptr;
  const void* spq_chain_hist = nullptr;
  bool is_reset = false;
  // staging for *_host entry points
  void* stage = nullptr;
  size_t stage_bytes = 0;
  // chunked host pipeline (hb_env_step_host_async / hb_rules_act_host_async / hb_env_host_wait)
  static constexpr int kMaxChunks = 16;
  cudaStream_t hs_step = nullptr, hs_out = nullptr, hs_act = nullptr;
  cudaEvent_t hev_step[kMaxChunks] = {};
  cudaEvent_t hev_done = nullptr;
  int h_chunks = 0;
  // double-buffered variant (hb_env_step_host_begin / _outputs / hb_env_host_wait_actions): staging slot of
  // the step in flight, "outputs of slot s are on the host" and "the agent has read the state" events
  int h_slot = 0;
  cudaEvent_t hev_out[2] = {nullptr, nullptr};
  cudaEvent_t hev_rules = nullptr, hev_acts = nullptr;   // hev_acts: the next actions have left for the host
  // Rule lists seen so far, with their comparison tables already built: an agent passes the same
  // (ids, fargs, iargs, seed) on every call, and the tables (52 x 64 float64 divisions per threshold
  // rule) cost more host time than a 20-step launch takes on the device.
  struct RuleKey {
    int n = -1;
    unsigned long long seed = 0;
    int32_t ids[hbg::kMaxRules] = {};
    int32_t iarg[hbg::kMaxRules] = {};
    double farg[hbg::kMaxRules] = {};
  };
  static constexpr int kRuleCache = 8;
  RuleKey rule_key[kRuleCache];
  hbg::RuleSet rule_set[kRuleCache];
  bool rule_ext[kRuleCache] = {};
  int rule_cached = 0, rule_next = 0;
  unsigned long long rule_fills = 0;    // hb_env_info(e, 9): tables built so far
};

namespace {

static_assert(hbg::kRuleBlock == 64, "rule tables are sized for 64-thread CTAs");
constexpr int kThreads = 64;   // games per CTA: 65,536 games -> 1024 CTAs = 6.9 per SM (148 SMs)

__device__ __forceinline__ long long game_index() {
  return (long long)blockIdx.x * kThreads + threadIdx.x;
}

// Writes the step outputs of one game; returns the legal-move mask of the player to move.
// `obs_stride` (words) is kObsStride (16-byte row stores) or kObsWords: rows without the alignment
// padding, for consumers on the far side of PCIe (word stores).
template <class SH>
__device__ __forceinline__ unsigned long long emit(const Game<SH>& g, long long row,
                                                   uint32_t* __restrict__ obs,
                                                   uint32_t* __restrict__ legal,
                                                   uint8_t* __restrict__ step_type,
                                                   int obs_stride = SH::kObsStride) {
  if (obs != nullptr) {
    if (obs_stride == SH::kObsStride) {
      encode_obs(g, SH::P - 1, obs + row * SH::kObsStride);
    } else {
      uint32_t w[SH::kObsStride];
      encode_row(g, SH::P - 1, w);
#pragma unroll
      for (int i = 0; i < SH::kObsWords; ++i) obs[row * obs_stride + i] = w[i];
    }
  }
  const unsigned long long m = legal_mask(g);
  if (legal != nullptr) {
    legal[row * SH::kLegalWords] = (uint32_t)m;
    if (SH::kLegalWords > 1) legal[row * SH::kLegalWords + 1] = (uint32_t)(m >> 32);
  }
  if (step_type != nullptr) step_type[row] = (uint8_t)g.step_type;
  return m;
}

template <class SH>
__global__ void __launch_bounds__(kThreads)
env_reset_kernel(Rt rt, uint32_t* __restrict__ state, const uint8_t* __restrict__ decks,
                 uint32_t* __restrict__ obs, uint32_t* __restrict__ legal,
                 uint8_t* __restrict__ step_type) {
  const long long game = game_index();
  if (game >= rt.n_games) return;
  Game<SH> g;
  g.episode = 0;
  reset_game(g, deck_of<SH>(rt, decks, game, 0));
  emit(g, game, obs, legal, step_type);
  store_game(g, state, game);
}

// One env step for one game (shared by the step kernel and the fused self-play kernel).
// A game whose previous step was LAST is reset onto its next deck (action ignored, reward 0,
// FIRST) -- otherwise rl_env.step: apply, deal, reward = score delta.
// `row` is the deck-pool row of the current episode, (game_id + episode * n_total) mod n_decks; the
// next episode's row is row + (n_total mod n_decks) with one conditional subtraction (no 64-bit
// modulo in the step loop).  `next_card` = deck[g.deck_pos], loaded by the caller ahead of time.
template <class SH>
__device__ __forceinline__ float step_game(Game<SH>& g, const Rt& rt, const uint8_t* decks,
                                           const uint8_t*& deck, unsigned long long& row,
                                           unsigned long long row_step, int action, uint32_t next_card,
                                           unsigned long long lm, unsigned long long* illegal,
                                           const MoveTable* mt = nullptr) {
  if (g.step_type == LAST) {
    ++g.episode;
    row += row_step;
    if (row >= (unsigned long long)rt.n_decks) row -= (unsigned long long)rt.n_decks;
    deck = decks + row * kDeckPitch;
    reset_game(g, deck);
    return 0.0f;
  }
  bool legal_action;                                               // lm: legal moves of g
  if constexpr (SH::A <= 32) legal_action = (unsigned)action < (unsigned)SH::A && (((uint32_t)lm >> action) & 1u);
  else legal_action = (unsigned)action < (unsigned)SH::A && ((lm >> action) & 1ull);
  if (!legal_action) {
    atomicAdd(illegal, 1ull);   // HLE REQUIREs legality; we count and leave the game as it is
    return 0.0f;
  }
  const int r = apply_move(g, action, next_card, mt);
  g.step_type = g.terminal() ? LAST : MID;
  if (g.step_type == LAST) {
    // the next step deals from the next episode's deck row: ask for it now (in the fused self-play kernel
    // the deal's 32 bytes were a cold miss that a whole warp waited for: 6 % of its stall samples)
    unsigned long long nrow = row + row_step;
    if (nrow >= (unsigned long long)rt.n_decks) nrow -= (unsigned long long)rt.n_decks;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(decks + nrow * kDeckPitch));
  }
  return (float)r;
}

// deck[g.deck_pos] (the pitch leaves room to read one past a 50-card deck)
template <class SH>
__device__ __forceinline__ uint32_t peek_card(const Game<SH>& g, const uint8_t* deck) {
  return __ldg(deck + g.deck_pos);
}

template <class SH>
__global__ void __launch_bounds__(kThreads)
env_step_kernel(Rt rt, uint32_t* __restrict__ state, const uint8_t* __restrict__ decks,
                const int32_t* __restrict__ action, uint32_t* __restrict__ obs,
                uint32_t* __restrict__ legal, float* __restrict__ reward,
                uint8_t* __restrict__ step_type, unsigned long long* illegal,
                long long first, long long count, int obs_stride) {
  const long long game = first + game_index();          // games [first, first + count) of this env
  if (game >= first + count) return;
  Game<SH> g;
  load_game(g, state, game);
  unsigned long long row = deck_row<SH>(rt, game, g.episode);
  const uint8_t* deck = decks + row * kDeckPitch;
  const float r = step_game(g, rt, decks, deck, row, rt.row_step, action[game], peek_card(g, deck),
                            legal_mask(g), illegal);
  if (reward != nullptr) reward[game] = r;
  emit(g, game, obs, legal, step_type, obs_stride);
  store_game(g, state, game);
}

// HanabiObservation of the player `offset` seats after the player to move.
template <class SH>
__global__ void __launch_bounds__(kThreads)
env_observe_kernel(Rt rt, const uint32_t* __restrict__ state, int offset,
                   uint32_t* __restrict__ obs, uint32_t* __restrict__ legal) {
  const long long game = game_index();
  if (game >= rt.n_games) return;
  Game<SH> g;
  load_game(g, state, game);
  if (legal != nullptr) {   // HanabiState::LegalMoves: empty unless it is that player's turn
    const unsigned long long m = offset == 0 ? legal_mask(g) : 0ull;
    legal[game * SH::kLegalWords] = (uint32_t)m;
    if (SH::kLegalWords > 1) legal[game * SH::kLegalWords + 1] = (uint32_t)(m >> 32);
  }
  if (obs != nullptr) {
    rotate_seats(g, offset);
    encode_obs(g, SH::P - 1 - offset, obs + game * SH::kObsStride);   // mover's seat seen from there
  }
}

// The other HLE observation types from the same state (HanabiGame::AgentObservationType; only
// kCardKnowledge is used by the reference's environments, run_experiment.py:114-128):
//   0 kMinimal  the canonical row without its last section, the card knowledge;
//   2 kSeer     the observer's own cards are encoded too, as the FIRST block of the hands section
//               (EncodeHands with show_own_cards): the canonical row moves up by H * BPC bits and the
//               one-hots of seat 0's slots fill the gap.
// Rows are assembled in registers; `stride` (words) comes from the host (16-byte multiples).
template <class SH>
__global__ void __launch_bounds__(kThreads)
env_observe_as_kernel(Rt rt, const uint32_t* __restrict__ state, int offset, int obs_type,
                      uint32_t* __restrict__ obs, int stride) {
  const long long game = game_index();
  if (game >= rt.n_games) return;
  Game<SH> g;
  load_game(g, state, game);
  rotate_seats(g, offset);
  uint32_t row[SH::kObsStride];
  encode_row(g, SH::P - 1 - offset, row);
  uint32_t* dst = obs + game * stride;
  if (obs_type == 2) {
    constexpr int kShift = SH::H * SH::BPC;                     // bits the canonical row moves up by
    constexpr int kWords = (SH::kObsLen + kShift + 31) / 32;
    constexpr int ws = kShift / 32, bs = kShift % 32;
    uint32_t out[kWords];
#pragma unroll
    for (int j = 0; j < kWords; ++j) {
      uint32_t v = 0;
      if (j - ws >= 0 && j - ws < SH::kObsStride) v = row[j - ws] << bs;
      if (bs != 0 && j - ws - 1 >= 0 && j - ws - 1 < SH::kObsStride) v |= row[j - ws - 1] >> (32 - bs);
      out[j] = v;
    }
#pragma unroll
    for (int i = 0; i < SH::H; ++i) {                           // own cards: slot i at bit i * BPC
      const unsigned long long oh = (unsigned long long)(g.oh[0][i] & SH::kGridAll) << ((i * SH::BPC) % 32);
      constexpr int dummy = 0; (void)dummy;
      const int w0 = (i * SH::BPC) / 32;
      out[w0] |= (uint32_t)oh;
      if (w0 + 1 < kWords) out[w0 + 1] |= (uint32_t)(oh >> 32);
    }
    for (int j = 0; j < stride; ++j) dst[j] = j < kWords ? out[j] : 0u;
  } else {
    const int bits = obs_type == 0 ? SH::kOffKnow : SH::kObsLen;
    for (int j = 0; j < stride; ++j) {
      uint32_t v = j < SH::kObsStride ? row[j] : 0u;
      const int lo = 32 * j;
      if (lo >= bits) v = 0u;
      else if (lo + 32 > bits) v &= (1u << (bits - lo)) - 1u;
      dst[j] = v;
    }
  }
}

// RulebasedAgent.explore over the whole batch (rule_based.py:27-31)
template <class SH, bool EXT>
__global__ void __launch_bounds__(kThreads)
rules_act_kernel(Rt rt, const uint32_t* __restrict__ state, RuleSet rs, unsigned long long step,
                 int32_t* __restrict__ action, int32_t* __restrict__ rule_hit,
                 unsigned long long* __restrict__ hist, long long first, long long count) {
  __shared__ unsigned long long sthr[kThrSlots * kThrRows];
  stage_thresholds(rs, sthr);
  const long long game = first + game_index();
  if (game >= first + count) return;
  Game<SH> g;
  load_game(g, state, game);
  int a = 0, hit = -1;
  if (g.step_type != LAST)
    a = rules_act<SH, EXT>(g, rs, sthr, (unsigned long long)(rt.game_id_base + game), step, &hit);
  action[game] = a;
  if (rule_hit != nullptr) rule_hit[game] = hit;
  // rule-hit histogram: one global atomic per (warp, rule that fired in it) -- one per game put
  // 65,536 atomics on ten addresses and made this kernel 27 us (profiles/r01_launches_bench_v5.csv)
  if (hist != nullptr) {
    const unsigned int peers = __match_any_sync(__activemask(), hit);
    if (hit >= 0 && (threadIdx.x & 31) == __ffs((int)peers) - 1)
      atomicAdd(&hist[hit], (unsigned long long)__popc(peers));
  }
}

// ---- work queue of the fused self-play kernel ---------------------------------------------------
// 65,536 games are 2048 warps on 592 schedulers: an SM's sub-partitions hold 4 / 4 / 3 / 3 of them, the
// ones with four set the duration of a launch and the others idle for the last fifth of it
// (profiles/r01_selfplay_history.md, v6).  Consecutive launches are therefore allowed to OVERLAP
// (programmatic dependent launch: the next launch's CTAs become resident as soon as slots are free) and
// the block index no longer names the games: a CTA takes the next group of 64 games from a FIFO ring that
// the previous launch's CTAs refill as they finish.  A group that finished early (light sub-partition)
// starts its next 64 steps early, wherever a slot is free, so over a run of launches every group sees
// the average load instead of some groups always the worst.  Every launch pops each group exactly once
// (FIFO: the i-th CTA of launch k+1 to start takes the i-th group launch k finished; the push index is
// counted per launch, because a CTA of launch k+1 may well finish before the last CTA of launch k);
// results do not depend on the order.
constexpr int kSpqLaunches = 8;   // push counters: launch seq uses counter seq % 8
struct SpQueue {
  unsigned long long* head;     // next pop ticket (launch `seq` takes tickets [seq * G, (seq + 1) * G))
  unsigned long long* pushes;   // [kSpqLaunches] groups handed on so far by the launches of each residue
  int* group;                   // [G] ring: group ids
  int* state;                   // [G] 2r+1: filled for launch r, 2r+2: taken by launch r
  int n_groups;
  unsigned long long seq;       // number of this launch (self-play launches of this env so far)
};

__global__ void spq_init_kernel(SpQueue q) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) *q.head = 0ull;
  if (i < kSpqLaunches) q.pushes[i] = 0ull;
  if (i < q.n_groups) { q.group[i] = i; q.state[i] = 1; }
}

__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
// bounded wait (a queue that never fills is a bug: trap instead of hanging the device)
__device__ __forceinline__ void spq_wait(const int* p, int want) {
  for (long long spins = 0; ld_acquire_gpu(p) != want; ++spins) {
    __nanosleep(64);
    if (spins > (1ll << 24)) __trap();
  }
}

// Fused self-play: `n_steps` x {rule-based move, env step, observation} with the game state
// held in registers; only the per-step outputs touch HBM.  Output row (t, game).
// Pass t of the loop is {observation of state t -> row t-1, rule bodies on state t} as ONE basic block
// (both only read the state; stores are predicated, not branched around), then {rule list, transition
// t -> row t}.  The observation of state 0 was written by the previous launch (pass 0 stores
// nothing); pass n_steps only writes the last observation.
template <class SH, bool EXT>
__global__ void __launch_bounds__(kThreads, 7)
selfplay_kernel(Rt rt, uint32_t* __restrict__ state, const uint8_t* __restrict__ decks, RuleSet rs,
                unsigned long long step0, int n_steps, uint32_t* __restrict__ obs,
                uint32_t* __restrict__ legal, float* __restrict__ reward,
                uint8_t* __restrict__ step_type, int32_t* __restrict__ action,
                unsigned long long* __restrict__ hist, unsigned long long* illegal, SpQueue q) {
  __shared__ unsigned long long sthr[kThrSlots * kThrRows];
  __shared__ EncTables tabs;
  __shared__ int s_group;
  unsigned long long ticket = 0;
  if (threadIdx.x == 0) ticket = atomicAdd(q.head, 1ull);
  __syncthreads();
  // every CTA of this launch has its ticket before the next launch may start taking tickets
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (threadIdx.x == 0) {
    const int slot = (int)(ticket % (unsigned long long)q.n_groups);
    const int filled = 2 * (int)(ticket / (unsigned long long)q.n_groups) + 1;
    spq_wait(q.state + slot, filled);
    s_group = __ldcg(q.group + slot);
    st_release_gpu(q.state + slot, filled + 1);
  }
  // rule-hit histogram: per-CTA counters in shared memory (one shared-memory atomic per game step on
  // the load/store pipe; packed 8-bit counters in registers were 20 ALU instructions per step)
  __shared__ unsigned int shist[kMaxRules + 1];
  __shared__ MoveTable mtab;
  stage_enc_tables<SH>(tabs);
  stage_move_table<SH>(mtab);
  if (threadIdx.x <= kMaxRules) shist[threadIdx.x] = 0;
  stage_thresholds(rs, sthr);                                   // ends with the barrier
  const int group = s_group;
  const long long game = (long long)group * kThreads + threadIdx.x;
  const bool live = game < rt.n_games;
  const unsigned long long game_id = (unsigned long long)(rt.game_id_base + game);
  if (live) {
  Game<SH> g;
  load_game<SH, true>(g, state, game);   // L2 only: the previous launch may still be running on this SM
  unsigned long long row = deck_row<SH>(rt, game, g.episode);   // one 64-bit modulo per launch
  const uint8_t* deck = decks + row * kDeckPitch;
#pragma unroll 1
  for (int t = 0;; ++t) {
    const long long orow = (long long)(t - 1) * rt.n_games + game;     // outputs of state t
    const bool out = t > 0;
    const uint32_t next_card = peek_card(g, deck);
    const unsigned long long lm = legal_mask(g);
    {
      uint32_t orow_w[SH::kObsStride];
      encode_row(g, SH::P - 1, orow_w, &tabs);
      uint32_t* dst = obs + orow * SH::kObsStride;
#pragma unroll
      for (int i = 0; i < SH::kObsStride / 4; ++i)
        st_v4_if(out && obs != nullptr, dst + 4 * i, orow_w[4 * i], orow_w[4 * i + 1], orow_w[4 * i + 2],
                 orow_w[4 * i + 3]);
    }
    st_u32_if(out && legal != nullptr, legal + orow * SH::kLegalWords, (uint32_t)lm);
    if (SH::kLegalWords > 1)
      st_u32_if(out && legal != nullptr, legal + orow * SH::kLegalWords + 1, (uint32_t)(lm >> 32));
    st_u8_if(out && step_type != nullptr, step_type + orow, (uint32_t)g.step_type);
    Cores co;
    rule_cores(g, rs.used, co);
    if (t == n_steps) break;

    int a = 0, hit = -1;
    if (g.step_type != LAST) a = rules_resolve<SH, EXT>(g, co, rs, sthr, game_id, step0 + t, &hit);
    const float r = step_game(g, rt, decks, deck, row, rt.row_step, a, next_card, lm, illegal, &mtab);
    const long long arow = (long long)t * rt.n_games + game;
    if (action != nullptr) action[arow] = a;
    if (reward != nullptr) reward[arow] = r;
    if (hit >= 0) atomicAdd(&shist[hit], 1u);
  }
  store_game(g, state, game);
  }
  __threadfence();                                              // the group's state, device-wide ...
  __syncthreads();
  if (hist != nullptr && threadIdx.x <= rs.n && shist[threadIdx.x] != 0)
    atomicAdd(&hist[threadIdx.x], (unsigned long long)shist[threadIdx.x]);
  if (threadIdx.x == 0) {                                       // ... before the group is handed on
    // i-th group this launch finishes -> slot i, for launch seq + 1.  The counter of residue seq % 8 has
    // seen (seq / 8) * G pushes once launch seq - 8 is through (it nearly always is: bounded wait).
    unsigned long long* ctr = q.pushes + (q.seq % kSpqLaunches);
    const unsigned long long base = (q.seq / kSpqLaunches) * (unsigned long long)q.n_groups;
    for (long long spins = 0; *(volatile unsigned long long*)ctr < base; ++spins) {
      __nanosleep(64);
      if (spins > (1ll << 24)) __trap();
    }
    const int slot = (int)(atomicAdd(ctr, 1ull) - base);
    const int next = (int)(q.seq + 1);
    spq_wait(q.state + slot, 2 * next);                         // this launch has taken the slot's last entry
    __stcg(q.group + slot, group);
    st_release_gpu(q.state + slot, 2 * next + 1);
  }
}

// canonical int32 dump, one thread per game, ABSOLUTE player order (tests; layout = oracle
// ho_env_dump).  The state is seat-relative: absolute player p sits at seat (p - cur) mod P.
template <class SH>
__global__ void env_dump_kernel(Rt rt, const uint32_t* __restrict__ state, int32_t* __restrict__ out,
                                int dump_len) {
  const long long game = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (game >= rt.n_games) return;
  const uint32_t* w = state + game * SH::kStateWords;
  const uint32_t a = w[SH::S], b = w[SH::S + 1];
  const unsigned long long disc = (unsigned long long)w[SH::S + 2] | ((unsigned long long)w[SH::S + 3] << 32);
  const uint32_t last = w[SH::S + 4];
  int32_t* o = out + game * dump_len;
  int k = 0;
  const int life = (a >> 29) & 3, cur = (b >> 21) & 7;
  const int score = __popc(a & kGridMask);
  o[k++] = cur; o[k++] = (a >> 25) & 15; o[k++] = life;
  o[k++] = SH::DECK - (int)(b & 63); o[k++] = (b >> 24) & 7; o[k++] = (b >> 27) & 3;
  o[k++] = life <= 0 ? 0 : score; o[k++] = (int32_t)(w[SH::S + 5] & 0x7fffffff);
  for (int c = 0; c < SH::C; ++c) o[k++] = __popc((a >> (5 * c)) & 31u);
  for (int p = 0; p < SH::P; ++p) {
    const int seat = (p - cur + SH::P) % SH::P;
    o[k++] = (b >> (6 + 3 * seat)) & 7;
  }
  for (int p = 0; p < SH::P; ++p) {
    const int seat = (p - cur + SH::P) % SH::P;
    for (int i = 0; i < SH::H; ++i) {
      const uint32_t sw = w[seat * SH::H + i];
      const bool have = slot_valid(sw);
      const uint32_t grid = slot_grid(sw);
      const int card = slot_card(sw);
      int cp = 0, rp = 0;
      for (int c = 0; c < 5; ++c) {
        const uint32_t rowbits = (grid >> (5 * c)) & 31u;
        if (rowbits) cp |= 1 << c;
        rp |= (int)rowbits;
      }
      o[k++] = have ? card : -1;
      o[k++] = have ? cp : 0;
      o[k++] = have ? rp : 0;
      o[k++] = have && slot_chint(sw) ? card_color(card) : -1;
      o[k++] = have && slot_rhint(sw) ? card_rank(card) : -1;
    }
  }
  for (int t = 0; t < SH::BPC; ++t) o[k++] = (int)((disc >> (2 * t)) & 3ull);
  if (last & 1u) {
    const int uid = (last >> 1) & 63;
    const bool card_move = uid < 2 * SH::H;
    int type, idx = -1, off = -1, mc = -1, mr = -1;
    if (card_move) {
      type = uid >= SH::H ? PLAY : DISCARD;
      idx = uid >= SH::H ? uid - SH::H : uid;
    } else {
      int u = uid - 2 * SH::H;
      const bool by_color = u < (SH::P - 1) * SH::C;
      if (!by_color) u -= (SH::P - 1) * SH::C;
      const int span = by_color ? SH::C : SH::R;
      off = 1 + u / span;
      type = by_color ? REVEAL_COLOR : REVEAL_RANK;
      if (by_color) mc = u % span; else mr = u % span;
    }
    const int card = (last >> 7) & 31;
    o[k++] = 1; o[k++] = (cur + SH::P - 1) % SH::P; o[k++] = type; o[k++] = idx; o[k++] = off;
    o[k++] = mc; o[k++] = mr;
    o[k++] = card_move ? card_color(card) : -1;
    o[k++] = card_move ? card_rank(card) : -1;
    o[k++] = (last >> 12) & 1; o[k++] = (last >> 13) & 1; o[k++] = (last >> 14) & 31;
  } else {
    for (int j = 0; j < 12; ++j) o[k++] = 0;
  }
}

// bit-packed rows -> one byte per bit ([n, bits] int8, the reference's layout:
// experience_buffer.py:9-12 stores obs / legal moves as np.byte)
__global__ void unpack_bits_kernel(const uint32_t* __restrict__ packed, int stride_words, int bits,
                                   int8_t* __restrict__ out, long long n) {
  const long long total = n * bits;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / bits;
    const int b = (int)(i - row * bits);
    out[i] = (int8_t)((packed[row * stride_words + (b >> 5)] >> (b & 31)) & 1u);
  }
}

}  // namespace

// ------------------------------------------------------------------ host side ----
namespace {

// The kernels are instantiated for the game shapes the reference's environments produce
// (Hanabi-Full 2..5 players, Hanabi-Small, Hanabi-Very-Small; run_experiment.py:115-128).
#define HB_SHAPES(X)                  \
  X(0, 2, 5, 5, 5, 8, 3)              \
  X(1, 3, 5, 5, 5, 8, 3)              \
  X(2, 4, 4, 5, 5, 8, 3)              \
  X(3, 5, 4, 5, 5, 8, 3)              \
  X(4, 2, 2, 2, 5, 3, 1)              \
  X(5, 2, 2, 1, 5, 3, 1)

int find_shape(int P, int H, int C, int R, int MI, int ML) {
#define X(id, p, h, c, r, mi, ml) \
  if (P == p && H == h && C == c && R == r && MI == mi && ML == ml) return id;
  HB_SHAPES(X)
#undef X
  return -1;
}

#define HB_DISPATCH(env, CALL)                                                   \
  switch ((env)->shape) {                                                        \
    case 0: { using SH = Shape<2, 5, 5, 5, 8, 3>; CALL; break; }                       \
    case 1: { using SH = Shape<3, 5, 5, 5, 8, 3>; CALL; break; }                       \
    case 2: { using SH = Shape<4, 4, 5, 5, 8, 3>; CALL; break; }                       \
    case 3: { using SH = Shape<5, 4, 5, 5, 8, 3>; CALL; break; }                       \
    case 4: { using SH = Shape<2, 2, 2, 5, 3, 1>; CALL; break; }                       \
    case 5: { using SH = Shape<2, 2, 1, 5, 3, 1>; CALL; break; }                       \
    default: hb::set_error("unsupported game shape"); return HB_ERR_INVALID;     \
  }

inline unsigned grid_for(const hb_env* e) {
  return (unsigned)((e->rt.n_games + kThreads - 1) / kThreads);
}

int env_stage(hb_env* e, size_t bytes) {
  if (e->stage_bytes >= bytes) return HB_OK;
  if (e->stage) cudaFree(e->stage);
  e->stage = nullptr;
  e->stage_bytes = 0;
  HB_CUDA(cudaMalloc(&e->stage, bytes));
  e->stage_bytes = bytes;
  return HB_OK;
}

// rule lists with one of the rules that are dead in the reference run the EXT kernels
inline bool uses_ext_rules(const RuleSet& rs) {
  for (int k = 0; k < rs.n; ++k)
    if (rs.id[k] >= R_TELL_ANYONE_USELESS_CARD) return true;
  return false;
}

int fill_ruleset(RuleSet& rs, const int32_t* ids, const double* farg, const int32_t* iarg, int n,
                 unsigned long long seed) {
  HB_REQUIRE(n >= 0 && n <= kMaxRules, "rule list too long (%d > %d)", n, kMaxRules);
  HB_REQUIRE(n == 0 || ids != nullptr, "rule ids are NULL");
  rs.n = n;
  rs.seed = seed;
  rs.used = 1u << R_LEGAL_RANDOM;
  for (int k = 0; k < n; ++k) {
    HB_REQUIRE(ids[k] >= 0 && ids[k] < R_NUM_RULES, "unknown rule id %d", ids[k]);
    rs.id[k] = ids[k];
    rs.used |= 1u << ids[k];
    if (ids[k] == R_OSAWA_DISCARD && iarg && iarg[k] != 0) rs.used |= kOsawaIntended;   // applies to the whole list
    rs.farg[k] = farg ? farg[k] : 0.0;
    rs.iarg[k] = iarg ? iarg[k] : 0;
  }
  rs.id[n] = R_LEGAL_RANDOM;   // Ruleset.legal_random fallback (rule_based.py:24)
  rs.farg[n] = 0.0;
  rs.iarg[n] = 0;
  // comparison tables of the threshold rules (the reference's float64 division, done here once) and
  // the rows of rules_resolve
  int slots = 0, td_rows = 0, td_max = -1;
  rs.lazy = 0;
  for (int s = 0; s < kThrSlots; ++s) rs.slot_extra[s] = 0;
  for (int r = 0; r < 16; ++r) rs.key_hi[r] = kKeyAbsent;
  rs.td_ilt[0] = rs.td_ilt[1] = 0;
  auto own = [&](int row, int k) { if (rs.key_hi[row] == kKeyAbsent) rs.key_hi[row] = (unsigned int)k << 8; };
  for (int k = 0; k <= n; ++k) {
    rs.tslot[k] = -1;
    const int id = rs.id[k];
    const bool thr_rule = id == R_PLAY_PROBABLY_SAFE || id == R_DISCARD_PROBABLY_USELESS;
    if (thr_rule && slots < kThrSlots) {
      for (int t = 0; t < kThrRows; ++t) {
        unsigned long long m = 0;
        for (int p = 0; p < 64 && t > 0; ++p) {
          volatile double q = (double)p / (double)t;    // rounded to double, no extended precision
          if (q >= rs.farg[k]) m |= 1ull << p;
        }
        rs.thr[slots][t] = m;
      }
      rs.tslot[k] = slots++;
    }
    switch (id) {
      case R_DISCARD_OLDEST_FIRST: own(ROW_DISCARD_OLDEST, k); break;
      case R_OSAWA_DISCARD: own(ROW_OSAWA, k); break;
      case R_TELL_UNKNOWN: own(ROW_TELL_UNKNOWN, k); break;
      case R_PLAY_SAFE_CARD: own(ROW_PLAY_SAFE, k); break;
      case R_PLAY_IF_CERTAIN: own(ROW_PLAY_CERTAIN, k); break;
      case R_TELL_PLAYABLE_CARD_OUTER:
      case R_TELL_ANYONE_USEFUL_CARD: own(ROW_TELL_PLAYABLE, k); break;
      case R_HAIL_MARY: own(ROW_HAIL_MARY, k); break;
      case R_TELL_RANDOMLY: own(ROW_TELL_RANDOMLY, k); break;
      case R_DISCARD_RANDOMLY: own(ROW_DISCARD_RANDOMLY, k); break;
      case R_TELL_MOST_INFORMATION: break;               // never fires
      case R_TELL_DISPENSABLE:
        if (rs.iarg[k] <= td_max) break;                 // an earlier one fires whenever this would
        if (td_rows < 2) {
          rs.td_ilt[td_rows] = rs.iarg[k];
          own(ROW_TELL_DISPENSABLE0 + td_rows, k);
          ++td_rows;
        } else {
          rs.lazy |= 1u << k;
        }
        td_max = rs.iarg[k];
        break;
      case R_PLAY_PROBABLY_SAFE:
        if (rs.tslot[k] >= 0 && rs.farg[k] > 0.0) {      // decided by the table alone
          rs.slot_extra[rs.tslot[k]] = rs.iarg[k] != 0;
          own(ROW_PPS0 + rs.tslot[k], k);
        } else {
          rs.lazy |= 1u << k;
        }
        break;
      default:                                           // random rules, rules dead in the reference
        rs.lazy |= 1u << k;
        break;
    }
  }
  rs.n_slots = slots;
  return HB_OK;
}

// The RuleSet of a rule list, built on first sight and kept with the env (hb_env::rule_set).
int cached_ruleset(hb_env* e, const int32_t* ids, const double* farg, const int32_t* iarg, int n,
                   unsigned long long seed, const RuleSet** out, bool* ext) {
  HB_REQUIRE(n >= 0 && n <= kMaxRules, "rule list too long (%d > %d)", n, kMaxRules);
  HB_REQUIRE(n == 0 || ids != nullptr, "rule ids are NULL");
  hb_env::RuleKey key;
  key.n = n;
  key.seed = seed;
  for (int k = 0; k < n; ++k) {
    key.ids[k] = ids[k];
    key.iarg[k] = iarg ? iarg[k] : 0;
    key.farg[k] = farg ? farg[k] : 0.0;
  }
  for (int c = 0; c < e->rule_cached; ++c) {
    const hb_env::RuleKey& k = e->rule_key[c];
    if (k.n == n && k.seed == seed && memcmp(k.ids, key.ids, sizeof(key.ids)) == 0 &&
        memcmp(k.iarg, key.iarg, sizeof(key.iarg)) == 0 && memcmp(k.farg, key.farg, sizeof(key.farg)) == 0) {
      *out = &e->rule_set[c];
      *ext = e->rule_ext[c];
      return HB_OK;
    }
  }
  const int slot = e->rule_next;
  e->rule_key[slot].n = -1;                                  // not a hit while half built
  int rc = fill_ruleset(e->rule_set[slot], ids, farg, iarg, n, seed);
  if (rc) return rc;
  e->rule_key[slot] = key;
  e->rule_ext[slot] = uses_ext_rules(e->rule_set[slot]);
  e->rule_next = (slot + 1) % hb_env::kRuleCache;
  if (e->rule_cached < hb_env::kRuleCache) ++e->rule_cached;
  ++e->rule_fills;
  *out = &e->rule_set[slot];
  *ext = e->rule_ext[slot];
  return HB_OK;
}

}  // namespace

extern "C" {

int hb_env_create(int players, int colors, int ranks, int hand_size, int max_info, int max_life,
                  int64_t n_games, int device, hb_env** out) {
  HB_REQUIRE(out != nullptr, "hb_env_create: out is NULL");
  HB_REQUIRE(n_games >= 1, "hb_env_create: n_games must be >= 1");
  const int shape = find_shape(players, hand_size, colors, ranks, max_info, max_life);
  HB_REQUIRE(shape >= 0,
             "hb_env_create: unsupported game players=%d hand=%d colors=%d ranks=%d info=%d life=%d "
             "(kernels exist for Hanabi-Full 2-5p, Hanabi-Small and Hanabi-Very-Small)",
             players, hand_size, colors, ranks, max_info, max_life);
  hb::DeviceGuard g(device);
  if (!g.ok) { hb::set_error("hb_env_create: cannot select CUDA device %d", device); return HB_ERR_CUDA; }
  hb_env* e = new (std::nothrow) hb_env();
  HB_REQUIRE(e != nullptr, "hb_env_create: out of host memory");
  e->device = device;
  e->P = players; e->H = hand_size; e->C = colors; e->R = ranks; e->shape = shape;
  const int bpc = colors * ranks;
  int per_color = 0;
  for (int r = 0; r < ranks; ++r) per_color += r == 0 ? 3 : (r == ranks - 1 ? 1 : 2);
  e->deck_size = colors * per_color;
  e->num_moves = 2 * hand_size + (players - 1) * (colors + ranks);
  e->legal_words = (e->num_moves + 31) / 32;
  e->state_words = (players * hand_size + 6 <= 16) ? 16 : 32;
  Rt& rt = e->rt;
  const int hands_len = (players - 1) * hand_size * bpc + players;
  const int board_len = (e->deck_size - players * hand_size) + bpc + max_info + max_life;
  const int last_len = players + 4 + players + colors + ranks + 2 * hand_size + bpc + 2;
  const int know_len = players * hand_size * (bpc + colors + ranks);
  e->obs_len = hands_len + board_len + e->deck_size + last_len + know_len;
  e->obs_stride = (((e->obs_len + 31) / 32) + 3) / 4 * 4;
  rt.n_games = n_games; rt.n_total = n_games; rt.game_id_base = 0; rt.n_decks = 0;
  cudaError_t err = cudaMalloc(&e->state, (size_t)n_games * e->state_words * 4);
  if (err == cudaSuccess) err = cudaMemset(e->state, 0, (size_t)n_games * e->state_words * 4);
  if (err == cudaSuccess) err = cudaMalloc(&e->illegal, sizeof(unsigned long long));
  if (err == cudaSuccess) err = cudaMemset(e->illegal, 0, sizeof(unsigned long long));
  if (err != cudaSuccess) {
    if (e->state) cudaFree(e->state);
    if (e->illegal) cudaFree(e->illegal);
    delete e;
    return hb::cuda_fail(err, "hb_env_create alloc", __FILE__, __LINE__);
  }
  *out = e;
  return HB_OK;
}

int hb_env_destroy(hb_env* e) {
  if (!e) return HB_OK;
  hb::DeviceGuard g(e->device);
  if (e->state) cudaFree(e->state);
  if (e->decks) cudaFree(e->decks);
  if (e->illegal) cudaFree(e->illegal);
  if (e->spq) cudaFree(e->spq);
  if (e->stage) cudaFree(e->stage);
  if (e->hs_step) cudaStreamDestroy(e->hs_step);
  if (e->hs_out) cudaStreamDestroy(e->hs_out);
  if (e->hs_act) cudaStreamDestroy(e->hs_act);
  for (int c = 0; c < hb_env::kMaxChunks; ++c)
    if (e->hev_step[c]) cudaEventDestroy(e->hev_step[c]);
  if (e->hev_done) cudaEventDestroy(e->hev_done);
  if (e->hev_out[0]) cudaEventDestroy(e->hev_out[0]);
  if (e->hev_out[1]) cudaEventDestroy(e->hev_out[1]);
  if (e->hev_rules) cudaEventDestroy(e->hev_rules);
  if (e->hev_acts) cudaEventDestroy(e->hev_acts);
  delete e;
  return HB_OK;
}

int64_t hb_env_info(const hb_env* e, int what) {
  if (!e) return -1;
  switch (what) {
    case 0: return e->obs_len;
    case 1: return e->obs_stride;
    case 2: return e->num_moves;
    case 3: return e->legal_words;
    case 4: return e->deck_size;
    case 5: return e->state_words;
    case 6: return e->rt.n_games;
    case 7: return 8 + e->C + e->P + 5 * e->P * e->H + e->C * e->R + 12;  // dump length
    case 8: return e->P;
    case 9: return (int64_t)e->rule_fills;
    default: return -1;
  }
}

int hb_env_set_shard(hb_env* e, int64_t game_id_base, int64_t n_total) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && game_id_base >= 0 && n_total >= game_id_base + e->rt.n_games,
             "hb_env_set_shard: bad shard [%lld, +%lld) of %lld", (long long)game_id_base,
             e ? (long long)e->rt.n_games : 0ll, (long long)n_total);
  e->rt.game_id_base = game_id_base;
  e->rt.n_total = n_total;
  if (e->rt.n_decks > 0) e->rt.row_step = n_total % e->rt.n_decks;
  return HB_OK;
}

int hb_env_set_decks(hb_env* e, const uint8_t* d_decks, int64_t n_decks, void* stream) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && d_decks && n_decks >= 1, "hb_env_set_decks: bad arguments");
  hb::DeviceGuard g(e->device);
  if (e->decks) { HB_CUDA(cudaStreamSynchronize((cudaStream_t)stream)); cudaFree(e->decks); e->decks = nullptr; }
  // one deck per kDeckPitch-byte row (aligned 16-byte loads of the initial deal; the tail is zero)
  const size_t bytes = (size_t)n_decks * kDeckPitch;
  HB_CUDA(cudaMalloc(&e->decks, bytes));
  HB_CUDA(cudaMemsetAsync(e->decks, 0, bytes, (cudaStream_t)stream));
  HB_CUDA(cudaMemcpy2DAsync(e->decks, kDeckPitch, d_decks, e->deck_size, e->deck_size, (size_t)n_decks,
                            cudaMemcpyDefault, (cudaStream_t)stream));
  e->rt.n_decks = n_decks;
  e->rt.row_step = e->rt.n_total % n_decks;
  return HB_OK;
}

int hb_env_set_decks_host(hb_env* e, const uint8_t* h_decks, int64_t n_decks) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && h_decks && n_decks >= 1, "hb_env_set_decks_host: bad arguments");
  // every deck must be a permutation of the card multiset (HLE deals from exactly that multiset)
  const int bpc = e->C * e->R;
  for (int64_t d = 0; d < n_decks; ++d) {
    int cnt[32] = {0};
    for (int i = 0; i < e->deck_size; ++i) {
      const int c = h_decks[d * e->deck_size + i];
      HB_REQUIRE(c < bpc, "deck %lld holds card index %d >= %d", (long long)d, c, bpc);
      cnt[c]++;
    }
    for (int c = 0; c < bpc; ++c) {
      const int r = c % e->R;
      const int want = r == 0 ? 3 : (r == e->R - 1 ? 1 : 2);
      HB_REQUIRE(cnt[c] == want, "deck %lld is not a permutation of the card multiset", (long long)d);
    }
  }
  int rc = hb_env_set_decks(e, h_decks, n_decks, nullptr);
  if (rc) return rc;
  hb::DeviceGuard g(e->device);
  HB_CUDA(cudaStreamSynchronize(nullptr));
  return HB_OK;
}

int hb_env_reset(hb_env* e, uint32_t* d_obs, uint32_t* d_legal, uint8_t* d_step_type, void* stream) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e, "hb_env_reset: NULL env");
  if (!e->decks) { hb::set_error("hb_env_reset: call hb_env_set_decks first"); return HB_ERR_STATE; }
  hb::DeviceGuard g(e->device);
  HB_DISPATCH(e, (env_reset_kernel<SH><<<grid_for(e), kThreads, 0, (cudaStream_t)stream>>>(
                     e->rt, e->state, e->decks, d_obs, d_legal, d_step_type)));
  HB_CUDA(cudaGetLastError());
  e->is_reset = true;
  return HB_OK;
}

int hb_env_step(hb_env* e, const int32_t* d_action, uint32_t* d_obs, uint32_t* d_legal,
                float* d_reward, uint8_t* d_step_type, void* stream) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && d_action, "hb_env_step: bad arguments");
  if (!e->is_reset) { hb::set_error("hb_env_step: reset the environment first"); return HB_ERR_STATE; }
  hb::DeviceGuard g(e->device);
  HB_DISPATCH(e, (env_step_kernel<SH><<<grid_for(e), kThreads, 0, (cudaStream_t)stream>>>(
                     e->rt, e->state, e->decks, d_action, d_obs, d_legal, d_reward, d_step_type,
                     e->illegal, 0, e->rt.n_games, SH::kObsStride)));
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_env_observe(hb_env* e, int offset, uint32_t* d_obs, uint32_t* d_legal, void* stream) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && offset >= 0 && offset < e->P, "hb_env_observe: bad arguments");
  if (!e->is_reset) { hb::set_error("hb_env_observe: reset the environment first"); return HB_ERR_STATE; }
  hb::DeviceGuard g(e->device);
  HB_DISPATCH(e, (env_observe_kernel<SH><<<grid_for(e), kThreads, 0, (cudaStream_t)stream>>>(
                     e->rt, e->state, offset, d_obs, d_legal)));
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int64_t hb_env_obs_len_as(const hb_env* e, int obs_type) {
  if (!e || obs_type < 0 || obs_type > 2) return -1;
  const int bpc = e->C * e->R;
  if (obs_type == 0) return e->obs_len - e->P * e->H * (bpc + e->C + e->R);
  if (obs_type == 2) return e->obs_len + e->H * bpc;
  return e->obs_len;
}

int hb_env_observe_as(hb_env* e, int obs_type, int offset, uint32_t* d_obs, int stride_words, void* stream) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && d_obs && offset >= 0 && offset < e->P && obs_type >= 0 && obs_type <= 2,
             "hb_env_observe_as: bad arguments (obs_type 0 minimal, 1 card knowledge, 2 seer)");
  HB_REQUIRE((int64_t)stride_words * 32 >= hb_env_obs_len_as(e, obs_type),
             "hb_env_observe_as: rows of %d words are shorter than %lld bits", stride_words,
             (long long)hb_env_obs_len_as(e, obs_type));
  if (!e->is_reset) { hb::set_error("hb_env_observe_as: reset the environment first"); return HB_ERR_STATE; }
  hb::DeviceGuard g(e->device);
  HB_DISPATCH(e, (env_observe_as_kernel<SH><<<grid_for(e), kThreads, 0, (cudaStream_t)stream>>>(
                     e->rt, e->state, offset, obs_type, d_obs, stride_words)));
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_env_illegal_count(hb_env* e, int64_t* out, void* stream) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && out, "hb_env_illegal_count: bad arguments");
  hb::DeviceGuard g(e->device);
  unsigned long long v = 0;
  HB_CUDA(cudaMemcpyAsync(&v, e->illegal, sizeof(v), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  HB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  *out = (int64_t)v;
  return HB_OK;
}

int hb_env_dump(hb_env* e, int32_t* d_out, void* stream) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && d_out, "hb_env_dump: bad arguments");
  hb::DeviceGuard g(e->device);
  const int len = (int)hb_env_info(e, 7);
  const unsigned blocks = (unsigned)((e->rt.n_games + 127) / 128);
  HB_DISPATCH(e, (env_dump_kernel<SH><<<blocks, 128, 0, (cudaStream_t)stream>>>(e->rt, e->state, d_out, len)));
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_env_unpack_obs(hb_env* e, const uint32_t* d_packed, int8_t* d_out, int64_t n, void* stream) {
  HB_REQUIRE(e && d_packed && d_out && n >= 0, "hb_env_unpack_obs: bad arguments");
  if (n == 0) return HB_OK;
  hb::DeviceGuard g(e->device);
  const long long total = n * e->obs_len;
  const unsigned blocks = (unsigned)min((long long)148 * 16, (total + 255) / 256);
  unpack_bits_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(d_packed, e->obs_stride,
                                                               e->obs_len, d_out, n);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_env_unpack_legal(hb_env* e, const uint32_t* d_legal, int8_t* d_out, int64_t n, void* stream) {
  HB_REQUIRE(e && d_legal && d_out && n >= 0, "hb_env_unpack_legal: bad arguments");
  if (n == 0) return HB_OK;
  hb::DeviceGuard g(e->device);
  const long long total = n * e->num_moves;
  const unsigned blocks = (unsigned)min((long long)148 * 16, (total + 255) / 256);
  unpack_bits_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(d_legal, e->legal_words, e->num_moves,
                                                               d_out, n);
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_rules_act(hb_env* e, const int32_t* rule_ids, const double* rule_farg, const int32_t* rule_iarg,
                 int n_rules, uint64_t seed, uint64_t step, int32_t* d_action, int32_t* d_rule_hit,
                 uint64_t* d_hist, void* stream) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && d_action, "hb_rules_act: bad arguments");
  if (!e->is_reset) { hb::set_error("hb_rules_act: reset the environment first"); return HB_ERR_STATE; }
  const RuleSet* rsp = nullptr;
  bool ext = false;
  int rc = cached_ruleset(e, rule_ids, rule_farg, rule_iarg, n_rules, seed, &rsp, &ext);
  if (rc) return rc;
  const RuleSet& rs = *rsp;
  hb::DeviceGuard g(e->device);
  if (ext) {
    HB_DISPATCH(e, (rules_act_kernel<SH, true><<<grid_for(e), kThreads, 0, (cudaStream_t)stream>>>(
                       e->rt, e->state, rs, step, d_action, d_rule_hit,
                       (unsigned long long*)d_hist, 0, e->rt.n_games)));
  } else {
    HB_DISPATCH(e, (rules_act_kernel<SH, false><<<grid_for(e), kThreads, 0, (cudaStream_t)stream>>>(
                       e->rt, e->state, rs, step, d_action, d_rule_hit,
                       (unsigned long long*)d_hist, 0, e->rt.n_games)));
  }
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

int hb_env_selfplay(hb_env* e, const int32_t* rule_ids, const double* rule_farg,
                    const int32_t* rule_iarg, int n_rules, uint64_t seed, uint64_t step0, int n_steps,
                    uint32_t* d_obs, uint32_t* d_legal, float* d_reward, uint8_t* d_step_type,
                    int32_t* d_action, uint64_t* d_hist, void* stream) {
  HB_REQUIRE(e && n_steps >= 1 && n_steps <= 255, "hb_env_selfplay: n_steps must be 1..255");
  if (!e->is_reset) { hb::set_error("hb_env_selfplay: reset the environment first"); return HB_ERR_STATE; }
  const RuleSet* rsp = nullptr;
  bool ext = false;
  int rc = cached_ruleset(e, rule_ids, rule_farg, rule_iarg, n_rules, seed, &rsp, &ext);
  if (rc) return rc;
  const RuleSet& rs = *rsp;
  hb::DeviceGuard g(e->device);
  const int n_groups = (int)grid_for(e);
  if (e->spq == nullptr) {
    HB_CUDA(cudaMalloc(&e->spq, 8 * (1 + kSpqLaunches) + (size_t)n_groups * 8));
    e->spq_groups = n_groups;
    e->spq_seq = 0;
    SpQueue q0{e->spq, e->spq + 1, (int*)(e->spq + 1 + kSpqLaunches),
               (int*)(e->spq + 1 + kSpqLaunches) + n_groups, n_groups, 0ull};
    spq_init_kernel<<<(n_groups + 255) / 256, 256, 0, (cudaStream_t)stream>>>(q0);
    HB_CUDA(cudaGetLastError());
  }
  SpQueue q{e->spq, e->spq + 1, (int*)(e->spq + 1 + kSpqLaunches),
            (int*)(e->spq + 1 + kSpqLaunches) + n_groups, n_groups, e->spq_seq};
  // consecutive self-play launches may overlap (see SpQueue); any other kernel is ordered as usual
  const bool chained = e->spq_chain && e->spq_chain_stream == stream && e->spq_chain_hist == (const void*)d_hist;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = chained ? 1 : 0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)n_groups);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = (cudaStream_t)stream;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  unsigned long long* hist = (unsigned long long*)d_hist;
  const uint8_t* decks = e->decks;
  const unsigned long long s0 = step0;
  cudaError_t lerr = cudaSuccess;
  if (ext) {
    HB_DISPATCH(e, (lerr = cudaLaunchKernelEx(&cfg, selfplay_kernel<SH, true>, e->rt, e->state, decks, rs, s0,
                                              n_steps, d_obs, d_legal, d_reward, d_step_type, d_action, hist,
                                              e->illegal, q)));
  } else {
    HB_DISPATCH(e, (lerr = cudaLaunchKernelEx(&cfg, selfplay_kernel<SH, false>, e->rt, e->state, decks, rs, s0,
                                              n_steps, d_obs, d_legal, d_reward, d_step_type, d_action, hist,
                                              e->illegal, q)));
  }
  HB_CUDA(lerr);
  ++e->spq_seq;
  e->spq_chain = true;
  e->spq_chain_stream = stream;
  e->spq_chain_hist = (const void*)d_hist;
  HB_CUDA(cudaGetLastError());
  return HB_OK;
}

// Host-buffer step: what a binding with numpy arguments calls (copies inside).
int hb_env_step_host(hb_env* e, const int32_t* h_action, uint32_t* h_obs, uint32_t* h_legal,
                     float* h_reward, uint8_t* h_step_type) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && h_action, "hb_env_step_host: bad arguments");
  hb::DeviceGuard g(e->device);
  const size_t n = (size_t)e->rt.n_games;
  const size_t b_act = n * 4, b_obs = n * e->obs_stride * 4, b_leg = n * e->legal_words * 4,
               b_rew = n * 4, b_st = (n + 3) / 4 * 4;
  int rc = env_stage(e, b_act + b_obs + b_leg + b_rew + b_st);
  if (rc) return rc;
  char* p = (char*)e->stage;
  int32_t* d_act = (int32_t*)p; p += b_act;
  uint32_t* d_obs = (uint32_t*)p; p += b_obs;
  uint32_t* d_leg = (uint32_t*)p; p += b_leg;
  float* d_rew = (float*)p; p += b_rew;
  uint8_t* d_st = (uint8_t*)p;
  HB_CUDA(cudaMemcpyAsync(d_act, h_action, b_act, cudaMemcpyHostToDevice, nullptr));
  rc = hb_env_step(e, d_act, h_obs ? d_obs : nullptr, h_legal ? d_leg : nullptr,
                   h_reward ? d_rew : nullptr, h_step_type ? d_st : nullptr, nullptr);
  if (rc) return rc;
  if (h_obs) HB_CUDA(cudaMemcpyAsync(h_obs, d_obs, b_obs, cudaMemcpyDeviceToHost, nullptr));
  if (h_legal) HB_CUDA(cudaMemcpyAsync(h_legal, d_leg, b_leg, cudaMemcpyDeviceToHost, nullptr));
  if (h_reward) HB_CUDA(cudaMemcpyAsync(h_reward, d_rew, b_rew, cudaMemcpyDeviceToHost, nullptr));
  if (h_step_type) HB_CUDA(cudaMemcpyAsync(h_step_type, d_st, n, cudaMemcpyDeviceToHost, nullptr));
  HB_CUDA(cudaStreamSynchronize(nullptr));
  return HB_OK;
}


// ---- chunked host pipeline ---------------------------------------------------------------------
// The reference's loop is `observations = env.step(actions); actions = agent.explore(observations)`
// with numpy arrays (run_experiment.py:293-308).  With HOST buffers that loop is bound by PCIe: 93 bytes
// of outputs per game and step.  The three calls below keep the link busy: the games are cut into
// chunks; chunk c's actions go up, its step kernel runs, and while its outputs travel down chunk c+1 is
// already stepping and the agent is choosing chunk c's next move on a third stream.  Observation rows
// are written without the 16-byte alignment padding (obs_words = ceil(obs_len / 32) words per game).
// All host pointers should be pinned (cudaHostAlloc / torch pin_memory) -- pageable memory works but the
// copies then serialise.

namespace {
struct StagePlan {
  int32_t* act; uint32_t* obs; uint32_t* leg; float* rew; uint8_t* st; int32_t* act_out;
};
int host_pipe_init(hb_env* e, int obs_words, StagePlan* p, int slot = 0) {
  if (!e->hs_step) {
    HB_CUDA(cudaStreamCreateWithFlags(&e->hs_step, cudaStreamNonBlocking));
    HB_CUDA(cudaStreamCreateWithFlags(&e->hs_out, cudaStreamNonBlocking));
    HB_CUDA(cudaStreamCreateWithFlags(&e->hs_act, cudaStreamNonBlocking));
    for (int c = 0; c < hb_env::kMaxChunks; ++c)
      HB_CUDA(cudaEventCreateWithFlags(&e->hev_step[c], cudaEventDisableTiming));
    HB_CUDA(cudaEventCreateWithFlags(&e->hev_done, cudaEventDisableTiming));
    HB_CUDA(cudaEventCreateWithFlags(&e->hev_out[0], cudaEventDisableTiming));
    HB_CUDA(cudaEventCreateWithFlags(&e->hev_out[1], cudaEventDisableTiming));
    HB_CUDA(cudaEventCreateWithFlags(&e->hev_rules, cudaEventDisableTiming));
    HB_CUDA(cudaEventCreateWithFlags(&e->hev_acts, cudaEventDisableTiming));
  }
  const size_t n = (size_t)e->rt.n_games;
  auto up = [](size_t b) { return (b + 255) / 256 * 256; };
  // the four outputs sit back to back (all but the last are whole words), so that host arrays laid
  // out the same way -- HanabiParallelEnv.host_buffers -- are filled by ONE copy per chunk
  const size_t b_act = up(n * 4), b_obs = n * obs_words * 4, b_leg = n * e->legal_words * 4,
               b_rew = n * 4, b_st = up(n);
  const size_t b_all = 2 * b_act + b_obs + b_leg + b_rew + b_st;       // two staging sets
  int rc = env_stage(e, 2 * b_all);
  if (rc) return rc;
  char* q = (char*)e->stage + (size_t)(slot & 1) * b_all;
  p->act = (int32_t*)q; q += b_act;
  p->obs = (uint32_t*)q; q += b_obs;
  p->leg = (uint32_t*)q; q += b_leg;
  p->rew = (float*)q; q += b_rew;
  p->st = (uint8_t*)q; q += b_st;
  p->act_out = (int32_t*)q;
  return HB_OK;
}
inline long long chunk_games(const hb_env* e, int chunks) {
  const long long per = (e->rt.n_games + chunks - 1) / chunks;
  return (per + kThreads - 1) / kThreads * kThreads;
}
}  // namespace

int hb_env_step_host_async(hb_env* e, const int32_t* h_action, uint32_t* h_obs, uint32_t* h_legal,
                           float* h_reward, uint8_t* h_step_type, int chunks) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && h_action && chunks >= 1 && chunks <= hb_env::kMaxChunks,
             "hb_env_step_host_async: bad arguments");
  if (!e->is_reset) { hb::set_error("hb_env_step_host_async: reset the environment first"); return HB_ERR_STATE; }
  hb::DeviceGuard g(e->device);
  const int obs_words = (e->obs_len + 31) / 32;
  StagePlan p;
  int rc = host_pipe_init(e, obs_words, &p, e->h_slot);
  if (rc) return rc;
  // what the previous pipeline pass still reads or writes must be done before the staging is reused
  HB_CUDA(cudaStreamWaitEvent(e->hs_step, e->hev_done, 0));
  HB_CUDA(cudaStreamWaitEvent(e->hs_step, e->hev_out[e->h_slot], 0));
  const long long n = e->rt.n_games, per = chunk_games(e, chunks);
  int used = 0;
  for (long long first = 0; first < n; first += per, ++used) {
    const long long cnt = std::min(per, n - first);
    const unsigned grid = (unsigned)((cnt + kThreads - 1) / kThreads);
    HB_CUDA(cudaMemcpyAsync(p.act + first, h_action + first, (size_t)cnt * 4, cudaMemcpyHostToDevice, e->hs_step));
    HB_DISPATCH(e, (env_step_kernel<SH><<<grid, kThreads, 0, e->hs_step>>>(
                       e->rt, e->state, e->decks, p.act, h_obs ? p.obs : nullptr, h_legal ? p.leg : nullptr,
                       h_reward ? p.rew : nullptr, h_step_type ? p.st : nullptr, e->illegal, first, cnt,
                       obs_words)));
    HB_CUDA(cudaGetLastError());
    HB_CUDA(cudaEventRecord(e->hev_step[used], e->hs_step));
    HB_CUDA(cudaStreamWaitEvent(e->hs_out, e->hev_step[used], 0));
    const bool whole = cnt == n && h_obs && h_legal && h_reward && h_step_type &&
                       (char*)h_legal == (char*)h_obs + (size_t)n * obs_words * 4 &&
                       (char*)h_reward == (char*)h_legal + (size_t)n * e->legal_words * 4 &&
                       (char*)h_step_type == (char*)h_reward + (size_t)n * 4;
    if (whole) {     // one chunk into back-to-back host arrays: a single copy (each copy costs ~5 us)
      HB_CUDA(cudaMemcpyAsync(h_obs, p.obs, (size_t)n * (obs_words * 4 + e->legal_words * 4 + 4 + 1),
                              cudaMemcpyDeviceToHost, e->hs_out));
      continue;
    }
    if (h_obs)
      HB_CUDA(cudaMemcpyAsync(h_obs + first * obs_words, p.obs + first * obs_words, (size_t)cnt * obs_words * 4,
                              cudaMemcpyDeviceToHost, e->hs_out));
    if (h_legal)
      HB_CUDA(cudaMemcpyAsync(h_legal + first * e->legal_words, p.leg + first * e->legal_words,
                              (size_t)cnt * e->legal_words * 4, cudaMemcpyDeviceToHost, e->hs_out));
    if (h_reward)
      HB_CUDA(cudaMemcpyAsync(h_reward + first, p.rew + first, (size_t)cnt * 4, cudaMemcpyDeviceToHost, e->hs_out));
    if (h_step_type)
      HB_CUDA(cudaMemcpyAsync(h_step_type + first, p.st + first, (size_t)cnt, cudaMemcpyDeviceToHost, e->hs_out));
  }
  e->h_chunks = used;
  return HB_OK;
}

int hb_rules_act_host_async(hb_env* e, const int32_t* rule_ids, const double* rule_farg,
                            const int32_t* rule_iarg, int n_rules, uint64_t seed, uint64_t step,
                            int32_t* h_action, uint64_t* d_hist) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && h_action, "hb_rules_act_host_async: bad arguments");
  HB_REQUIRE(e->h_chunks >= 1, "hb_rules_act_host_async: call hb_env_step_host_async first");
  const RuleSet* rsp = nullptr;
  bool ext = false;
  int rc = cached_ruleset(e, rule_ids, rule_farg, rule_iarg, n_rules, seed, &rsp, &ext);
  if (rc) return rc;
  const RuleSet& rs = *rsp;
  hb::DeviceGuard g(e->device);
  StagePlan p;
  rc = host_pipe_init(e, (e->obs_len + 31) / 32, &p, e->h_slot);
  if (rc) return rc;
  const long long n = e->rt.n_games, per = chunk_games(e, e->h_chunks);
  int c = 0;
  for (long long first = 0; first < n; first += per, ++c) {
    const long long cnt = std::min(per, n - first);
    const unsigned grid = (unsigned)((cnt + kThreads - 1) / kThreads);
    HB_CUDA(cudaStreamWaitEvent(e->hs_act, e->hev_step[c], 0));
    if (ext) {
      HB_DISPATCH(e, (rules_act_kernel<SH, true><<<grid, kThreads, 0, e->hs_act>>>(
                         e->rt, e->state, rs, step, p.act_out, nullptr, (unsigned long long*)d_hist, first, cnt)));
    } else {
      HB_DISPATCH(e, (rules_act_kernel<SH, false><<<grid, kThreads, 0, e->hs_act>>>(
                         e->rt, e->state, rs, step, p.act_out, nullptr, (unsigned long long*)d_hist, first, cnt)));
    }
    HB_CUDA(cudaGetLastError());
    HB_CUDA(cudaEventRecord(e->hev_rules, e->hs_act));          // the state has been read
    HB_CUDA(cudaMemcpyAsync(h_action + first, p.act_out + first, (size_t)cnt * 4, cudaMemcpyDeviceToHost, e->hs_act));
  }
  HB_CUDA(cudaEventRecord(e->hev_acts, e->hs_act));
  return HB_OK;
}

// ---- double-buffered host loop ------------------------------------------------------------------
// One step of the host-in-the-loop pattern is bound by its 6 MB of outputs crossing PCIe (114 us) plus
// ~40 us of dependent small work (actions up, step kernel, agent kernel, next actions down).  With two
// staging sets the next step's small work runs while this step's outputs are still on the wire:
//   hb_env_step_host_begin(actions)    actions H2D + env_step_kernel into the other staging set
//   hb_rules_act_host_async(next)      agent kernel + the next actions D2H (small, queued FIRST)
//   hb_env_step_host_outputs(bufs)     the step's outputs D2H (large), into host buffers the caller alternates
//   hb_env_host_wait_actions()         returns when `next` is on the host; the outputs may still be in flight
// hb_env_host_wait_outputs(slot) / hb_env_host_wait() tell when host buffers are complete.
int hb_env_step_host_begin(hb_env* e, const int32_t* h_action) {
  if (e) e->spq_chain = false;
  HB_REQUIRE(e && h_action, "hb_env_step_host_begin: bad arguments");
  if (!e->is_reset) { hb::set_error("hb_env_step_host_begin: reset the environment first"); return HB_ERR_STATE; }
  hb::DeviceGuard g(e->device);
  const int obs_words = (e->obs_len + 31) / 32;
  e->h_slot ^= 1;
  StagePlan p;
  int rc = host_pipe_init(e, obs_words, &p, e->h_slot);
  if (rc) return rc;
  // this staging set was last used two steps ago: its outputs must have left, and the agent of the previous
  // step must have read the state this step changes
  HB_CUDA(cudaStreamWaitEvent(e->hs_step, e->hev_out[e->h_slot], 0));
  HB_CUDA(cudaStreamWaitEvent(e->hs_step, e->hev_rules, 0));
  HB_CUDA(cudaStreamWaitEvent(e->hs_step, e->hev_done, 0));
  const long long n = e->rt.n_games;
  HB_CUDA(cudaMemcpyAsync(p.act, h_action, (size_t)n * 4, cudaMemcpyHostToDevice, e->hs_step));
  HB_DISPATCH(e, (env_step_kernel<SH><<<grid_for(e), kThreads, 0, e->hs_step>>>(
                     e->rt, e->state, e->decks, p.act, p.obs, p.leg, p.rew, p.st, e->illegal, 0, n, obs_words)));
  HB_CUDA(cudaGetLastError());
  HB_CUDA(cudaEventRecord(e->hev_step[0], e->hs_step));
  e->h_chunks = 1;
  return HB_OK;
}

int hb_env_step_host_outputs(hb_env* e, uint32_t* h_obs, uint32_t* h_legal, float* h_reward,
                             uint8_t* h_step_type) {
  HB_REQUIRE(e && e->h_chunks == 1, "hb_env_step_host_outputs: call hb_env_step_host_begin first");
  hb::DeviceGuard g(e->device);
  const int obs_words = (e->obs_len + 31) / 32;
  StagePlan p;
  int rc = host_pipe_init(e, obs_words, &p, e->h_slot);
  if (rc) return rc;
  const size_t n = (size_t)e->rt.n_games;
  HB_CUDA(cudaStreamWaitEvent(e->hs_out, e->hev_step[0], 0));
  // The large copy is ready before the agent's small one (it needs only the step kernel) and would take
  // the copy engine first: the host would then get its next actions only after 6 MB had crossed.  Behind
  // the actions it costs the loop nothing, and the next step's work overlaps it.
  HB_CUDA(cudaStreamWaitEvent(e->hs_out, e->hev_acts, 0));
  const bool whole = h_obs && h_legal && h_reward && h_step_type &&
                     (char*)h_legal == (char*)h_obs + n * obs_words * 4 &&
                     (char*)h_reward == (char*)h_legal + n * e->legal_words * 4 &&
                     (char*)h_step_type == (char*)h_reward + n * 4;
  if (whole) {
    HB_CUDA(cudaMemcpyAsync(h_obs, p.obs, n * (obs_words * 4 + e->legal_words * 4 + 4 + 1),
                            cudaMemcpyDeviceToHost, e->hs_out));
  } else {
    if (h_obs) HB_CUDA(cudaMemcpyAsync(h_obs, p.obs, n * obs_words * 4, cudaMemcpyDeviceToHost, e->hs_out));
    if (h_legal) HB_CUDA(cudaMemcpyAsync(h_legal, p.leg, n * e->legal_words * 4, cudaMemcpyDeviceToHost, e->hs_out));
    if (h_reward) HB_CUDA(cudaMemcpyAsync(h_reward, p.rew, n * 4, cudaMemcpyDeviceToHost, e->hs_out));
    if (h_step_type) HB_CUDA(cudaMemcpyAsync(h_step_type, p.st, n, cudaMemcpyDeviceToHost, e->hs_out));
  }
  HB_CUDA(cudaEventRecord(e->hev_out[e->h_slot], e->hs_out));
  return HB_OK;
}

int hb_env_host_wait_actions(hb_env* e) {
  HB_REQUIRE(e, "hb_env_host_wait_actions: NULL env");
  if (!e->hs_act) return HB_OK;
  hb::DeviceGuard g(e->device);
  HB_CUDA(cudaStreamSynchronize(e->hs_act));
  return HB_OK;
}

int hb_env_host_wait_outputs(hb_env* e, int slot) {
  HB_REQUIRE(e && (slot == 0 || slot == 1), "hb_env_host_wait_outputs: bad arguments");
  if (!e->hs_out) return HB_OK;
  hb::DeviceGuard g(e->device);
  HB_CUDA(cudaEventSynchronize(e->hev_out[slot]));
  return HB_OK;
}

int hb_env_host_slot(const hb_env* e) { return e ? e->h_slot : -1; }

int hb_env_host_wait(hb_env* e) {
  HB_REQUIRE(e, "hb_env_host_wait: NULL env");
  if (!e->hs_step) return HB_OK;
  hb::DeviceGuard g(e->device);
  HB_CUDA(cudaStreamSynchronize(e->hs_step));
  HB_CUDA(cudaStreamSynchronize(e->hs_out));
  HB_CUDA(cudaStreamSynchronize(e->hs_act));
  HB_CUDA(cudaEventRecord(e->hev_done, e->hs_act));
  return HB_OK;
}

}  // extern "C"
