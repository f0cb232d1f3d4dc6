// Warp-per-game Hanabi engine: state layout and device functions.
//
// Replaces (for N games at once) the third-party HLE calls the reference makes through
// `environment.step` / `observation['vectorized']` / `legal_moves_as_int`
// (hanabi_agents/rainbow/run_experiment.py:266-273,308; consumed as observations[1] at
// hanabi_agents/rlax_dqn/rlax_rainbow.py:278,285,298).  Semantics: SURVEY.md Appendix A.
//
// One warp owns one game.  Lane s < S (S = players*hand_size) owns CARD SLOT s =
// player*H + index: its card and the hint knowledge about it live in one register.
// Everything else (fireworks, tokens, discard counts, last move ...) is warp-uniform and
// kept redundantly in every lane.  A game is `kStateWords` 32-bit words in HBM, lane j
// loads/stores word j, so a game's state moves as one coalesced 64 B / 128 B access.
//
// Slot word:   [4:0] card (color*R+rank)  [9:5] colours still plausible  [14:10] ranks still
//              plausible  [15] colour hinted  [16] rank hinted  [17] slot occupied
// Scalar words (at word S + k):
//   k=0  [14:0] fireworks 3b x5 | [18:15] info | [20:19] life | [23:21] cur player |
//        [26:24] turns_to_play | [28:27] step_type
//   k=1  [5:0] deck_pos | [20:6] hand_n 3b x5
//   k=2,3  discard counts, 2 bits per card type (25 types -> 50 bits)
//   k=4  last non-deal move: [0] valid [3:1] player [9:4] move uid [12:10] card colour
//        [15:13] card rank [16] scored [17] info token [22:18] reveal bitmask
//   k=5  episode counter (selects the deck: (game_id + episode*n_total) mod n_decks)
#pragma once
#include <stdint.h>

namespace hbg {

constexpr unsigned kFull = 0xffffffffu;
enum { FIRST = 0, MID = 1, LAST = 2 };
enum { PLAY = 1, DISCARD = 2, REVEAL_COLOR = 3, REVEAL_RANK = 4 };

template <int P_, int H_, int C_, int R_, int MI_, int ML_>
struct Shape {
  static_assert(R_ == 5, "rank-dependent bit tricks assume 5 ranks (3/2/2/2/1 instances)");
  static constexpr int P = P_, H = H_, C = C_, R = R_;
  static constexpr int MAX_INFO = MI_, MAX_LIFE = ML_;
  static constexpr int S = P * H;                 // card slots
  static constexpr int BPC = C * R;               // bits per card
  static constexpr int kPerColor = (R == 1) ? 3 : (3 + 2 * (R - 2) + 1);
  static constexpr int DECK = C * kPerColor;
  static constexpr int A = 2 * H + (P - 1) * (C + R);       // num moves
  static constexpr int kStateWords = (S + 6 <= 16) ? 16 : 32;
  static constexpr int kLegalWords = (A + 31) / 32;
  // canonical observation layout, all in bits
  static constexpr int kHandsCards = (P - 1) * H * BPC;      // other players' cards
  static constexpr int kOffBoard = kHandsCards + P;           // after the "short hand" bits
  static constexpr int kDeckThermo = DECK - P * H;
  static constexpr int kOffFireworks = kOffBoard + kDeckThermo;
  static constexpr int kOffInfo = kOffFireworks + BPC;
  static constexpr int kOffDisc = kOffInfo + MAX_INFO + MAX_LIFE;
  static constexpr int kOffLast = kOffDisc + DECK;
  static constexpr int kLastLen = 2 * P + 4 + C + R + 2 * H + BPC + 2;
  static constexpr int kOffKnow = kOffLast + kLastLen;
  static constexpr int kKnowWidth = BPC + C + R;
  static constexpr int kObsLen = kOffKnow + P * H * kKnowWidth;
  static constexpr int kObsWords = (kObsLen + 31) / 32;
  static constexpr int kObsStride = (kObsWords + 3) / 4 * 4;  // row = multiple of 16 B
  __host__ __device__ static constexpr int instances(int r) {
    return r == 0 ? 3 : (r == R - 1 ? 1 : 2);
  }
  // bit offset of rank r inside one colour's discard thermometer block
  __host__ __device__ static constexpr int disc_off(int r) {
    int o = 0;
    for (int i = 0; i < r; ++i) o += instances(i);
    return o;
  }
};

// runtime part of the configuration (kernel argument, by value)
struct Rt {
  long long n_games;                // games on this device
  long long n_total;                // games in the whole job (deck selection)
  long long game_id_base;           // global id of local game 0
  long long n_decks;
};

template <class SH>
struct Game {
  // per-lane
  uint32_t slot;       // this lane's card slot word (0 for lanes >= S)
  // warp-uniform
  uint32_t fw;         // 3 bits per colour
  int info, life, cur, ttp, step_type, deck_pos;
  uint32_t hn;         // 3 bits per player
  unsigned long long disc;  // 2 bits per card type
  uint32_t last;
  uint32_t episode;

  __device__ __forceinline__ int firework(int c) const { return (fw >> (3 * c)) & 7; }
  __device__ __forceinline__ int hand_n(int p) const { return (hn >> (3 * p)) & 7; }
  __device__ __forceinline__ int disc_n(int t) const { return (int)((disc >> (2 * t)) & 3ull); }
  __device__ __forceinline__ int score() const {
    if (life <= 0) return 0;
    int s = 0;
#pragma unroll
    for (int c = 0; c < SH::C; ++c) s += firework(c);
    return s;
  }
  __device__ __forceinline__ bool terminal() const {
    return life < 1 || score() >= SH::C * SH::R || ttp <= 0;
  }
};

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ int slot_card(uint32_t w) { return w & 31; }
__device__ __forceinline__ int slot_cplaus(uint32_t w) { return (w >> 5) & 31; }
__device__ __forceinline__ int slot_rplaus(uint32_t w) { return (w >> 10) & 31; }
__device__ __forceinline__ bool slot_chint(uint32_t w) { return (w >> 15) & 1; }
__device__ __forceinline__ bool slot_rhint(uint32_t w) { return (w >> 16) & 1; }
__device__ __forceinline__ bool slot_valid(uint32_t w) { return (w >> 17) & 1; }

template <class SH>
__device__ __forceinline__ uint32_t fresh_slot(int card) {
  return (uint32_t)card | (((1u << SH::C) - 1) << 5) | (((1u << SH::R) - 1) << 10) | (1u << 17);
}

// ---- state <-> HBM -----------------------------------------------------------------
template <class SH>
__device__ __forceinline__ void load_game(Game<SH>& g, const uint32_t* __restrict__ state,
                                          long long game) {
  const int lane = lane_id();
  uint32_t w = 0;
  if (lane < SH::kStateWords) w = state[game * SH::kStateWords + lane];
  g.slot = lane < SH::S ? w : 0u;
  const uint32_t a = __shfl_sync(kFull, w, SH::S + 0);
  const uint32_t b = __shfl_sync(kFull, w, SH::S + 1);
  const uint32_t d0 = __shfl_sync(kFull, w, SH::S + 2);
  const uint32_t d1 = __shfl_sync(kFull, w, SH::S + 3);
  g.last = __shfl_sync(kFull, w, SH::S + 4);
  g.episode = __shfl_sync(kFull, w, SH::S + 5);
  g.fw = a & 0x7fff;
  g.info = (a >> 15) & 15;
  g.life = (a >> 19) & 3;
  g.cur = (a >> 21) & 7;
  g.ttp = (a >> 24) & 7;
  g.step_type = (a >> 27) & 3;
  g.deck_pos = b & 63;
  g.hn = (b >> 6) & 0x7fff;
  g.disc = (unsigned long long)d0 | ((unsigned long long)d1 << 32);
}

template <class SH>
__device__ __forceinline__ void store_game(const Game<SH>& g, uint32_t* __restrict__ state,
                                           long long game) {
  const int lane = lane_id();
  uint32_t w = g.slot;
  const uint32_t a = g.fw | ((uint32_t)g.info << 15) | ((uint32_t)g.life << 19) |
                     ((uint32_t)g.cur << 21) | ((uint32_t)g.ttp << 24) |
                     ((uint32_t)g.step_type << 27);
  const uint32_t b = (uint32_t)g.deck_pos | (g.hn << 6);
  if (lane == SH::S + 0) w = a;
  if (lane == SH::S + 1) w = b;
  if (lane == SH::S + 2) w = (uint32_t)g.disc;
  if (lane == SH::S + 3) w = (uint32_t)(g.disc >> 32);
  if (lane == SH::S + 4) w = g.last;
  if (lane == SH::S + 5) w = g.episode;
  if (lane >= SH::S + 6) w = 0;
  if (lane < SH::kStateWords) state[game * SH::kStateWords + lane] = w;
}

template <class SH>
__device__ __forceinline__ const uint8_t* deck_of(const Rt& rt, const uint8_t* __restrict__ decks,
                                                  long long game, uint32_t episode) {
  const long long row = (rt.game_id_base + game + (long long)episode * rt.n_total) % rt.n_decks;
  return decks + row * SH::DECK;
}

// HanabiState ctor + initial deals (A.3): player 0 is filled first, then player 1, ...
template <class SH>
__device__ __forceinline__ void reset_game(Game<SH>& g, const Rt& rt, const uint8_t* deck) {
  const int lane = lane_id();
  g.slot = lane < SH::S ? fresh_slot<SH>(deck[lane]) : 0u;
  g.fw = 0;
  g.info = SH::MAX_INFO;
  g.life = SH::MAX_LIFE;
  g.cur = 0;
  g.ttp = SH::P;
  g.step_type = FIRST;
  g.deck_pos = SH::S;
  g.hn = 0;
#pragma unroll
  for (int p = 0; p < SH::P; ++p) g.hn |= (uint32_t)SH::H << (3 * p);
  g.disc = 0;
  g.last = 0;
}

// ---- legality (A.4) -------------------------------------------------------------------
// 64-bit mask over move uids for player q (the player to move).
template <class SH>
__device__ __forceinline__ unsigned long long legal_mask(const Game<SH>& g, const Rt& rt, int q) {
  const int lane = lane_id();
  const unsigned long long hand_bits = (1ull << g.hand_n(q)) - 1;
  unsigned long long m = hand_bits << SH::H;                 // Play(i), i < hand size
  if (g.info < SH::MAX_INFO) m |= hand_bits;                  // Discard(i)
  uint32_t lo = 0, hi = 0;
  if (g.info > 0 && lane < SH::S && slot_valid(g.slot)) {
    const int p = lane / SH::H;
    int k = p - q;
    if (k < 0) k += SH::P;
    if (k >= 1) {
      const int card = slot_card(g.slot);
      const int c = card / SH::R, r = card - c * SH::R;
      const unsigned long long b =
          (1ull << (2 * SH::H + (k - 1) * SH::C + c)) |
          (1ull << (2 * SH::H + (SH::P - 1) * SH::C + (k - 1) * SH::R + r));
      lo = (uint32_t)b;
      hi = (uint32_t)(b >> 32);
    }
  }
  lo = __reduce_or_sync(kFull, lo);
  if (SH::A > 32) hi = __reduce_or_sync(kFull, hi);
  return m | (unsigned long long)lo | ((unsigned long long)hi << 32);
}

// MoveIsLegal for ONE move uid of the player to move (cheaper than the full mask).
template <class SH>
__device__ __forceinline__ bool move_is_legal(const Game<SH>& g, int uid) {
  if (uid < 0 || uid >= SH::A) return false;
  const int me = g.cur;
  if (uid < SH::H) return g.info < SH::MAX_INFO && uid < g.hand_n(me);
  if (uid < 2 * SH::H) return uid - SH::H < g.hand_n(me);
  if (g.info <= 0) return false;
  int u = uid - 2 * SH::H;
  const bool by_color = u < (SH::P - 1) * SH::C;
  if (!by_color) u -= (SH::P - 1) * SH::C;
  const int span = by_color ? SH::C : SH::R;
  const int k = 1 + u / span, v = u - (k - 1) * span;
  int t = me + k;
  if (t >= SH::P) t -= SH::P;
  const int lane = lane_id();
  bool match = false;
  if (lane < SH::S && slot_valid(g.slot) && lane / SH::H == t) {
    const int card = slot_card(g.slot);
    const int c = card / SH::R, r = card - c * SH::R;
    match = by_color ? (c == v) : (r == v);
  }
  return __any_sync(kFull, match);
}

// ---- transition (A.3): ApplyMove + chance deal + AdvanceToNextPlayer -------------------
// `uid` must be legal.  Returns the reward (score delta).
template <class SH>
__device__ __forceinline__ int apply_move(Game<SH>& g, const Rt& rt, int uid, const uint8_t* deck) {
  const int lane = lane_id();
  const int before = g.score();
  if (g.deck_pos >= SH::DECK) --g.ttp;          // evaluated BEFORE the move
  const int me = g.cur;
  uint32_t last = 1u | ((uint32_t)me << 1) | ((uint32_t)uid << 4);
  if (uid < 2 * SH::H) {
    const bool is_play = uid >= SH::H;
    const int idx = is_play ? uid - SH::H : uid;
    const int card = slot_card(__shfl_sync(kFull, g.slot, me * SH::H + idx));
    const int c = card / SH::R, r = card - c * SH::R;
    last |= ((uint32_t)c << 10) | ((uint32_t)r << 13);
    bool to_discard = true;
    if (!is_play) {
      if (g.info < SH::MAX_INFO) { ++g.info; last |= 1u << 17; }
    } else if (r == g.firework(c)) {             // AddToFireworks
      g.fw += 1u << (3 * c);
      last |= 1u << 16;
      to_discard = false;
      if (g.firework(c) == SH::R && g.info < SH::MAX_INFO) { ++g.info; last |= 1u << 17; }
    } else {
      --g.life;
    }
    if (to_discard) g.disc += 1ull << (2 * card);
    // RemoveFromHand: later cards (and their knowledge) shift down
    const uint32_t up = __shfl_down_sync(kFull, g.slot, 1);
    const int p = lane / SH::H, i = lane - p * SH::H;
    if (lane < SH::S && p == me && i >= idx) g.slot = (i == SH::H - 1) ? 0u : up;
    int n = g.hand_n(me) - 1;
    // chance node: the only short hand is `me`; the new card goes to the END of the hand
    if (g.deck_pos < SH::DECK) {
      const int nc = deck[g.deck_pos++];
      if (lane == me * SH::H + n) g.slot = fresh_slot<SH>(nc);
      ++n;
    }
    g.hn = (g.hn & ~(7u << (3 * me))) | ((uint32_t)n << (3 * me));
  } else {
    --g.info;
    int u = uid - 2 * SH::H;
    const bool by_color = u < (SH::P - 1) * SH::C;
    if (!by_color) u -= (SH::P - 1) * SH::C;
    const int span = by_color ? SH::C : SH::R;
    const int k = 1 + u / span, v = u - (k - 1) * span;
    int t = me + k;
    if (t >= SH::P) t -= SH::P;
    bool match = false;
    if (lane < SH::S && slot_valid(g.slot) && lane / SH::H == t) {
      const int card = slot_card(g.slot);
      const int c = card / SH::R, r = card - c * SH::R;
      if (by_color) {                            // HanabiHand::RevealColor
        match = (c == v);
        if (match) g.slot = (g.slot & ~(31u << 5)) | (1u << (5 + v)) | (1u << 15);
        else g.slot &= ~(1u << (5 + v));
      } else {                                   // HanabiHand::RevealRank
        match = (r == v);
        if (match) g.slot = (g.slot & ~(31u << 10)) | (1u << (10 + v)) | (1u << 16);
        else g.slot &= ~(1u << (10 + v));
      }
    }
    const uint32_t mask = (__ballot_sync(kFull, match) >> (t * SH::H)) & ((1u << SH::H) - 1);
    last |= mask << 18;
  }
  g.last = last;
  g.cur = me + 1 == SH::P ? 0 : me + 1;
  return g.score() - before;
}

}  // namespace hbg
