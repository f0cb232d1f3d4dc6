// Thread-per-game Hanabi engine: state layout, legality and transition.
//
// Replaces (for N games at once) the third-party HLE calls the reference makes through
// `environment.step` / `observation['vectorized']` / `legal_moves_as_int`
// (hanabi_agents/rainbow/run_experiment.py:266-273,308; consumed as observations[1] at
// hanabi_agents/rlax_dqn/rlax_rainbow.py:278,285,298).  Semantics: SURVEY.md Appendix A.
//
// One THREAD owns one game; the whole game lives in its registers.  Every branch of the
// rules (4 move types, N agent rules) is small, so a warp simply walks all of them with the
// non-participating games predicated off: no instruction is spent on warp-uniform bookkeeping
// (the first, warp-per-game version of this file spent >70 % of its issue slots there;
// profiles/r01_selfplay_warp_per_game.md).
//
// Hands are kept in SEAT-RELATIVE order: hand[0] is always the player to move, hand[k] the
// player k seats later.  After every move the hands rotate by one seat, so all agent rules
// and the whole observation encoder index with compile-time constants.
//
// A game is kStateWords 32-bit words in HBM (16 for 2 players = 64 B, else 32 = 128 B):
//   words 0..S-1  card slots, seat-relative, slot (k, i) at word k*H + i:
//        [24:0] plausible(colour) x plausible(rank) grid, bit colour*5+rank (0 = empty slot)
//        [25] colour hinted  [26] rank hinted  [31:27] card index colour*5+rank
//   word S+0  [24:0] fireworks as a thermometer grid (bit colour*5+rank set iff rank < stack height)
//             | [28:25] info | [30:29] life
//   word S+1  [5:0] deck_pos | [20:6] hand sizes 3b x5 (seat-relative) | [23:21] player to move
//             | [26:24] turns_to_play | [28:27] step_type
//   word S+2,3  discard counts, 2 bits per card type (25 types -> 50 bits)
//   word S+4  last move (made by the player now at seat P-1): [0] valid [6:1] move uid
//             [11:7] card index [12] scored [13] info token [18:14] reveal bitmask
//   word S+5  episode counter (selects the deck: (game_id + episode*n_total) mod n_decks)
#pragma once
#include <stdint.h>

namespace hbg {

enum { FIRST = 0, MID = 1, LAST = 2 };
enum { PLAY = 1, DISCARD = 2, REVEAL_COLOR = 3, REVEAL_RANK = 4 };

template <int P_, int H_, int C_, int R_, int MI_, int ML_>
struct Shape {
  static_assert(R_ == 5, "bit tricks assume 5 ranks with 3/2/2/2/1 instances");
  static constexpr int P = P_, H = H_, C = C_, R = R_;
  static constexpr int MAX_INFO = MI_, MAX_LIFE = ML_;
  static constexpr int S = P * H;                 // card slots
  static constexpr int BPC = C * R;               // bits per card
  static constexpr int kPerColor = 10;            // 3+2+2+2+1 cards per colour
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
  static constexpr uint32_t kGridAll = (1u << (C * R)) - 1;
};

// runtime part of the configuration (kernel argument, by value)
struct Rt {
  long long n_games;                // games on this device
  long long n_total;                // games in the whole job (deck selection)
  long long game_id_base;           // global id of local game 0
  long long n_decks;
  long long row_step;               // n_total mod n_decks: deck-row increment from one episode to the next
};

constexpr uint32_t kGridMask = 0x1ffffffu;
__device__ __forceinline__ uint32_t slot_grid(uint32_t w) { return w & kGridMask; }
__device__ __forceinline__ bool slot_valid(uint32_t w) { return (w & kGridMask) != 0; }
__device__ __forceinline__ uint32_t slot_chint(uint32_t w) { return (w >> 25) & 1u; }
__device__ __forceinline__ uint32_t slot_rhint(uint32_t w) { return (w >> 26) & 1u; }
__device__ __forceinline__ int slot_card(uint32_t w) { return (int)(w >> 27); }
__device__ __forceinline__ int card_color(int card) { return (card * 52) >> 8; }   // card / 5, card < 32
__device__ __forceinline__ int card_rank(int card) { return card - 5 * card_color(card); }
// all cells of colour c / of rank r in the 5x5 grid
__device__ __forceinline__ uint32_t color_cells(int c) { return 31u << (5 * c); }
__device__ __forceinline__ uint32_t rank_cells(int r) { return 0x108421u << r; }

// one-hot grid bit of the card in a slot (0 for an empty slot: the true identity of a card is
// always among its plausible ones, and an empty slot has an empty grid)
__device__ __forceinline__ uint32_t slot_onehot(uint32_t w) { return (1u << (w >> 27)) & w & kGridMask; }

template <class SH>
__device__ __forceinline__ uint32_t fresh_slot(int card) {
  return SH::kGridAll | ((uint32_t)card << 27);   // everything plausible, nothing hinted
}

// ---- 25-bit grids (bit colour*5 + rank) -------------------------------------------------------
// Everything the rules and the transition ask about a card ("is it playable", "has it been played",
// "is its colour finished", "is its rank below every stack") is ONE AND of the card's one-hot grid
// bit with a grid derived from the fireworks / discards once per step.
constexpr uint32_t kLsb5 = 0x0108421u;       // bit 0 of every colour group
constexpr uint32_t kLow4 = 0x0F7BDEFu;       // bits 0..3 of every colour group
constexpr uint32_t kRank4 = 0x1084210u;      // bit 4 (the last rank) of every colour group
// bit 0 / bit 1 of the number of copies per card type (3,2,2,2,1): ranks {0,4} / ranks {0,1,2,3}
constexpr uint32_t kInst0 = kLsb5 | kRank4;
constexpr uint32_t kInst1 = kLow4;

template <class SH>
struct Game {
  uint32_t hand[SH::P][SH::H];   // seat-relative card slots
  uint32_t oh[SH::P][SH::H];     // one-hot grid bit of each slot's card, 0 for an empty slot (derived, not
                                 // stored: 3 ALU instructions per slot and step when computed on demand)
  uint32_t fw;                   // fireworks thermometer grid: bit 5c+r <=> r < height of stack c
  uint32_t hn;                   // 3 bits per seat
  int info, life, cur, ttp, step_type, deck_pos;
  unsigned long long disc;       // 2 bits per card type (count of discarded copies)
  uint32_t d0, d1;               // the same counts as two bit planes in grid layout (derived, not stored)
  uint32_t last;
  uint32_t episode;

  __device__ __forceinline__ int firework(int c) const { return __popc((fw >> (5 * c)) & 31u); }
  __device__ __forceinline__ int hand_n(int k) const { return (hn >> (3 * k)) & 7; }
  __device__ __forceinline__ int score() const { return life <= 0 ? 0 : __popc(fw); }
  // HanabiState::EndOfGameStatus
  __device__ __forceinline__ bool terminal() const {
    return life < 1 || __popc(fw) >= SH::C * SH::R || ttp <= 0;
  }
  // grid of the cards playable right now: bit 5c + height(c)   (playable_card, ruleset.py:14-33).
  // (fw << 1) | lsb is the thermometer one notch higher; what leaks across a group boundary lands on
  // a bit the lsb mask sets anyway.
  __device__ __forceinline__ uint32_t playable() const {
    return ((fw << 1) | kLsb5) & ~fw & SH::kGridAll;
  }
};

// every second bit of x (bits 0, 2, 4, ...) packed into the low 32 bits
__device__ __forceinline__ uint32_t even_bits(unsigned long long x) {
  x &= 0x5555555555555555ull;
  x = (x | (x >> 1)) & 0x3333333333333333ull;
  x = (x | (x >> 2)) & 0x0f0f0f0f0f0f0f0full;
  x = (x | (x >> 4)) & 0x00ff00ff00ff00ffull;
  x = (x | (x >> 8)) & 0x0000ffff0000ffffull;
  x = (x | (x >> 16)) & 0x00000000ffffffffull;
  return (uint32_t)x;
}

// ---- state <-> HBM (16-byte vector accesses; a game is 64 B or 128 B contiguous) ----------
template <class SH, bool L2_ONLY = false>
__device__ __forceinline__ void load_game(Game<SH>& g, const uint32_t* __restrict__ state,
                                          long long game) {
  uint32_t w[SH::kStateWords];
  const uint4* src = reinterpret_cast<const uint4*>(state + game * SH::kStateWords);
#pragma unroll
  for (int i = 0; i < SH::kStateWords / 4; ++i) {
    if (4 * i < SH::S + 6) {
      const uint4 v = L2_ONLY ? __ldcg(src + i) : src[i];
      w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
    }
  }
#pragma unroll
  for (int k = 0; k < SH::P; ++k)
#pragma unroll
    for (int i = 0; i < SH::H; ++i) {
      g.hand[k][i] = w[k * SH::H + i];
      g.oh[k][i] = slot_onehot(g.hand[k][i]);
    }
  const uint32_t a = w[SH::S], b = w[SH::S + 1];
  g.fw = a & kGridMask;
  g.info = (a >> 25) & 15;
  g.life = (a >> 29) & 3;
  g.deck_pos = b & 63;
  g.hn = (b >> 6) & 0x7fff;
  g.cur = (b >> 21) & 7;
  g.ttp = (b >> 24) & 7;
  g.step_type = (b >> 27) & 3;
  g.disc = (unsigned long long)w[SH::S + 2] | ((unsigned long long)w[SH::S + 3] << 32);
  g.d0 = even_bits(g.disc);
  g.d1 = even_bits(g.disc >> 1);
  g.last = w[SH::S + 4];
  g.episode = w[SH::S + 5];
}

template <class SH>
__device__ __forceinline__ void store_game(const Game<SH>& g, uint32_t* __restrict__ state,
                                           long long game) {
  uint32_t w[SH::kStateWords];
#pragma unroll
  for (int i = 0; i < SH::kStateWords; ++i) w[i] = 0;
#pragma unroll
  for (int k = 0; k < SH::P; ++k)
#pragma unroll
    for (int i = 0; i < SH::H; ++i) w[k * SH::H + i] = g.hand[k][i];
  w[SH::S] = g.fw | ((uint32_t)g.info << 25) | ((uint32_t)g.life << 29);
  w[SH::S + 1] = (uint32_t)g.deck_pos | (g.hn << 6) | ((uint32_t)g.cur << 21) |
                 ((uint32_t)g.ttp << 24) | ((uint32_t)g.step_type << 27);
  w[SH::S + 2] = (uint32_t)g.disc;
  w[SH::S + 3] = (uint32_t)(g.disc >> 32);
  w[SH::S + 4] = g.last;
  w[SH::S + 5] = g.episode;
  uint4* dst = reinterpret_cast<uint4*>(state + game * SH::kStateWords);
#pragma unroll
  for (int i = 0; i < SH::kStateWords / 4; ++i)
    if (4 * i < SH::S + 6) dst[i] = make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
}

// The deck pool on the device: one deck per 64-byte row (hb_env_set_decks pads the rows), so that the
// cards of the initial deal are one aligned 16-byte load.
constexpr int kDeckPitch = 64;

template <class SH>
__device__ __forceinline__ unsigned long long deck_row(const Rt& rt, long long game, uint32_t episode) {
  return (unsigned long long)(rt.game_id_base + game + (long long)episode * rt.n_total) %
         (unsigned long long)rt.n_decks;
}

template <class SH>
__device__ __forceinline__ const uint8_t* deck_of(const Rt& rt, const uint8_t* __restrict__ decks,
                                                  long long game, uint32_t episode) {
  return decks + deck_row<SH>(rt, game, episode) * kDeckPitch;
}

// HanabiState ctor + initial deals (A.3): player 0 is filled first, then player 1, ...
// Player 0 moves first, so seat-relative order == absolute order here.
template <class SH>
__device__ __forceinline__ void reset_game(Game<SH>& g, const uint8_t* __restrict__ deck) {
  static_assert(SH::S <= 32, "initial deal read as two 16-byte loads");
  uint32_t q[8];
  {
    const uint4 v = *reinterpret_cast<const uint4*>(deck);
    q[0] = v.x; q[1] = v.y; q[2] = v.z; q[3] = v.w;
  }
  if constexpr (SH::S > 16) {
    const uint4 v = *reinterpret_cast<const uint4*>(deck + 16);
    q[4] = v.x; q[5] = v.y; q[6] = v.z; q[7] = v.w;
  }
#pragma unroll
  for (int k = 0; k < SH::P; ++k)
#pragma unroll
    for (int i = 0; i < SH::H; ++i) {
      const int j = k * SH::H + i;
      const uint32_t card = (q[j >> 2] >> (8 * (j & 3))) & 0xffu;
      g.hand[k][i] = fresh_slot<SH>((int)card);
      g.oh[k][i] = 1u << card;
    }
  g.fw = 0;
  g.info = SH::MAX_INFO;
  g.life = SH::MAX_LIFE;
  g.cur = 0;
  g.ttp = SH::P;
  g.step_type = FIRST;
  g.deck_pos = SH::S;
  g.hn = 0;
#pragma unroll
  for (int k = 0; k < SH::P; ++k) g.hn |= (uint32_t)SH::H << (3 * k);
  g.disc = 0;
  g.d0 = 0;
  g.d1 = 0;
  g.last = 0;
}

// union of the one-hots of hand k
template <class SH>
__device__ __forceinline__ uint32_t hand_union(const Game<SH>& g, int k) {
  uint32_t u = 0;
#pragma unroll
  for (int i = 0; i < SH::H; ++i) u |= g.oh[k][i];
  return u;
}

// which colours / ranks occur in a grid, as C-bit / R-bit masks
template <class SH>
__device__ __forceinline__ uint32_t grid_ranks(uint32_t u) {
  uint32_t r = u;
#pragma unroll
  for (int c = 1; c < SH::C; ++c) r |= u >> (5 * c);
  return r & 31u;
}
template <class SH>
__device__ __forceinline__ uint32_t grid_colors(uint32_t u) {
  // bit 4 of a group <- "group is not zero", then bits 4, 9, 14, ... gathered into bits 0..4 by one
  // multiplication (the 25 partial products land on distinct bits: 5j - 4i is injective on 0..4)
  const uint32_t t = (((u & kLow4) + kLow4) | u) & kRank4;
  return (((t >> 4) * 0x11111u) >> 16) & ((1u << SH::C) - 1);
}

// ---- legality (A.4): 64-bit mask over move uids for the player to move ---------------------
template <class SH>
__device__ __forceinline__ unsigned long long legal_mask(const Game<SH>& g) {
  const unsigned long long hand_bits = (1ull << g.hand_n(0)) - 1;
  unsigned long long m = hand_bits << SH::H;                 // Play(i), i < hand size
  if (g.info < SH::MAX_INFO) m |= hand_bits;                 // Discard(i)
  unsigned long long hints = 0;
#pragma unroll
  for (int k = 1; k < SH::P; ++k) {
    const uint32_t u = hand_union(g, k);                     // what the target hand contains
    hints |= (unsigned long long)grid_colors<SH>(u) << (2 * SH::H + (k - 1) * SH::C);
    hints |= (unsigned long long)grid_ranks<SH>(u) << (2 * SH::H + (SH::P - 1) * SH::C + (k - 1) * SH::R);
  }
  if (g.info > 0) m |= hints;
  return m;
}

// ---- transition (A.3): ApplyMove + chance deal + AdvanceToNextPlayer -------------------
template <class SH>
__device__ __forceinline__ void rotate_once(Game<SH>& g) {
  uint32_t first[SH::H], first_oh[SH::H];
#pragma unroll
  for (int i = 0; i < SH::H; ++i) { first[i] = g.hand[0][i]; first_oh[i] = g.oh[0][i]; }
#pragma unroll
  for (int k = 0; k + 1 < SH::P; ++k)
#pragma unroll
    for (int i = 0; i < SH::H; ++i) { g.hand[k][i] = g.hand[k + 1][i]; g.oh[k][i] = g.oh[k + 1][i]; }
#pragma unroll
  for (int i = 0; i < SH::H; ++i) { g.hand[SH::P - 1][i] = first[i]; g.oh[SH::P - 1][i] = first_oh[i]; }
  g.hn = (g.hn >> 3) | ((g.hn & 7u) << (3 * (SH::P - 1)));
}

// What a move uid means to the transition: lo = cells of the hinted colour / rank | hinted flag
// (bit 25 colour, bit 26 rank), 0 for a card move; hi = [2:0] hand index (H: not a card move),
// [3] card move, [4] play, [7:5] target seat.  The fused kernel looks this up in shared memory
// (one LDS.64 instead of ~20 ALU instructions per game step).
template <class SH>
__device__ __forceinline__ unsigned long long decode_move(int uid) {
  constexpr int P = SH::P, H = SH::H, C = SH::C, R = SH::R;
  const bool card_move = uid < 2 * H;
  const bool is_play = card_move && uid >= H;
  const int idx = card_move ? (is_play ? uid - H : uid) : H;
  int u = uid - 2 * H;
  const bool by_color = u < (P - 1) * C;
  if (!by_color) u -= (P - 1) * C;
  const int span = by_color ? C : R;
  const int k = 1 + u / span, v = u - (k - 1) * span;
  uint32_t lo = (by_color ? color_cells(v) : rank_cells(v)) | (by_color ? (1u << 25) : (1u << 26));
  uint32_t hi = (uint32_t)idx | ((uint32_t)k << 5);
  if (card_move) { lo = 0; hi = (uint32_t)idx | 8u | (is_play ? 16u : 0u); }
  return (unsigned long long)lo | ((unsigned long long)hi << 32);
}
struct MoveTable { unsigned long long e[64]; };
template <class SH>
__device__ __forceinline__ void stage_move_table(MoveTable& mt) {   // caller supplies the barrier
  for (int j = threadIdx.x; j < 64; j += blockDim.x) mt.e[j] = j < SH::A ? decode_move<SH>(j) : 0ull;
}

// `uid` must be legal.  Returns the reward (score delta).  Ends with the seat rotation.
// `next_card` = deck[g.deck_pos] (whatever when the deck is empty): the caller loads it ahead of time
// so that the chance node does not wait for memory.
// Written without branches: a warp nearly always holds games making every kind of move, so both
// sides would be walked anyway; as selects they are one basic block the scheduler can interleave.
template <class SH>
__device__ __forceinline__ int apply_move(Game<SH>& g, int uid, uint32_t next_card,
                                          const MoveTable* mt = nullptr) {
  constexpr int P = SH::P, H = SH::H;
  const int before = __popc(g.fw);                 // the game is running: life >= 1
  const bool deck_empty = g.deck_pos >= SH::DECK;
  if (deck_empty) --g.ttp;                         // evaluated BEFORE the move
  const uint32_t pg = g.playable();
  const unsigned long long mv = mt != nullptr ? mt->e[uid] : decode_move<SH>(uid);
  const uint32_t mv_lo = (uint32_t)mv, mv_hi = (uint32_t)(mv >> 32);
  const bool card_move = (mv_hi & 8u) != 0;
  const bool is_play = (mv_hi & 16u) != 0;
  const int idx = (int)(mv_hi & 7u);
  uint32_t last = 1u | ((uint32_t)uid << 1);

  // ---- Play / Discard of hand[0][idx] (idx = H: not a card move, nothing shifts)
  uint32_t w = 0;
#pragma unroll
  for (int i = 0; i < H; ++i)
    if (i == idx) w = g.hand[0][i];
  const uint32_t card = w >> 27;
  const uint32_t oh = card_move ? (1u << card) : 0u;
  const bool fits = (oh & pg) != 0;
  const bool succ = is_play && fits;               // AddToFireworks
  const bool fail = is_play && !fits;
  const bool room = g.info < SH::MAX_INFO;
  // IncrementInformationTokens: a discard, or the last card of a stack
  const bool token = room && ((card_move && !is_play) || (succ && (oh & kRank4) != 0));
  if (card_move) last |= card << 7;
  if (succ) { g.fw |= oh; last |= 1u << 12; }
  if (token) { ++g.info; last |= 1u << 13; }
  if (fail) --g.life;
  if (card_move && !succ) {                        // to the discard pile
    g.disc += 1ull << (2 * card);
    const uint32_t carry = g.d0 & oh;
    g.d0 ^= oh;
    g.d1 ^= carry;
  }
  // RemoveFromHand: later cards (and their knowledge) shift down; the chance node puts the new card
  // at the END of the mover's hand.  While the deck lasts every hand is full, so that is slot H-1.
#pragma unroll
  for (int i = 0; i < H - 1; ++i)
    if (i >= idx) { g.hand[0][i] = g.hand[0][i + 1]; g.oh[0][i] = g.oh[0][i + 1]; }
  if (card_move) {
    g.hand[0][H - 1] = deck_empty ? 0u : fresh_slot<SH>((int)next_card);
    g.oh[0][H - 1] = deck_empty ? 0u : 1u << next_card;
    if (deck_empty) g.hn -= 1u; else ++g.deck_pos;
  }

  // ---- RevealColor / RevealRank: matching cards keep only the hinted cells and get the hinted
  // flag; the others lose those cells (HanabiHand::RevealColor / RevealRank)
  const uint32_t cells = mv_lo & kGridMask, flag = mv_lo & ~kGridMask;   // 0 for a card move
  const int k = (int)(mv_hi >> 5);
  if (!card_move) --g.info;
  uint32_t mask = 0;
#pragma unroll
  for (int kk = 1; kk < P; ++kk) {
    const uint32_t kcells = (P == 2 || kk == k) ? cells : 0u;
    const uint32_t kflag = (P == 2 || kk == k) ? flag : 0u;
#pragma unroll
    for (int i = 0; i < H; ++i) {
      const uint32_t s = g.hand[kk][i];
      const bool match = (g.oh[kk][i] & kcells) != 0;
      g.hand[kk][i] = match ? ((s & (kcells | ~kGridMask)) | kflag) : (s & ~kcells);
      if (match) mask |= 1u << i;
    }
  }
  last |= mask << 14;

  g.last = last;
  // AdvanceToNextPlayer: the next player becomes seat 0
  g.cur = g.cur + 1 == P ? 0 : g.cur + 1;
  rotate_once(g);
  return g.score() - before;
}

// View from another seat: rotate a copy so that seat `offset` becomes hand[0].
template <class SH>
__device__ __forceinline__ void rotate_seats(Game<SH>& g, int offset) {
#pragma unroll
  for (int step = 0; step + 1 < SH::P; ++step)
    if (step < offset) rotate_once(g);
}

}  // namespace hbg
