// Canonical observation encoder (HLE CanonicalObservationEncoder),
// bit-packed: observation bit b lives in word b>>5, bit b&31 of the game's row.
//
// Thread-per-game, observer = seat 0 (hands are seat-relative, see hanabi_game.cuh), so the
// bit offset of EVERY field of the vector is a compile-time constant: the row is assembled in
// registers by OR-ing each field into one or two words with constant shifts, and leaves as
// 16-byte vector stores (row stride is a multiple of 16 B).
#pragma once
#include "hanabi_game.cuh"

namespace hbg {

// multiply-add forms of "shift a field into place and merge it": v * 2^SH + acc.  The fields of a row are
// disjoint, so + is |, and IMAD / IMAD.HI issue on the FMA pipe, which the bit logic of this kernel
// leaves idle, instead of SHF + LOP3 on the ALU pipe that bounds it (profiles/r01_selfplay_history.md).
// The multipliers come from constant memory: a literal power of two is strength-reduced by ptxas back
// into SHF / LEA on the ALU pipe; an operand it cannot see through stays an IMAD.
__constant__ uint32_t g_pow2[32] = {
    1u << 0,  1u << 1,  1u << 2,  1u << 3,  1u << 4,  1u << 5,  1u << 6,  1u << 7,
    1u << 8,  1u << 9,  1u << 10, 1u << 11, 1u << 12, 1u << 13, 1u << 14, 1u << 15,
    1u << 16, 1u << 17, 1u << 18, 1u << 19, 1u << 20, 1u << 21, 1u << 22, 1u << 23,
    1u << 24, 1u << 25, 1u << 26, 1u << 27, 1u << 28, 1u << 29, 1u << 30, 1u << 31};
template <int SH>
__device__ __forceinline__ uint32_t shl_add(uint32_t v, uint32_t acc) {      // (v << SH) + acc
  if constexpr (SH == 0) return v + acc;
  return v * g_pow2[SH] + acc;
}
template <int SH>
__device__ __forceinline__ uint32_t shr_add(uint32_t v, uint32_t acc) {      // (v >> (32 - SH)) + acc
  static_assert(SH >= 1 && SH <= 31, "shr_add");
  return __umulhi(v, g_pow2[SH]) + acc;
}

// Merge a WIDTH-bit field into the row at compile-time bit offset OFF (the target bits are still 0).
template <int OFF, int WIDTH, int N>
__device__ __forceinline__ void put(uint32_t (&row)[N], unsigned long long v) {
  constexpr int idx = OFF / 32, sh = OFF % 32;
  static_assert(WIDTH >= 1 && WIDTH <= 64 && idx < N, "field outside the row");
  const uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
  row[idx] = shl_add<sh>(lo, row[idx]);
  if constexpr (sh + WIDTH > 32) {
    static_assert(idx + 1 < N, "field outside the row");
    if constexpr (sh > 0) row[idx + 1] = shr_add<sh>(lo, row[idx + 1]);
    if constexpr (WIDTH > 32) row[idx + 1] = shl_add<sh>(hi, row[idx + 1]);
  }
  if constexpr (sh + WIDTH > 64) {
    static_assert(idx + 2 < N && sh > 0, "field outside the row");
    row[idx + 2] = shr_add<sh>(hi, row[idx + 2]);
  }
}

// EncodeDiscards for every colour at once.  Counts are 2-bit fields at bit 10c+2r; the
// thermometers sit at 10c+{0,3,5,7,9} with 3/2/2/2/1 levels: rank 0 keeps its position,
// ranks 1..4 move up by one, level L sits L bits higher.
template <class SH>
__device__ __forceinline__ unsigned long long discard_thermometers(unsigned long long disc) {
  unsigned long long m0 = 0, m14 = 0, m13 = 0;
#pragma unroll
  for (int c = 0; c < SH::C; ++c) {
    m0 |= 1ull << (10 * c);
    m14 |= 0x154ull << (10 * c);
    m13 |= 0x054ull << (10 * c);
  }
  const unsigned long long lo = disc & 0x5555555555555555ull, hi = (disc >> 1) & 0x5555555555555555ull;
  const unsigned long long ge1 = lo | hi, ge2 = hi, eq3 = lo & hi;
  return (ge1 & m0) | ((ge1 & m14) << 1) | ((ge2 & m0) << 1) | ((ge2 & m13) << 2) | ((eq3 & m0) << 2);
}

template <class SH, int K, int I, int N>
__device__ __forceinline__ void put_hand_card(uint32_t (&row)[N], uint32_t oh) {
  // EncodeHands: one-hot card index of another player's card
  put<(K - 1) * SH::H * SH::BPC + I * SH::BPC, SH::BPC>(row, oh);
}

// ---- lookup tables in shared memory ------------------------------------------------------------
// Two pieces of the encoder are pure functions of a few bits: the hinted colour / rank one-hots of a
// card slot (of the slot word's top 7 bits: card index, hinted flags) and the static part of the last
// action (of the move uid).  Computed, they are ~8 and ~30 LOP3 / SHF / SEL each per game step on the
// ALU pipe that bounds the fused kernel; looked up, one LDS.64 on the otherwise idle load/store pipe.
// The tables are filled at kernel start by the SAME functions the table-free path evaluates per step.
struct EncTables {
  unsigned long long know[128];   // [w >> 25]: hint one-hots at bits BPC.. of the knowledge field
  unsigned long long last[128];   // [last & 127] (= 1 | uid << 1, or 0): lo = bits of `a`;
                                  // hi = [2H-1:0] card-position bit, [12] card move, [14:13] play mask
};

// hinted colour / rank one-hots of a slot word, positioned inside the card's knowledge field
template <class SH>
__device__ __forceinline__ unsigned long long know_hints(uint32_t w) {
  // A hint names the card's true colour / rank, so the one-hots come from the card itself.
  const int card = slot_card(w);
  return ((unsigned long long)slot_chint(w) << (SH::BPC + card_color(card))) |
         ((unsigned long long)slot_rhint(w) << (SH::BPC + SH::C + card_rank(card)));
}

// static part of EncodeLastAction for move `uid` made by the player at observer-relative `mover_seat`:
// a = acting player, move type, target player, colour / rank; pos = card position bit (card moves)
template <class SH>
__device__ __forceinline__ void last_action_static(int uid, int mover_seat, uint32_t& a, uint32_t& pos,
                                                   bool& card_move, bool& is_play) {
  constexpr int P = SH::P, H = SH::H, C = SH::C, R = SH::R;
  a = 1u << mover_seat;
  pos = 0;
  card_move = uid < 2 * H;
  is_play = card_move && uid >= H;
  if (card_move) {
    a |= 1u << (P + (is_play ? 0 : 1));               // move type: Play, Discard
    pos = 1u << (H + (is_play ? uid - H : uid));      // card position
  } else {
    int u = uid - 2 * H;
    const bool by_color = u < (P - 1) * C;
    if (!by_color) u -= (P - 1) * C;
    const int span = by_color ? C : R;
    const int k = 1 + u / span, val = u - (k - 1) * span;
    a |= 1u << (P + (by_color ? 2 : 3));              // RevealColor, RevealRank
    int tgt = mover_seat + k;                         // observer-relative target
    if (tgt >= P) tgt -= P;
    a |= 1u << (P + 4 + tgt);
    a |= 1u << (2 * P + 4 + (by_color ? 0 : C) + val);
  }
}

// Fill the tables (mover of the last action = the player before the observer, as in env.step).  Every
// thread of the CTA must call it; the caller supplies the barrier.
template <class SH>
__device__ __forceinline__ void stage_enc_tables(EncTables& tb) {
  for (int j = threadIdx.x; j < 128; j += blockDim.x) {
    tb.know[j] = know_hints<SH>((uint32_t)j << 25);
    unsigned long long e = 0;
    const int uid = j >> 1;
    if ((j & 1) && uid < SH::A) {
      uint32_t a, pos;
      bool card_move, is_play;
      last_action_static<SH>(uid, SH::P - 1, a, pos, card_move, is_play);
      const uint32_t hi = pos | (card_move ? 1u << 12 : 0u) | (is_play ? 3u << 13 : 0u);
      e = (unsigned long long)a | ((unsigned long long)hi << 32);
    }
    tb.last[j] = e;
  }
}
template <class SH, int K, int I, int N>
__device__ __forceinline__ void put_knowledge(uint32_t (&row)[N], uint32_t w, const EncTables* tb) {
  // EncodeCardKnowledge: plausible grid | one-hot hinted colour | one-hot hinted rank.
  const unsigned long long f = slot_grid(w) | (tb != nullptr ? tb->know[w >> 25] : know_hints<SH>(w));
  put<SH::kOffKnow + (K * SH::H + I) * SH::kKnowWidth, SH::kKnowWidth>(row, f);
}

template <class SH, int K, int N>
__device__ __forceinline__ void put_seat(uint32_t (&row)[N], const Game<SH>& g, const EncTables* tb) {
  if constexpr (K < SH::P) {
#define HB_CARD(I)                                                        \
    if constexpr (I < SH::H) {                                            \
      if constexpr (K >= 1) put_hand_card<SH, K, I>(row, g.oh[K][I]);     \
      put_knowledge<SH, K, I>(row, g.hand[K][I], tb);                     \
    }
    HB_CARD(0) HB_CARD(1) HB_CARD(2) HB_CARD(3) HB_CARD(4)
#undef HB_CARD
    put_seat<SH, K + 1>(row, g, tb);
  }
}

// Observation of the player at seat 0, `mover_seat` = seat of the player who made g.last
// (P-1 for the player to move), assembled in registers: kObsStride words.
// `tb` (a compile-time null or not at every call site): the shared-memory tables, valid for
// mover_seat == P-1 only.
template <class SH>
__device__ __forceinline__ void encode_row(const Game<SH>& g, int mover_seat, uint32_t (&row)[SH::kObsStride],
                                           const EncTables* tb = nullptr) {
  constexpr int P = SH::P, H = SH::H, C = SH::C, R = SH::R, BPC = SH::BPC;
#pragma unroll
  for (int i = 0; i < SH::kObsStride; ++i) row[i] = 0;

  put_seat<SH, 0>(row, g, tb);                          // hands + card knowledge

  {
    uint32_t v = 0;                                     // EncodeHands: "hand is short" bits
#pragma unroll
    for (int k = 0; k < P; ++k)
      if (g.hand_n(k) < H) v |= 1u << k;
    put<SH::kHandsCards, P>(row, v);
  }
  {                                                     // EncodeBoard
    const int rem = SH::DECK - g.deck_pos;
    put<SH::kOffBoard, SH::kDeckThermo>(row, (1ull << rem) - 1);
    // fireworks: one-hot of height-1 = the top bit of each colour's thermometer
    put<SH::kOffFireworks, BPC>(row, g.fw & ~((g.fw >> 1) & kLow4));
    put<SH::kOffInfo, SH::MAX_INFO + SH::MAX_LIFE>(
        row, ((1u << g.info) - 1) | (((1u << g.life) - 1) << SH::MAX_INFO));
  }
  put<SH::kOffDisc, SH::DECK>(row, discard_thermometers<SH>(g.disc));   // EncodeDiscards

  {                                                     // EncodeLastAction
    // g.last: [0] valid [6:1] uid [11:7] card [12] scored [13] info token [18:14] reveal bitmask
    // (card / reveal bits are 0 unless the move was a card move / a hint)
    const uint32_t last = g.last;
    uint32_t a, b, c3;
    if (tb != nullptr) {
      const unsigned long long e = tb->last[last & 127u];
      const uint32_t hi = (uint32_t)(e >> 32);
      a = (uint32_t)e;
      b = (hi & ((1u << (2 * H)) - 1)) | ((last >> 14) & ((1u << H) - 1));
      c3 = (((hi >> 12) & 1u) << ((last >> 7) & 31)) | (((last & (hi >> 1)) >> 12) << BPC);
    } else {
      uint32_t pos;
      bool card_move, is_play;
      last_action_static<SH>((last >> 1) & 63, mover_seat, a, pos, card_move, is_play);
      b = pos | ((last >> 14) & ((1u << H) - 1));       // card position | reveal outcome
      c3 = card_move ? 1u << ((last >> 7) & 31) : 0u;   // card identity
      if (is_play) c3 |= ((last >> 12) & 3u) << BPC;    // scored, information token
      if (!(last & 1u)) a = b = c3 = 0;
    }
    put<SH::kOffLast, 2 * P + 4 + C + R>(row, a);
    put<SH::kOffLast + 2 * P + 4 + C + R, 2 * H>(row, b);
    put<SH::kOffLast + 2 * P + 4 + C + R + 2 * H, BPC + 2>(row, c3);
  }
}

// ... and written as kObsStride words at `out` (16-byte stores).
template <class SH>
__device__ __forceinline__ void encode_obs(const Game<SH>& g, int mover_seat, uint32_t* __restrict__ out) {
  uint32_t row[SH::kObsStride];
  encode_row(g, mover_seat, row);
  uint4* dst = reinterpret_cast<uint4*>(out);
#pragma unroll
  for (int i = 0; i < SH::kObsStride / 4; ++i)
    dst[i] = make_uint4(row[4 * i], row[4 * i + 1], row[4 * i + 2], row[4 * i + 3]);
}

// Stores that are switched off by a predicate instead of a branch: the fused self-play kernel keeps
// {encoder, legal moves, rule bodies} in ONE basic block so that the scheduler can interleave their
// independent dependency chains; `if (ptr != nullptr)` around the encoder would cut the block.
__device__ __forceinline__ void st_v4_if(bool p, uint32_t* ptr, uint32_t a, uint32_t b, uint32_t c,
                                         uint32_t d) {
  asm volatile("{ .reg .pred q; setp.ne.u32 q, %5, 0; @q st.global.v4.u32 [%0], {%1, %2, %3, %4}; }"
               :: "l"(ptr), "r"(a), "r"(b), "r"(c), "r"(d), "r"((uint32_t)p));
}
__device__ __forceinline__ void st_u32_if(bool p, uint32_t* ptr, uint32_t a) {
  asm volatile("{ .reg .pred q; setp.ne.u32 q, %2, 0; @q st.global.u32 [%0], %1; }"
               :: "l"(ptr), "r"(a), "r"((uint32_t)p));
}
__device__ __forceinline__ void st_u8_if(bool p, uint8_t* ptr, uint32_t a) {
  asm volatile("{ .reg .pred q; setp.ne.u32 q, %2, 0; @q st.global.u8 [%0], %1; }"
               :: "l"(ptr), "r"(a), "r"((uint32_t)p));
}

}  // namespace hbg
