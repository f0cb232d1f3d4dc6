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

// OR a WIDTH-bit field into the row at compile-time bit offset OFF.
template <int OFF, int WIDTH, int N>
__device__ __forceinline__ void put(uint32_t (&row)[N], unsigned long long v) {
  constexpr int idx = OFF / 32, sh = OFF % 32;
  static_assert(WIDTH >= 1 && WIDTH <= 64 && idx < N, "field outside the row");
  row[idx] |= (uint32_t)(v << sh);
  if constexpr (sh + WIDTH > 32) {
    static_assert(idx + 1 < N, "field outside the row");
    row[idx + 1] |= (uint32_t)(v >> (32 - sh));
  }
  if constexpr (sh + WIDTH > 64) {
    static_assert(idx + 2 < N, "field outside the row");
    row[idx + 2] |= (uint32_t)(v >> (64 - sh));
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
__device__ __forceinline__ void put_hand_card(uint32_t (&row)[N], uint32_t w) {
  // EncodeHands: one-hot card index of another player's card
  put<(K - 1) * SH::H * SH::BPC + I * SH::BPC, SH::BPC>(row, slot_onehot(w));
}

template <class SH, int K, int I, int N>
__device__ __forceinline__ void put_knowledge(uint32_t (&row)[N], uint32_t w) {
  // EncodeCardKnowledge: plausible grid | one-hot hinted colour | one-hot hinted rank.
  // A hint names the card's true colour / rank, so the one-hots come from the card itself.
  const int card = slot_card(w);
  unsigned long long f = slot_grid(w);
  f |= (unsigned long long)slot_chint(w) << (SH::BPC + card_color(card));
  f |= (unsigned long long)slot_rhint(w) << (SH::BPC + SH::C + card_rank(card));
  put<SH::kOffKnow + (K * SH::H + I) * SH::kKnowWidth, SH::kKnowWidth>(row, f);
}

template <class SH, int K, int N>
__device__ __forceinline__ void put_seat(uint32_t (&row)[N], const Game<SH>& g) {
  if constexpr (K < SH::P) {
#define HB_CARD(I)                                                        \
    if constexpr (I < SH::H) {                                            \
      if constexpr (K >= 1) put_hand_card<SH, K, I>(row, g.hand[K][I]);   \
      put_knowledge<SH, K, I>(row, g.hand[K][I]);                         \
    }
    HB_CARD(0) HB_CARD(1) HB_CARD(2) HB_CARD(3) HB_CARD(4)
#undef HB_CARD
    put_seat<SH, K + 1>(row, g);
  }
}

// Observation of the player at seat 0, `mover_seat` = seat of the player who made g.last
// (P-1 for the player to move), assembled in registers: kObsStride words.
template <class SH>
__device__ __forceinline__ void encode_row(const Game<SH>& g, int mover_seat, uint32_t (&row)[SH::kObsStride]) {
  constexpr int P = SH::P, H = SH::H, C = SH::C, R = SH::R, BPC = SH::BPC;
#pragma unroll
  for (int i = 0; i < SH::kObsStride; ++i) row[i] = 0;

  put_seat<SH, 0>(row, g);                              // hands + card knowledge

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

  if (g.last & 1u) {                                    // EncodeLastAction
    const uint32_t last = g.last;
    const int uid = (last >> 1) & 63;
    uint32_t a = 1u << mover_seat, b, c3 = 0;
    if (uid < 2 * H) {
      const bool is_play = uid >= H;
      a |= 1u << (P + (is_play ? 0 : 1));               // move type: Play, Discard
      b = 1u << (H + (is_play ? uid - H : uid));        // card position
      c3 = 1u << ((last >> 7) & 31);                    // card identity
      if (is_play) c3 |= ((last >> 12) & 3u) << BPC;    // scored, information token
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
      b = (last >> 14) & ((1u << H) - 1);               // reveal outcome
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
