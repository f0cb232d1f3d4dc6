// Canonical observation encoder (HLE CanonicalObservationEncoder),
// bit-packed: observation bit b lives in word b>>5, bit b&31 of the game's row.
//
// Lane W assembles output word W directly in a register -- no shared memory, no atomics:
//   * the warp-uniform middle of the vector (short-hand bits, board, discards, last action)
//     is built once as a handful of 32-bit words with compile-time bit offsets;
//   * the card-dependent ends (hands section, card-knowledge section) are gathered from the
//     slot lanes with shuffles: a 32-bit window overlaps at most kHandTaps one-hot card
//     fields and kKnowTaps knowledge fields.
// The finished row leaves with one coalesced store (row stride is a multiple of 16 B).
#pragma once
#include "hanabi_game.cuh"

namespace hbg {

template <int OFF, int WIDTH, int W0, int NU>
__device__ __forceinline__ void or_field(uint32_t (&U)[NU], uint32_t v) {
  constexpr int idx = OFF / 32 - W0, sh = OFF % 32;
  static_assert(idx >= 0 && idx < NU, "field outside the uniform window");
  U[idx] |= v << sh;
  if constexpr (sh + WIDTH > 32) {
    static_assert(idx + 1 < NU, "field crosses the end of the uniform window");
    U[idx + 1] |= v >> (32 - sh);
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

// Observation of player q (observer-relative offsets k = (p - q) mod P), written to `row`.
template <class SH>
__device__ __forceinline__ void encode_obs(const Game<SH>& g, int q, uint32_t* __restrict__ row) {
  constexpr int P = SH::P, H = SH::H, C = SH::C, R = SH::R, BPC = SH::BPC;
  constexpr int KW = SH::kKnowWidth;
  constexpr int W0 = SH::kHandsCards / 32, W1 = (SH::kOffKnow - 1) / 32, NU = W1 - W0 + 1;
  constexpr int kHandTaps = (32 + BPC - 2) / BPC + 1;
  constexpr int kKnowTaps = (32 + KW - 2) / KW + 1;
  const int lane = lane_id();

  // ---- this lane's card-knowledge field: plausible grid | hinted colour | hinted rank ----
  unsigned long long know = 0;
  if (lane < SH::S && slot_valid(g.slot)) {
    const uint32_t cp = slot_cplaus(g.slot), rp = slot_rplaus(g.slot);
    uint32_t grid = 0;
#pragma unroll
    for (int c = 0; c < C; ++c)
      if ((cp >> c) & 1) grid |= rp << (c * R);
    const uint32_t hints = (slot_chint(g.slot) ? cp : 0u) | ((slot_rhint(g.slot) ? rp : 0u) << C);
    know = (unsigned long long)grid | ((unsigned long long)hints << BPC);
  }
  const uint32_t know_lo = (uint32_t)know, know_hi = (uint32_t)(know >> 32);

  // ---- warp-uniform words W0..W1 ---------------------------------------------------------
  uint32_t U[NU];
#pragma unroll
  for (int i = 0; i < NU; ++i) U[i] = 0;
  {
    uint32_t v = 0;                                     // EncodeHands: "hand is short" bits
#pragma unroll
    for (int k = 0; k < P; ++k) {
      int p = q + k;
      if (p >= P) p -= P;
      if (g.hand_n(p) < H) v |= 1u << k;
    }
    or_field<SH::kHandsCards, P, W0>(U, v);
  }
  {
    const int rem = SH::DECK - g.deck_pos;              // EncodeBoard: deck thermometer
    constexpr int LO = SH::kDeckThermo < 32 ? SH::kDeckThermo : 32;
    or_field<SH::kOffBoard, LO, W0>(U, rem >= 32 ? 0xffffffffu : ((1u << rem) - 1));
    if constexpr (SH::kDeckThermo > 32)
      or_field<SH::kOffBoard + 32, SH::kDeckThermo - 32, W0>(U, rem > 32 ? ((1u << (rem - 32)) - 1) : 0u);
  }
  {
    uint32_t v = 0;                                     // fireworks: one-hot of height-1
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int f = g.firework(c);
      if (f > 0) v |= 1u << (c * R + f - 1);
    }
    or_field<SH::kOffFireworks, BPC, W0>(U, v);
    or_field<SH::kOffInfo, SH::MAX_INFO + SH::MAX_LIFE, W0>(
        U, ((1u << g.info) - 1) | (((1u << g.life) - 1) << SH::MAX_INFO));
  }
  {
    const unsigned long long t = discard_thermometers<SH>(g.disc);   // EncodeDiscards
    constexpr int LO = SH::DECK < 32 ? SH::DECK : 32;
    or_field<SH::kOffDisc, LO, W0>(U, (uint32_t)t);
    if constexpr (SH::DECK > 32) or_field<SH::kOffDisc + 32, SH::DECK - 32, W0>(U, (uint32_t)(t >> 32));
  }
  if (g.last & 1u) {                                    // EncodeLastAction
    const uint32_t last = g.last;
    const int uid = (last >> 4) & 63;
    int rel = (int)((last >> 1) & 7) - q;
    if (rel < 0) rel += P;
    uint32_t a = 1u << rel, b, c3 = 0;
    if (uid < 2 * H) {
      const bool is_play = uid >= H;
      a |= 1u << (P + (is_play ? 0 : 1));               // move type: Play, Discard
      b = 1u << (H + (is_play ? uid - H : uid));        // card position
      c3 = 1u << (((last >> 10) & 7) * R + ((last >> 13) & 7));   // card identity
      if (is_play) c3 |= ((last >> 16) & 3u) << BPC;    // scored, information token
    } else {
      int u = uid - 2 * H;
      const bool by_color = u < (P - 1) * C;
      if (!by_color) u -= (P - 1) * C;
      const int span = by_color ? C : R;
      const int k = 1 + u / span, val = u - (k - 1) * span;
      a |= 1u << (P + (by_color ? 2 : 3));              // RevealColor, RevealRank
      int tgt = rel + k;
      if (tgt >= P) tgt -= P;
      a |= 1u << (P + 4 + tgt);
      a |= 1u << (2 * P + 4 + (by_color ? 0 : C) + val);
      b = (last >> 18) & ((1u << H) - 1);               // reveal outcome
    }
    or_field<SH::kOffLast, 2 * P + 4 + C + R, W0>(U, a);
    or_field<SH::kOffLast + 2 * P + 4 + C + R, 2 * H, W0>(U, b);
    or_field<SH::kOffLast + 2 * P + 4 + C + R + 2 * H, BPC + 2, W0>(U, c3);
  }

  // ---- one output word per lane (two rounds when the row is longer than 32 words) ---------
#pragma unroll
  for (int base = 0; base < SH::kObsStride; base += 32) {
    const int W = base + lane;
    uint32_t word = 0;
#pragma unroll
    for (int i = 0; i < NU; ++i)
      if (W == W0 + i) word = U[i];

    if (base * 32 < SH::kHandsCards) {                  // EncodeHands: others' cards, one-hot
      const int j0 = (32 * W) / BPC;
#pragma unroll
      for (int t = 0; t < kHandTaps; ++t) {
        const int j = j0 + t;
        const bool ok = j < (P - 1) * H;
        const int k = 1 + j / H, i = j - (k - 1) * H;
        int p = q + k;
        if (p >= P) p -= P;
        const uint32_t sw = __shfl_sync(kFull, g.slot, ok ? p * H + i : 0);
        const int bit = j * BPC + slot_card(sw) - 32 * W;
        if (ok && slot_valid(sw) && (unsigned)bit < 32u) word |= 1u << bit;
      }
    }
    if ((base + 32) * 32 > SH::kOffKnow) {              // EncodeCardKnowledge
      const int b0 = 32 * W - SH::kOffKnow;
      const int j0 = b0 > 0 ? b0 / KW : 0;
#pragma unroll
      for (int t = 0; t < kKnowTaps; ++t) {
        const int j = j0 + t;
        const bool ok = j < P * H;
        const int k = j / H, i = j - k * H;
        int p = q + k;
        if (p >= P) p -= P;
        const int src = ok ? p * H + i : 0;
        const uint32_t lo = __shfl_sync(kFull, know_lo, src);
        const uint32_t hi = KW > 32 ? __shfl_sync(kFull, know_hi, src) : 0u;
        const unsigned long long f = (unsigned long long)lo | ((unsigned long long)hi << 32);
        const int rel = j * KW - b0;                    // field bit 0 relative to this word
        uint32_t add = 0;
        if (rel >= 0) { if (rel < 32) add = (uint32_t)(f << rel); }
        else if (rel > -64) add = (uint32_t)(f >> (-rel));
        if (ok) word |= add;
      }
    }
    if (W < SH::kObsStride) row[W] = word;
  }
}

}  // namespace hbg
