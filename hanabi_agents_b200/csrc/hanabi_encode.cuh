// Canonical observation encoder (HLE CanonicalObservationEncoder),
// bit-packed: observation bit b lives in word b>>5, bit b&31 of the game's row.
//
// The warp assembles the row in a small shared-memory bit buffer: slot lanes OR in the bits
// that depend on their card (hands section, card-knowledge section), then lanes 0..12 each OR
// in one warp-uniform field (board, discards, last action).  The finished row leaves with one
// coalesced store (row stride is a multiple of 16 B).
#pragma once
#include "hanabi_game.cuh"

namespace hbg {

// OR the low bits of v into the bit buffer at bit offset `off` (v spans at most 2 words).
__device__ __forceinline__ void put_bits(uint32_t* buf, int off, uint32_t v) {
  if (v == 0) return;
  const int w = off >> 5, sh = off & 31;
  atomicOr(&buf[w], v << sh);
  if (sh != 0) {
    const uint32_t hi = v >> (32 - sh);
    if (hi) atomicOr(&buf[w + 1], hi);
  }
}

// Observation of player q (observer-relative offsets k = (p - q) mod P).
template <class SH>
__device__ __forceinline__ void encode_obs(const Game<SH>& g, const Rt& rt, int q, uint32_t* buf) {
  constexpr int P = SH::P, H = SH::H, C = SH::C, R = SH::R, BPC = SH::BPC;
  const int lane = lane_id();
  for (int j = lane; j < rt.obs_stride; j += 32) buf[j] = 0;
  __syncwarp();

  // ---- slot lanes: hands + card knowledge -------------------------------------------
  if (lane < SH::S && slot_valid(g.slot)) {
    const int p = lane / H, i = lane - p * H;
    int k = p - q;
    if (k < 0) k += P;
    const int card = slot_card(g.slot);
    if (k >= 1) {                                       // EncodeHands: others' cards only
      const int off = (k - 1) * H * BPC + i * BPC + card;
      atomicOr(&buf[off >> 5], 1u << (off & 31));
    }
    const int cp = slot_cplaus(g.slot), rp = slot_rplaus(g.slot);
    uint32_t grid = 0;                                  // EncodeCardKnowledge: plausible(c)&plausible(r)
#pragma unroll
    for (int c = 0; c < C; ++c)
      if ((cp >> c) & 1) grid |= (uint32_t)rp << (c * R);
    const int base = rt.off_know + (k * H + i) * (BPC + C + R);
    put_bits(buf, base, grid);
    const uint32_t hints = (slot_chint(g.slot) ? (uint32_t)cp : 0u) |
                           ((slot_rhint(g.slot) ? (uint32_t)rp : 0u) << C);
    put_bits(buf, base + BPC, hints);
  }

  // ---- warp-uniform fields, one per lane ----------------------------------------------
  const uint32_t last = g.last;
  const bool lv = last & 1u;
  const int l_uid = (last >> 4) & 63;
  const bool l_card_move = l_uid < 2 * H;
  switch (lane) {
    case 0: {                                           // "hand is short" bit per player
      uint32_t v = 0;
#pragma unroll
      for (int k = 0; k < P; ++k) {
        int p = q + k;
        if (p >= P) p -= P;
        if (g.hand_n(p) < H) v |= 1u << k;
      }
      put_bits(buf, (P - 1) * H * BPC, v);
      break;
    }
    case 1: {                                           // EncodeBoard: deck thermometer, low word
      const int n = min(SH::DECK - g.deck_pos, 32);
      put_bits(buf, rt.off_board, n >= 32 ? 0xffffffffu : ((1u << n) - 1));
      break;
    }
    case 2: {                                           // deck thermometer, high word
      const int n = SH::DECK - g.deck_pos - 32;
      if (n > 0) put_bits(buf, rt.off_board + 32, (1u << n) - 1);
      break;
    }
    case 3: {                                           // fireworks: one-hot of height-1
      uint32_t v = 0;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const int f = g.firework(c);
        if (f > 0) v |= 1u << (c * R + f - 1);
      }
      put_bits(buf, rt.off_board + rt.deck_thermo, v);
      break;
    }
    case 4: {                                           // info + life thermometers
      const uint32_t v = ((1u << g.info) - 1) | (((1u << g.life) - 1) << rt.max_info);
      put_bits(buf, rt.off_board + rt.deck_thermo + BPC, v);
      break;
    }
    case 5: case 6: case 7: case 8: case 9: {           // EncodeDiscards, one colour per lane
      const int c = lane - 5;
      if (c < C) {
        uint32_t v = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) v |= ((1u << g.disc_n(c * R + r)) - 1) << SH::disc_off(r);
        put_bits(buf, rt.off_disc + c * SH::kPerColor, v);
      }
      break;
    }
    case 10: {                                          // EncodeLastAction: who / what / whom
      if (lv) {
        int rel = (int)((last >> 1) & 7) - q;
        if (rel < 0) rel += P;
        uint32_t v = 1u << rel;
        if (l_card_move) {
          v |= 1u << (P + (l_uid >= H ? 0 : 1));        // Play, Discard
        } else {
          int u = l_uid - 2 * H;
          const bool by_color = u < (P - 1) * C;
          if (!by_color) u -= (P - 1) * C;
          const int span = by_color ? C : R;
          const int k = 1 + u / span, val = u - (k - 1) * span;
          v |= 1u << (P + (by_color ? 2 : 3));          // RevealColor, RevealRank
          int tgt = rel + k;
          if (tgt >= P) tgt -= P;
          v |= 1u << (P + 4 + tgt);
          v |= 1u << (2 * P + 4 + (by_color ? 0 : C) + val);
        }
        put_bits(buf, rt.off_last, v);
      }
      break;
    }
    case 11: {                                          // reveal outcome | card position
      if (lv) {
        const uint32_t v = l_card_move ? (1u << (H + (l_uid >= H ? l_uid - H : l_uid)))
                                       : ((last >> 18) & ((1u << H) - 1));
        put_bits(buf, rt.off_last + 2 * P + 4 + C + R, v);
      }
      break;
    }
    case 12: {                                          // card identity | scored, info token
      if (lv && l_card_move) {
        const int c = (last >> 10) & 7, r = (last >> 13) & 7;
        uint32_t v = 1u << (c * R + r);
        if (l_uid >= H) v |= ((last >> 16) & 3u) << BPC; // play only
        put_bits(buf, rt.off_last + 2 * P + 4 + C + R + 2 * H, v);
      }
      break;
    }
    default: break;
  }
  __syncwarp();
}

// buffer -> global row (coalesced; stride is a multiple of 4 words)
__device__ __forceinline__ void store_obs(const Rt& rt, const uint32_t* buf, uint32_t* __restrict__ obs,
                                          long long game) {
  uint32_t* row = obs + game * rt.obs_stride;
  for (int j = lane_id(); j < rt.obs_stride; j += 32) row[j] = buf[j];
  __syncwarp();
}

}  // namespace hbg
