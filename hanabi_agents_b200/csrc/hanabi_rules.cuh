// Rule-based agents, thread-per-game, on the seat-relative packed state (hand[0] = me).
//
// Mirrors hanabi_agents/rule_based/ruleset.py (rules) and rule_based.py:13-25 (first rule that
// returns a move wins, otherwise Ruleset.legal_random) of the reference.  The rule list is
// warp-uniform, so a warp walks the list once; a game that already has its move sits the
// remaining rules out.
//
// Documented divergences (DESIGN.md "rule-based agents"): python `random` draws are a
// counter-based stream keyed (seed, game id, step, draw); legal_random / discard_randomly use
// their intended semantics (the reference versions raise TypeError, ruleset.py:598-614);
// tell_unknown hints the real card's colour/rank (the reference builds an illegal move from
// the knowledge object, :291-307); osawa_discard keeps the fireworks[-1] quirk (:229-230);
// tell_anyone_useless_card (:522-535), tell_playable_card (:570-594) and
// discard_probably_useless (:638-649) raise TypeError in the reference and follow their intended
// semantics here.
#pragma once
#include "hanabi_game.cuh"

namespace hbg {

enum RuleId {
  R_DISCARD_OLDEST_FIRST = 0, R_OSAWA_DISCARD = 1, R_TELL_UNKNOWN = 2, R_TELL_RANDOMLY = 3,
  R_PLAY_SAFE_CARD = 4, R_PLAY_IF_CERTAIN = 5, R_TELL_PLAYABLE_CARD_OUTER = 6,
  R_TELL_DISPENSABLE = 7, R_TELL_ANYONE_USEFUL_CARD = 8, R_PLAY_PROBABLY_SAFE = 9,
  R_HAIL_MARY = 10, R_LEGAL_RANDOM = 11, R_DISCARD_RANDOMLY = 12, R_TELL_MOST_INFORMATION = 13,
  R_TELL_ANYONE_USELESS_CARD = 14, R_TELL_PLAYABLE_CARD = 15, R_DISCARD_PROBABLY_USELESS = 16,
  R_NUM_RULES = 17
};

constexpr int kMaxRules = 16;
constexpr int kThrSlots = 4;      // rules with a probability threshold that get a comparison table
constexpr int kThrRows = 52;      // denominators 0..51 (a deck holds 50 cards)
struct RuleSet {        // entry [n] is the legal_random fallback, filled by the host
  int n;
  int id[kMaxRules + 1];
  int iarg[kMaxRules + 1];
  double farg[kMaxRules + 1];
  unsigned long long seed;
  // `(double)p / (double)t >= farg` (ruleset.py:105,622 / :151,644) for every count pair, evaluated
  // once on the host in IEEE double: bit p of thr[slot][t].  The float64 division itself was a
  // ~100-instruction subroutine, called twice per game step: 22 % of all warp instructions of the
  // fused self-play kernel (profiles/r01_selfplay_history.md, v4).
  int tslot[kMaxRules + 1];       // -1: divide
  unsigned long long thr[kThrSlots][kThrRows];
};

// numerator / denominator >= threshold, exactly as the reference's float64 division decides it
__device__ __forceinline__ bool ratio_at_least(const RuleSet& rs, int slot, int num, int den,
                                               double threshold) {
  if (threshold <= 0.0) return true;                       // num >= 0, den > 0
  if (slot >= 0 && den < kThrRows && num < 64) return (rs.thr[slot][den] >> num) & 1ull;
  return (double)num / (double)den >= threshold;
}

__device__ __forceinline__ uint32_t rule_rand(unsigned long long seed, unsigned long long game,
                                              unsigned long long step, uint32_t k) {
  unsigned long long x = seed + 0x9E3779B97F4A7C15ull * (game + 1);
  x ^= step * 0xBF58476D1CE4E5B9ull + (unsigned long long)(k + 1) * 0x94D049BB133111EBull;
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return (uint32_t)(x >> 32);
}
__device__ __forceinline__ int rand_below(uint32_t r, int n) {
  return (int)(((unsigned long long)r * (unsigned long long)n) >> 32);
}

template <class SH> __device__ __forceinline__ int uid_discard(int i) { return i; }
template <class SH> __device__ __forceinline__ int uid_play(int i) { return SH::H + i; }
template <class SH> __device__ __forceinline__ int uid_color(int k, int c) {
  return 2 * SH::H + (k - 1) * SH::C + c;
}
template <class SH> __device__ __forceinline__ int uid_rank(int k, int r) {
  return 2 * SH::H + (SH::P - 1) * SH::C + (k - 1) * SH::R + r;
}

template <class SH>
__device__ __forceinline__ int min_firework(const Game<SH>& g) {
  int m = g.firework(0);
#pragma unroll
  for (int c = 1; c < SH::C; ++c) m = min(m, g.firework(c));
  return m;
}

// grid of the cards playable right now: bit c*5 + fireworks[c]   (playable_card :14-33)
template <class SH>
__device__ __forceinline__ uint32_t playable_grid(const Game<SH>& g) {
  uint32_t pg = 0;
#pragma unroll
  for (int c = 0; c < SH::C; ++c)
    if (g.firework(c) < SH::R) pg |= 1u << (c * SH::R + g.firework(c));
  return pg;
}

template <class SH>
__device__ __forceinline__ constexpr unsigned long long instances_packed() {
  unsigned long long k = 0;   // 2-bit fields 3,2,2,2,1 per colour = num_in_deck_by_rank (ruleset.py:9)
  for (int c = 0; c < SH::C; ++c)
    k |= (3ull | (2ull << 2) | (2ull << 4) | (2ull << 6) | (1ull << 8)) << (10 * c);
  return k;
}

// get_max_fireworks  ruleset.py:174-193: 5, or the lowest rank whose every copy is discarded.
// Returned as the grid of cards still worth keeping: bit c*5+r set iff r < max_fireworks[c];
// *mf_out gets the packed 3-bit values.
template <class SH>
__device__ __forceinline__ uint32_t useful_grid(const Game<SH>& g, uint32_t* mf_out) {
  const unsigned long long x = g.disc ^ instances_packed<SH>();
  const unsigned long long gone = ~(x | (x >> 1)) & 0x5555555555555555ull;   // count == in deck
  uint32_t grid = 0, mf = 0;
#pragma unroll
  for (int c = 0; c < SH::C; ++c) {
    const uint32_t bits = (uint32_t)(gone >> (10 * c)) & 0x155u;
    const int m = bits ? (__ffs(bits) - 1) >> 1 : 5;
    mf |= (uint32_t)m << (3 * c);
    grid |= ((1u << m) - 1) << (5 * c);
  }
  *mf_out = mf;
  return grid;
}

// 25-bit grid (bit 5c+r) -> 2-bit-field positions (bit 10c+2r)
__device__ __forceinline__ unsigned long long spread_grid(uint32_t grid) {
  unsigned long long x = grid;
  x = (x | (x << 16)) & 0x0000ffff0000ffffull;
  x = (x | (x << 8)) & 0x00ff00ff00ff00ffull;
  x = (x | (x << 4)) & 0x0f0f0f0f0f0f0f0full;
  x = (x | (x << 2)) & 0x3333333333333333ull;
  x = (x | (x << 1)) & 0x5555555555555555ull;
  return x;
}

// sum of the 2-bit fields of x selected by sel (sel has bits at even positions)
__device__ __forceinline__ int sum_fields(unsigned long long x, unsigned long long sel) {
  return __popcll(x & sel) + 2 * __popcll((x >> 1) & sel);
}

// play_probably_safe_factory  ruleset.py:617-635 with get_card_playability :82-125.
// playability = playable copies / plausible copies among the cards not visible to me; the
// first-maximum (np.argmax) is taken by exact cross-multiplication, the threshold test is the
// reference's float64 division.
template <class SH>
__device__ __forceinline__ int play_probably_safe(const Game<SH>& g, double threshold,
                                                  bool require_extra_lives, const RuleSet& rs, int slot) {
  const int n = g.hand_n(0);
  if (n == 0) return -1;
  // visible copies per card type: other hands + discards + fireworks  (:60-76)
  unsigned long long vis = g.disc;
#pragma unroll
  for (int k = 1; k < SH::P; ++k)
#pragma unroll
    for (int i = 0; i < SH::H; ++i)
      if (slot_valid(g.hand[k][i])) vis += 1ull << (2 * slot_card(g.hand[k][i]));
#pragma unroll
  for (int c = 0; c < SH::C; ++c)   // ranks below the stack height: one visible copy each
    vis += (0x155ull & ((1ull << (2 * g.firework(c))) - 1)) << (10 * c);
  const unsigned long long left = instances_packed<SH>() - vis;   // no borrows: visible <= instances
  const unsigned long long pg_s = spread_grid(playable_grid(g));
  int best = 0, bp = 0, bt = 1;
#pragma unroll
  for (int i = 0; i < SH::H; ++i) {
    if (i < n) {
      const unsigned long long gs = spread_grid(slot_grid(g.hand[0][i]));
      const int total = sum_fields(left, gs);
      const int playable = sum_fields(left, gs & pg_s);
      // p_i/t_i > p_b/t_b  <=>  p_i*t_b > p_b*t_i   (t > 0); first maximum wins
      if (i == 0 || playable * bt > bp * total) { bp = playable; bt = total; best = i; }
    }
  }
  if (!require_extra_lives || g.life > 1)
    if (ratio_at_least(rs, slot, bp, bt, threshold)) return uid_play<SH>(best);   // :105, :622
  return -1;
}

// grid of the cards that can no longer score: rank below the stack or at / beyond max_fireworks
// (useless_card  ruleset.py:35-42)
template <class SH>
__device__ __forceinline__ uint32_t useless_grid(const Game<SH>& g) {
  uint32_t mf;
  const uint32_t useful = useful_grid(g, &mf);
  uint32_t below = 0;
#pragma unroll
  for (int c = 0; c < SH::C; ++c) below |= ((1u << g.firework(c)) - 1) << (5 * c);
  return (~useful | below) & SH::kGridAll;
}

// discard_probably_useless_factory  ruleset.py:638-649 with get_probability_useless :127-170
// (intended semantics): same bookkeeping as play_probably_safe with the useless grid.
template <class SH>
__device__ __forceinline__ int discard_probably_useless(const Game<SH>& g, double threshold,
                                                        const RuleSet& rs, int slot) {
  const int n = g.hand_n(0);
  if (!(g.info < 8) || n == 0) return -1;
  unsigned long long vis = g.disc;
#pragma unroll
  for (int k = 1; k < SH::P; ++k)
#pragma unroll
    for (int i = 0; i < SH::H; ++i)
      if (slot_valid(g.hand[k][i])) vis += 1ull << (2 * slot_card(g.hand[k][i]));
#pragma unroll
  for (int c = 0; c < SH::C; ++c)
    vis += (0x155ull & ((1ull << (2 * g.firework(c))) - 1)) << (10 * c);
  const unsigned long long left = instances_packed<SH>() - vis;
  const unsigned long long ug_s = spread_grid(useless_grid(g));
  int best = 0, bu = 0, bt = 1;
#pragma unroll
  for (int i = 0; i < SH::H; ++i) {
    if (i < n) {
      const unsigned long long gs = spread_grid(slot_grid(g.hand[0][i]));
      const int total = sum_fields(left, gs);
      const int useless = sum_fields(left, gs & ug_s);
      if (i == 0 || useless * bt > bu * total) { bu = useless; bt = total; best = i; }
    }
  }
  if (ratio_at_least(rs, slot, bu, bt, threshold)) return uid_discard<SH>(best);
  return -1;
}

// tell_playable_card_outer  ruleset.py:413-451
template <class SH>
__device__ __forceinline__ int tell_playable_outer(const Game<SH>& g) {
  if (g.info <= 0) return -1;
  int mv = -1;
#pragma unroll
  for (int k = SH::P - 1; k >= 1; --k)          // reverse order, last write = first match
#pragma unroll
    for (int i = SH::H - 1; i >= 0; --i) {
      const uint32_t w = g.hand[k][i];
      const int card = slot_card(w), c = card_color(card), r = card_rank(card);
      if (slot_valid(w) && r == g.firework(c)) {
        if (!slot_rhint(w)) mv = uid_rank<SH>(k, r);
        else if (!slot_chint(w)) mv = uid_color<SH>(k, c);
      }
    }
  return mv;
}

// The three rules that are dead in the reference (ids >= R_TELL_ANYONE_USELESS_CARD).  Kept out of
// run_rule's switch and compiled only into the EXT instantiations of the kernels: with them in
// the same switch the fused self-play kernel of the headline benchmark lost 11 % (register
// allocation of the hot cases), although a piers agent never executes them.
template <class SH>
__device__ __noinline__ int run_ext_rule(const Game<SH>& g, int rule, double farg, const RuleSet& rs,
                                         int slot, unsigned long long seed, unsigned long long game_id,
                                         unsigned long long step) {
  constexpr int H = SH::H, P = SH::P;
  switch (rule) {
    case R_TELL_ANYONE_USELESS_CARD: {                 // :522-535 (intended semantics)
      if (!(g.info > 1)) return -1;
      const uint32_t ug = useless_grid(g);
      int mv = -1;
#pragma unroll
      for (int k = P - 1; k >= 1; --k)
#pragma unroll
        for (int i = H - 1; i >= 0; --i) {
          const uint32_t w = g.hand[k][i];
          const int card = slot_card(w);
          if (slot_valid(w) && ((ug >> card) & 1u)) {
            if (!slot_chint(w)) mv = uid_color<SH>(k, card_color(card));
            else if (!slot_rhint(w)) mv = uid_rank<SH>(k, card_rank(card));
          }
        }
      return mv;
    }

    case R_TELL_PLAYABLE_CARD: {                       // :570-594 (intended semantics)
      if (!(g.info > 0)) return -1;
      int fk = -1, fc = 0;
#pragma unroll
      for (int k = P - 1; k >= 1; --k)
#pragma unroll
        for (int i = H - 1; i >= 0; --i) {
          const uint32_t w = g.hand[k][i];
          const int card = slot_card(w);
          if (slot_valid(w) && card_rank(card) == g.firework(card_color(card))) { fk = k; fc = card; }
        }
      if (fk < 0) return -1;
      const int coin = rand_below(rule_rand(seed, game_id, step, 4), 2);   // random.randint(0, 1) :584
      return coin == 0 ? uid_rank<SH>(fk, card_rank(fc)) : uid_color<SH>(fk, card_color(fc));
    }

    case R_DISCARD_PROBABLY_USELESS:                   // :638-649
      return discard_probably_useless(g, farg, rs, slot);

    default:
      return -1;
  }
}

template <class SH, bool EXT = false>
__device__ __forceinline__ int run_rule(const Game<SH>& g, const RuleSet& rs, int k,
                                        unsigned long long game_id, unsigned long long step) {
  constexpr int C = SH::C, R = SH::R, H = SH::H, P = SH::P;
  const int rule = rs.id[k], iarg = rs.iarg[k], slot = rs.tslot[k];
  const double farg = rs.farg[k];
  const unsigned long long seed = rs.seed;
  if constexpr (EXT) {
    if (rule >= R_TELL_ANYONE_USELESS_CARD) return run_ext_rule(g, rule, farg, rs, slot, seed, game_id, step);
  }
  switch (rule) {
    case R_DISCARD_OLDEST_FIRST:                       // ruleset.py:206-216
      return g.info < 8 ? uid_discard<SH>(0) : -1;

    case R_OSAWA_DISCARD: {                            // :220-280
      if (g.info == 8) return -1;
      uint32_t mf;
      const uint32_t useful = useful_grid(g, &mf);
      const int minfw = min_firework(g);
      int first = -1, second = -1;
#pragma unroll
      for (int i = H - 1; i >= 0; --i) {
        const uint32_t w = g.hand[0][i];
        if (!slot_valid(w)) continue;
        const int card = slot_card(w), hc = card_color(card), hr = card_rank(card);
        const bool ch = slot_chint(w), rh = slot_rhint(w);
        const int cidx = ch ? hc : C - 1;              // fireworks[-1] quirk  :229-230
        bool a = g.firework(cidx) == 5;
        if (ch && rh) a = a || hr < g.firework(hc) || hr >= (int)((mf >> (3 * hc)) & 7);
        if (rh) a = a || hr < minfw;
        if (a) first = i;
        if ((slot_grid(w) & useful) == 0) second = i;  // no plausible identity is still useful
      }
      if (first >= 0) return uid_discard<SH>(first);
      return second >= 0 ? uid_discard<SH>(second) : -1;
    }

    case R_TELL_UNKNOWN: {                             // :285-309 (next player, colour first)
      if (g.info <= 0) return -1;
      int mv = -1;
#pragma unroll
      for (int i = H - 1; i >= 0; --i) {
        const uint32_t w = g.hand[1][i];
        const int card = slot_card(w);
        if (slot_valid(w)) {
          if (!slot_chint(w)) mv = uid_color<SH>(1, card_color(card));
          else if (!slot_rhint(w)) mv = uid_rank<SH>(1, card_rank(card));
        }
      }
      return mv;
    }

    case R_TELL_RANDOMLY: {                            // :314-347
      if (g.info <= 0) return -1;
      const int n = g.hand_n(1);
      if (n == 0) return -1;
      const int idx = rand_below(rule_rand(seed, game_id, step, 0), n);
      const int coin = rand_below(rule_rand(seed, game_id, step, 1), 2);
      uint32_t w = 0;
#pragma unroll
      for (int i = 0; i < H; ++i)
        if (i == idx) w = g.hand[1][i];
      const int card = slot_card(w);
      return coin == 0 ? uid_rank<SH>(1, card_rank(card)) : uid_color<SH>(1, card_color(card));
    }

    case R_PLAY_SAFE_CARD: {                           // :350-380: every plausible identity playable
      const uint32_t pg = playable_grid(g);
      int mv = -1;
#pragma unroll
      for (int i = H - 1; i >= 0; --i) {
        const uint32_t w = g.hand[0][i];
        if (slot_valid(w) && (slot_grid(w) & ~pg) == 0) mv = uid_play<SH>(i);
      }
      return mv;
    }

    case R_PLAY_IF_CERTAIN: {                          // :383-409
      int mv = -1;
#pragma unroll
      for (int i = H - 1; i >= 0; --i) {
        const uint32_t w = g.hand[0][i];
        const int card = slot_card(w);
        if (slot_valid(w) && slot_chint(w) && slot_rhint(w) &&
            card_rank(card) == g.firework(card_color(card)))
          mv = uid_play<SH>(i);
      }
      return mv;
    }

    case R_TELL_PLAYABLE_CARD_OUTER:                   // :413-451
    case R_TELL_ANYONE_USEFUL_CARD:                    // :518-519
      return tell_playable_outer(g);

    case R_TELL_DISPENSABLE: {                         // :454-514, iarg = min_information_tokens
      if (!(g.info < iarg) || g.info <= 0) return -1;
      const int minfw = min_firework(g);
      int mv = -1;
#pragma unroll
      for (int k = P - 1; k >= 1; --k)
#pragma unroll
        for (int i = H - 1; i >= 0; --i) {
          const uint32_t w = g.hand[k][i];
          if (!slot_valid(w)) continue;
          const int card = slot_card(w), c = card_color(card), r = card_rank(card);
          const bool ch = slot_chint(w), rh = slot_rhint(w);
          const int f = g.firework(c);
          if (!ch && f == 5) mv = uid_color<SH>(k, c);
          else if (!rh && r < minfw) mv = uid_rank<SH>(k, r);
          else if (r < f && !ch && rh) mv = uid_color<SH>(k, c);
          else if (r < f && ch && !rh) mv = uid_rank<SH>(k, r);
        }
      return mv;
    }

    case R_PLAY_PROBABLY_SAFE:                         // :617-635
    case R_HAIL_MARY: {                                // :653-655 = play_probably_safe(0.0) when the deck is
      const bool hm = rule == R_HAIL_MARY;             // empty and a life can be spared; ONE inlined copy
      if (hm && !(SH::DECK - g.deck_pos == 0 && g.life > 1)) return -1;
      return play_probably_safe(g, hm ? 0.0 : farg, hm ? false : iarg != 0, rs, hm ? -1 : slot);
    }

    case R_LEGAL_RANDOM: {                             // :598-604 (intended semantics)
      unsigned long long m = legal_mask(g);
      const int n = __popcll(m);
      if (n == 0) return -1;
      int j = rand_below(rule_rand(seed, game_id, step, 2), n);
      while (j-- > 0) m &= m - 1;
      return __ffsll((long long)m) - 1;
    }

    case R_DISCARD_RANDOMLY: {                         // :607-614 (intended semantics)
      const int n = g.hand_n(0);
      if (!(g.info < 8) || n == 0) return -1;
      return uid_discard<SH>(rand_below(rule_rand(seed, game_id, step, 3), n));
    }

    case R_TELL_MOST_INFORMATION:                      // :539-562 always None (:562)
    default:
      return -1;
  }
}

// RulebasedAgent.get_move  rule_based.py:13-25.  *hit = index of the rule that fired
// (rs.n = the legal_random fallback, stored by the host as entry [n]).
template <class SH, bool EXT = false>
__device__ __forceinline__ int rules_act(const Game<SH>& g, const RuleSet& rs,
                                         unsigned long long game_id, unsigned long long step,
                                         int* hit) {
  int a = -1, h = rs.n;
#pragma unroll 1
  for (int k = 0; k <= rs.n; ++k) {
    if (a < 0) {
      const int r = run_rule<SH, EXT>(g, rs, k, game_id, step);
      if (r >= 0) { a = r; h = k; }
    }
  }
  *hit = h;
  return a < 0 ? 0 : a;
}

}  // namespace hbg
