// Rule-based agents, evaluated warp-cooperatively on the packed game state.
//
// Mirrors hanabi_agents/rule_based/ruleset.py (rules) and rule_based.py:13-25 (first rule that
// returns a move wins, otherwise Ruleset.legal_random) of the reference.  Each rule is "the
// first card, in (player offset, hand index) order, that satisfies a predicate": every slot
// lane evaluates the predicate for its own card and a warp min-reduction picks the first.
//
// Documented divergences (DESIGN.md "rule-based agents"): python `random` draws are a
// counter-based stream keyed (seed, game id, step, draw); legal_random / discard_randomly use
// their intended semantics (the reference versions raise TypeError, ruleset.py:598-614);
// tell_unknown hints the real card's colour/rank (the reference builds an illegal move from
// the knowledge object, :291-307); osawa_discard keeps the fireworks[-1] quirk (:229-230).
#pragma once
#include "hanabi_game.cuh"

namespace hbg {

enum RuleId {
  R_DISCARD_OLDEST_FIRST = 0, R_OSAWA_DISCARD = 1, R_TELL_UNKNOWN = 2, R_TELL_RANDOMLY = 3,
  R_PLAY_SAFE_CARD = 4, R_PLAY_IF_CERTAIN = 5, R_TELL_PLAYABLE_CARD_OUTER = 6,
  R_TELL_DISPENSABLE = 7, R_TELL_ANYONE_USEFUL_CARD = 8, R_PLAY_PROBABLY_SAFE = 9,
  R_HAIL_MARY = 10, R_LEGAL_RANDOM = 11, R_DISCARD_RANDOMLY = 12, R_TELL_MOST_INFORMATION = 13,
  R_NUM_RULES = 14
};

constexpr int kMaxRules = 16;
struct RuleSet {
  int n;
  int id[kMaxRules];
  int iarg[kMaxRules];
  double farg[kMaxRules];
  unsigned long long seed;
};

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

// Per-lane view of the observer (= player to move) plus the warp-uniform quantities several
// rules share.  2-bit-per-card-type packing ("spread" form, type t at bit 2t) lets the
// playability sums run as popcounts.
template <class SH>
struct View {
  bool valid;       // this lane holds a card
  int k, i;         // owner's offset from the observer, index in hand
  int c, r;         // the real card (meaningful for k >= 1)
  int cp, rp;       // plausible colour / rank masks
  bool ch, rh;      // colour / rank explicitly hinted
  uint32_t key;     // order of iteration: (k, i)
  unsigned long long grid_s;   // plausible(colour) x plausible(rank), spread form
  unsigned long long pg_s;     // uniform: cards playable right now, spread form
  int minfw;                   // uniform: min(fireworks)
};

// 5-bit mask -> bits at 0,2,4,6,8
__device__ __forceinline__ uint32_t spread5(uint32_t x) {
  return (x & 1u) | ((x & 2u) << 1) | ((x & 4u) << 2) | ((x & 8u) << 3) | ((x & 16u) << 4);
}

template <class SH>
__device__ __forceinline__ View<SH> make_view(const Game<SH>& g) {
  View<SH> v;
  const int lane = lane_id();
  const int p = lane / SH::H;
  v.valid = lane < SH::S && slot_valid(g.slot);
  v.i = lane - p * SH::H;
  v.k = p - g.cur;
  if (v.k < 0) v.k += SH::P;
  const int card = slot_card(g.slot);
  v.c = card / SH::R;
  v.r = card - v.c * SH::R;
  v.cp = slot_cplaus(g.slot);
  v.rp = slot_rplaus(g.slot);
  v.ch = slot_chint(g.slot);
  v.rh = slot_rhint(g.slot);
  v.key = (uint32_t)(v.k * SH::H + v.i);
  const unsigned long long rs = spread5((uint32_t)v.rp);
  v.grid_s = 0;
  v.pg_s = 0;
  v.minfw = g.firework(0);
#pragma unroll
  for (int c = 0; c < SH::C; ++c) {
    if ((v.cp >> c) & 1) v.grid_s |= rs << (10 * c);
    const int f = g.firework(c);
    if (f < SH::R) v.pg_s |= 1ull << (10 * c + 2 * f);
    v.minfw = min(v.minfw, f);
  }
  return v;
}

// lane of the first card (in (k, i) order) whose predicate holds, or -1
template <class SH>
__device__ __forceinline__ int first_lane(const View<SH>& v, bool pred) {
  const uint32_t x = pred ? ((v.key << 5) | (uint32_t)lane_id()) : 0xffffffffu;
  const uint32_t m = __reduce_min_sync(kFull, x);
  return m == 0xffffffffu ? -1 : (int)(m & 31);
}

template <class SH> __device__ __forceinline__ int uid_discard(int i) { return i; }
template <class SH> __device__ __forceinline__ int uid_play(int i) { return SH::H + i; }
template <class SH> __device__ __forceinline__ int uid_color(int k, int c) {
  return 2 * SH::H + (k - 1) * SH::C + c;
}
template <class SH> __device__ __forceinline__ int uid_rank(int k, int r) {
  return 2 * SH::H + (SH::P - 1) * SH::C + (k - 1) * SH::R + r;
}

// 2-bit fields holding num_in_deck_by_rank = [3,2,2,2,1]  (ruleset.py:9)
template <class SH>
__device__ __forceinline__ constexpr unsigned long long in_deck_packed() {
  unsigned long long k = 0;
  for (int c = 0; c < SH::C; ++c)
    k |= (3ull | (2ull << 2) | (2ull << 4) | (2ull << 6) | (1ull << 8)) << (10 * c);
  return k;
}

// get_max_fireworks  ruleset.py:174-193: 5, or the lowest rank whose every copy is discarded.
// 3 bits per colour.
template <class SH>
__device__ __forceinline__ uint32_t max_fireworks(const Game<SH>& g) {
  const unsigned long long x = g.disc ^ in_deck_packed<SH>();
  const unsigned long long gone = ~(x | (x >> 1)) & 0x5555555555555555ull;   // count == in deck
  uint32_t out = 0;
#pragma unroll
  for (int c = 0; c < SH::C; ++c) {
    const uint32_t bits = (uint32_t)(gone >> (10 * c)) & 0x155u;
    const int m = bits ? (__ffs(bits) - 1) >> 1 : 5;
    out |= (uint32_t)m << (3 * c);
  }
  return out;
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
__device__ __forceinline__ int play_probably_safe(const Game<SH>& g, const View<SH>& v,
                                                  double threshold, bool require_extra_lives) {
  const int n = g.hand_n(g.cur);
  if (n == 0) return -1;
  // visible copies per card type: other hands + discards + fireworks  (:60-76)
  unsigned long long vis = (v.valid && v.k >= 1) ? (1ull << (2 * (v.c * SH::R + v.r))) : 0ull;
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) vis += __shfl_xor_sync(kFull, vis, o);
  vis += g.disc;
#pragma unroll
  for (int c = 0; c < SH::C; ++c)   // ranks below the stack height: one visible copy each
    vis += (0x155ull & ((1ull << (2 * g.firework(c))) - 1)) << (10 * c);
  unsigned long long inst = 0;
#pragma unroll
  for (int c = 0; c < SH::C; ++c)
    inst |= (3ull | (2ull << 2) | (2ull << 4) | (2ull << 6) | (1ull << 8)) << (10 * c);
  const unsigned long long left = inst - vis;            // no borrows: visible <= instances
  const int total = sum_fields(left, v.grid_s);
  const int playable = sum_fields(left, v.grid_s & v.pg_s);
  const uint32_t mine = (uint32_t)playable | ((uint32_t)total << 8);
  int best = 0;
  uint32_t bp = __shfl_sync(kFull, mine, g.cur * SH::H);
  for (int i = 1; i < n; ++i) {
    const uint32_t x = __shfl_sync(kFull, mine, g.cur * SH::H + i);
    // p_i/t_i > p_b/t_b  <=>  p_i*t_b > p_b*t_i   (t > 0)
    if ((x & 255u) * (bp >> 8) > (bp & 255u) * (x >> 8)) { bp = x; best = i; }
  }
  if (!require_extra_lives || g.life > 1)
    if ((double)(bp & 255u) / (double)(bp >> 8) >= threshold) return uid_play<SH>(best);   // :105,622
  return -1;
}
template <class SH>
__device__ __forceinline__ int run_rule(const Game<SH>& g, const Rt& rt, const View<SH>& v, int rule,
                                        double farg, int iarg, unsigned long long seed,
                                        unsigned long long game_id, unsigned long long step) {
  constexpr int C = SH::C, R = SH::R, H = SH::H, P = SH::P;
  const bool mine = v.valid && v.k == 0;
  const bool theirs = v.valid && v.k >= 1;
  switch (rule) {
    case R_DISCARD_OLDEST_FIRST:                       // ruleset.py:206-216
      return g.info < 8 ? uid_discard<SH>(0) : -1;

    case R_OSAWA_DISCARD: {                            // :220-280
      if (g.info == 8) return -1;
      const uint32_t mf = max_fireworks(g);
      const int minfw = v.minfw;
      const int hc = __ffs(v.cp) - 1, hr = __ffs(v.rp) - 1;   // hinted value = the single bit
      const int cidx = v.ch ? hc : C - 1;                     // fireworks[-1] quirk  :229-230
      bool a = g.firework(cidx) == 5;
      if (v.ch && v.rh) a = a || hr < g.firework(hc) || hr >= (int)((mf >> (3 * hc)) & 7);
      if (v.rh) a = a || hr < minfw;
      int l = first_lane(v, mine && a);
      if (l >= 0) return uid_discard<SH>(l - g.cur * H);
      bool eventually = false;                                 // :262-279
#pragma unroll
      for (int c = 0; c < C; ++c)
        if ((v.cp >> c) & 1) {
          const int lim = (mf >> (3 * c)) & 7;                 // ranks r < lim still useful
          if (v.rp & ((1 << lim) - 1)) eventually = true;
        }
      l = first_lane(v, mine && !eventually);
      return l >= 0 ? uid_discard<SH>(l - g.cur * H) : -1;
    }

    case R_TELL_UNKNOWN: {                             // :285-309 (next player, colour first)
      if (g.info <= 0) return -1;
      const bool mine1 = v.valid && v.k == 1;
      const int l = first_lane(v, mine1 && (!v.ch || !v.rh));
      if (l < 0) return -1;
      const int mv = !v.ch ? uid_color<SH>(1, v.c) : uid_rank<SH>(1, v.r);
      return __shfl_sync(kFull, mv, l);
    }

    case R_TELL_RANDOMLY: {                            // :314-347
      if (g.info <= 0) return -1;
      int t = g.cur + 1;
      if (t >= P) t -= P;
      const int n = g.hand_n(t);
      if (n == 0) return -1;
      const int i = rand_below(rule_rand(seed, game_id, step, 0), n);
      const int coin = rand_below(rule_rand(seed, game_id, step, 1), 2);
      const int mv = coin == 0 ? uid_rank<SH>(1, v.r) : uid_color<SH>(1, v.c);
      return __shfl_sync(kFull, mv, t * H + i);
    }

    case R_PLAY_SAFE_CARD: {                           // :350-380
      const bool safe = (v.grid_s & ~v.pg_s) == 0;
      const int l = first_lane(v, mine && safe);
      return l >= 0 ? uid_play<SH>(l - g.cur * H) : -1;
    }

    case R_PLAY_IF_CERTAIN: {                          // :383-409
      const int hc = __ffs(v.cp) - 1, hr = __ffs(v.rp) - 1;
      const bool ok = v.ch && v.rh && hr == g.firework(v.ch ? hc : 0);
      const int l = first_lane(v, mine && ok);
      return l >= 0 ? uid_play<SH>(l - g.cur * H) : -1;
    }

    case R_TELL_PLAYABLE_CARD_OUTER:                   // :413-451
    case R_TELL_ANYONE_USEFUL_CARD: {                  // :518-519
      if (g.info <= 0) return -1;
      const bool playable = v.r == g.firework(v.c);
      const int l = first_lane(v, theirs && playable && (!v.rh || !v.ch));
      if (l < 0) return -1;
      const int mv = !v.rh ? uid_rank<SH>(v.k, v.r) : uid_color<SH>(v.k, v.c);
      return __shfl_sync(kFull, mv, l);
    }

    case R_TELL_DISPENSABLE: {                         // :454-514, iarg = min_information_tokens
      if (!(g.info < iarg) || g.info <= 0) return -1;
      const int minfw = v.minfw;
      const int f = g.firework(v.c);
      int mv = -1;
      if (!v.ch && f == 5) mv = uid_color<SH>(v.k, v.c);
      else if (!v.rh && v.r < minfw) mv = uid_rank<SH>(v.k, v.r);
      else if (v.r < f && !v.ch && v.rh) mv = uid_color<SH>(v.k, v.c);
      else if (v.r < f && v.ch && !v.rh) mv = uid_rank<SH>(v.k, v.r);
      const int l = first_lane(v, theirs && mv >= 0);
      return l >= 0 ? __shfl_sync(kFull, mv, l) : -1;
    }

    case R_PLAY_PROBABLY_SAFE:                         // :617-635
      return play_probably_safe(g, v, farg, iarg != 0);

    case R_HAIL_MARY:                                  // :653-655
      if (SH::DECK - g.deck_pos == 0 && g.life > 1) return play_probably_safe(g, v, 0.0, false);
      return -1;

    case R_LEGAL_RANDOM: {                             // :598-604 (intended semantics)
      unsigned long long m = legal_mask(g, rt, g.cur);
      const int n = __popcll(m);
      if (n == 0) return -1;
      int j = rand_below(rule_rand(seed, game_id, step, 2), n);
      while (j-- > 0) m &= m - 1;
      return __ffsll((long long)m) - 1;
    }

    case R_DISCARD_RANDOMLY: {                         // :607-614 (intended semantics)
      const int n = g.hand_n(g.cur);
      if (!(g.info < 8) || n == 0) return -1;
      return uid_discard<SH>(rand_below(rule_rand(seed, game_id, step, 3), n));
    }

    case R_TELL_MOST_INFORMATION:                      // :539-562 always None (:562)
    default:
      return -1;
  }
}

// RulebasedAgent.get_move  rule_based.py:13-25.  *hit = index of the rule that fired
// (rules.n = the legal_random fallback).
template <class SH>
__device__ __forceinline__ int rules_act(const Game<SH>& g, const Rt& rt, const RuleSet& rs,
                                         unsigned long long game_id, unsigned long long step,
                                         int* hit) {
  const View<SH> v = make_view(g);
  int a = -1, k = 0;
#pragma unroll 1
  for (; k <= rs.n; ++k) {   // k == rs.n: Ruleset.legal_random fallback (rule_based.py:24)
    const bool fb = k == rs.n;
    a = run_rule(g, rt, v, fb ? (int)R_LEGAL_RANDOM : rs.id[fb ? 0 : k], fb ? 0.0 : rs.farg[fb ? 0 : k],
                 fb ? 0 : rs.iarg[fb ? 0 : k], rs.seed, game_id, step);
    if (a >= 0 || fb) break;
  }
  *hit = k;
  return a < 0 ? 0 : a;
}

}  // namespace hbg
