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
// bit of RuleSet::used (rule ids end at 16): osawa_discard with its iarg != 0, i.e. the INTENDED reading of
// ruleset.py:229-230 (colour test only on colour-hinted cards) instead of the fireworks[-1] quirk
constexpr unsigned int kOsawaIntended = 1u << 31;
constexpr int kThrSlots = 4;      // rules with a probability threshold that get a comparison table
constexpr int kThrRows = 52;      // denominators 0..51 (a deck holds 50 cards)
struct RuleSet {        // entry [n] is the legal_random fallback, filled by the host
  int n;
  unsigned int used;    // bit r: rule id r occurs in the list (its core is evaluated)
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
  // How rules_resolve finds the first rule that fires without walking the list: every rule whose
  // verdict is a plain function of the state owns a ROW; key_hi[row] = (position in the list) << 8, or
  // kKeyAbsent when the list does not hold it.  Entries of the `lazy` mask (random rules, rules that
  // are dead in the reference, anything beyond the rows) are evaluated only when a game gets there.
  unsigned int key_hi[16];
  int td_ilt[2];                  // tell_dispensable rows: fires when info < td_ilt
  unsigned int lazy;
  int n_slots;                    // comparison tables in use
  int slot_extra[kThrSlots];      // require_extra_lives of the rule that owns the table
};

enum RuleRow {
  ROW_DISCARD_OLDEST = 0, ROW_OSAWA = 1, ROW_TELL_UNKNOWN = 2, ROW_PLAY_SAFE = 3, ROW_PLAY_CERTAIN = 4,
  ROW_TELL_PLAYABLE = 5, ROW_TELL_DISPENSABLE0 = 6, ROW_TELL_DISPENSABLE1 = 7, ROW_HAIL_MARY = 8,
  ROW_TELL_RANDOMLY = 9, ROW_DISCARD_RANDOMLY = 10,
  ROW_PPS0 = 11, ROW_NUM = ROW_PPS0 + kThrSlots
};
static_assert(ROW_NUM <= 16, "RuleSet::key_hi");
constexpr unsigned int kKeyNone = 0x10000u;      // "no move" sorts after every position
constexpr unsigned int kKeyAbsent = 0x40000u;
constexpr int kRuleBlock = 64;    // threads per CTA of every kernel that resolves rule lists

// numerator / denominator >= threshold, exactly as the reference's float64 division decides it.
// The division is kept out of line: it is reached only by a rule list with more threshold rules than
// comparison tables, and inlined its ~40 instructions were walked (predicated off) by every game that
// evaluated a lazy rule.
__device__ __noinline__ bool ratio_by_division(int num, int den, double threshold) {
  return (double)num / (double)den >= threshold;
}
__device__ __forceinline__ bool ratio_at_least(const RuleSet& rs, int slot, int num, int den,
                                               double threshold) {
  if (threshold <= 0.0) return true;                       // num >= 0, den > 0
  if (slot >= 0 && den < kThrRows && num < 64) return (rs.thr[slot][den] >> num) & 1ull;
  return ratio_by_division(num, den, threshold);
}

// Counter-based uniform stream keyed (seed, game id, step, draw): a 64-bit key per (seed, game) -- loop
// invariant in the fused kernel -- and ONE 32-bit multiply-xorshift hash per (game, step).  At most one
// random rule decides a game's move in a step, so only the draws of one rule have to be independent of
// each other: index draws (0, 2, 3) read the hash from the top, coin draws (1, 4) from bit 15 down.
// ptxas evaluates the draws of the "lazy" rules ahead of the branches for every game; as five 64-bit
// splitmix hashes they were 143 of 1192 warp-instructions per step (profiles/r01_selfplay_history.md, v6).
// (The CPU restatement used by the tests defines the same stream.)
__device__ __forceinline__ unsigned long long rule_key(unsigned long long seed, unsigned long long game) {
  unsigned long long x = seed + 0x9E3779B97F4A7C15ull * (game + 1);
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return x;
}
__device__ __forceinline__ uint32_t rule_rand(unsigned long long seed, unsigned long long game,
                                              unsigned long long step, uint32_t k) {
  const unsigned long long key = rule_key(seed, game);
  uint32_t h = (uint32_t)key + (uint32_t)step * 0x9E3779B1u;
  h ^= h >> 16; h *= 0x7FEB352Du;
  h ^= h >> 15; h ^= (uint32_t)(key >> 32); h *= 0x846CA68Bu;
  h ^= h >> 16;
  return (k == 1 || k == 4) ? h << 16 : h;
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

// ---- grids derived from the board once per step -------------------------------------------------
// rows of the finished colours (stack height 5)
template <class SH>
__device__ __forceinline__ uint32_t done_rows(const Game<SH>& g) { return ((g.fw >> 4) & kLsb5) * 31u; }

// cards whose rank is below EVERY stack (rank < min(fireworks)), in every colour row
template <class SH>
__device__ __forceinline__ uint32_t below_all_grid(const Game<SH>& g) {
  uint32_t a = g.fw;
#pragma unroll
  for (int c = 1; c < SH::C; ++c) a &= g.fw >> (5 * c);
  return (a & 31u) * kLsb5;
}

// get_max_fireworks  ruleset.py:174-193: 5, or the lowest rank whose every copy is discarded.
// Returned as the grid of cards still worth keeping: bit c*5+r set iff r < max_fireworks[c].
template <class SH>
__device__ __forceinline__ uint32_t useful_grid(const Game<SH>& g) {
  uint32_t p = ~((g.d0 ^ kInst0) | (g.d1 ^ kInst1)) & SH::kGridAll;   // count == copies in the deck
  p |= (p << 1) & ~kLsb5;                                             // prefix-OR inside each colour
  p |= (p << 2) & ~(kLsb5 | (kLsb5 << 1));
  p |= (p << 1) & ~kLsb5;
  return ~p & SH::kGridAll;
}

// grid of the cards that can no longer score: rank below the stack or at / beyond max_fireworks
// (useless_card  ruleset.py:35-42)
template <class SH>
__device__ __forceinline__ uint32_t useless_grid(const Game<SH>& g) {
  return (~useful_grid(g) | g.fw) & SH::kGridAll;
}

// copies of every card type that I cannot see (deck + my own hand), as two bit planes:
// copies in the game - discards - other hands - cards on the stacks   (get_visible_cards :60-76)
template <class SH>
__device__ __forceinline__ void unseen_planes(const Game<SH>& g, uint32_t& l0, uint32_t& l1) {
  uint32_t v0 = g.d0 ^ g.fw, v1 = g.d1 ^ (g.d0 & g.fw);      // one visible copy per rank below a stack
#pragma unroll
  for (int k = 1; k < SH::P; ++k)
#pragma unroll
    for (int i = 0; i < SH::H; ++i) {
      const uint32_t oh = g.oh[k][i];
      v1 ^= v0 & oh;
      v0 ^= oh;
    }
  l0 = kInst0 ^ v0;                                          // copies - visible, no borrow out of a field
  l1 = kInst1 ^ v1 ^ (~kInst0 & v0);
}

__device__ __forceinline__ int count_planes(uint32_t sel, uint32_t l0, uint32_t l1) {
  return __popc(sel & l0) + 2 * __popc(sel & l1);
}

// What each rule would do, gates that depend on the rule's arguments left out.  A warp holds games
// in every situation, so it walks (nearly) every rule every step anyway: the bodies are evaluated
// once, side by side, as straight-line code that shares the board grids and the card one-hots,
// and the rule list is resolved afterwards (rules_resolve).
struct Cores {
  int play_safe, play_certain, tell_playable, tell_dispensable, osawa, tell_unknown;
  int best, bp, bt;     // play_probably_safe: first-maximum card, its playable / plausible copies
};

template <class SH>
__device__ __forceinline__ void rule_cores(const Game<SH>& g, const unsigned int used, Cores& o) {
  constexpr int C = SH::C, H = SH::H, P = SH::P;
  const uint32_t pg = g.playable();
  const uint32_t done = done_rows(g);
  const uint32_t ltmin = below_all_grid(g);
  const uint32_t useful = useful_grid(g);

  // Empty slots are the LAST slots of a hand (cards shift down, the deck refills the end) and every scan
  // below keeps the lowest matching index, so the tests may ignore whether a slot holds a card: a match
  // at an index >= the hand size means that no real card matched.
  const int n0 = g.hand_n(0);

  o.play_safe = -1;                                    // :350-380: every plausible identity playable
  if (used & (1u << R_PLAY_SAFE_CARD)) {
    int idx = H;
#pragma unroll
    for (int i = H - 1; i >= 0; --i)
      if ((slot_grid(g.hand[0][i]) & ~pg) == 0) idx = i;
    if (idx < n0) o.play_safe = uid_play<SH>(idx);
  }

  o.play_certain = -1;                                 // :383-409
  if (used & (1u << R_PLAY_IF_CERTAIN)) {
#pragma unroll
    for (int i = H - 1; i >= 0; --i) {
      const uint32_t w = g.hand[0][i];
      if ((w & (3u << 25)) == (3u << 25) && (g.oh[0][i] & pg) != 0) o.play_certain = uid_play<SH>(i);
    }
  }

  // play_probably_safe_factory  ruleset.py:617-635 with get_card_playability :82-125.
  // playability = playable copies / plausible copies among the cards not visible to me; the
  // first-maximum (np.argmax) is taken by exact cross-multiplication.  An empty slot counts 0 / 0 and
  // never beats the running maximum.
  o.best = 0; o.bp = 0; o.bt = 1;
  if (used & ((1u << R_PLAY_PROBABLY_SAFE) | (1u << R_HAIL_MARY))) {
    uint32_t l0, l1;
    unseen_planes(g, l0, l1);
    int p[H], t[H];
#pragma unroll
    for (int i = 0; i < H; ++i) {
      const uint32_t grid = slot_grid(g.hand[0][i]);
      t[i] = count_planes(grid, l0, l1);
      p[i] = count_planes(grid & pg, l0, l1);
    }
    // first maximum of p_i / t_i (np.argmax) as a tournament: p_j/t_j > p_i/t_i <=> p_j*t_i > p_i*t_j
    // (t > 0; an empty slot is 0 / 0 and never beats anything).  In a pair the later card wins only if
    // strictly greater, so ties go to the lowest index at every level.
    int bi[H];
#pragma unroll
    for (int i = 0; i < H; ++i) bi[i] = i;
#pragma unroll
    for (int span = 1; span < H; span *= 2)
#pragma unroll
      for (int i = 0; i + span < H; i += 2 * span) {
        const bool later = p[i + span] * t[i] > p[i] * t[i + span];
        if (later) { p[i] = p[i + span]; t[i] = t[i + span]; bi[i] = bi[i + span]; }
      }
    o.bp = p[0]; o.bt = t[0]; o.best = bi[0];
  }

  o.osawa = -1;                                        // :220-280
  if (used & (1u << R_OSAWA_DISCARD)) {
    // Per card: colour hinted -> its stack is finished; NOT hinted -> fireworks[-1] is finished (quirk
    // :229-230); both hinted -> rank below its stack or at / beyond max_fireworks; rank hinted -> rank
    // below every stack.  The four hint states pick one of four grids the card's one-hot is tested against.
    // (kOsawaIntended in `used`: the intended reading of :229, un-hinted colours are skipped -> `all` = 0)
    const uint32_t all = (!(used & kOsawaIntended) && ((g.fw >> (5 * (C - 1) + 4)) & 1u)) ? SH::kGridAll : 0u;
    const uint32_t g1 = done, g2 = ltmin | all, g3 = done | ltmin | ((g.fw | ~useful) & SH::kGridAll);
    int first = H, second = H;
#pragma unroll
    for (int i = H - 1; i >= 0; --i) {
      const uint32_t w = g.hand[0][i];
      const bool ch = (w & (1u << 25)) != 0, rh = (w & (1u << 26)) != 0;
      const uint32_t sel = rh ? (ch ? g3 : g2) : (ch ? g1 : all);
      if ((g.oh[0][i] & sel) != 0) first = i;             // an empty slot has no one-hot
      if ((slot_grid(w) & useful) == 0) second = i;    // no plausible identity is still useful
    }
    if (second >= n0) second = H;
    const int pick = first < H ? first : second;
    if (pick < H) o.osawa = uid_discard<SH>(pick);
  }

  // The three "tell" scans pick the FIRST card (seat, then slot order) that satisfies their test;
  // each keeps the slot word of that card and the move is decoded once, after the scan.
  o.tell_playable = -1;                                // tell_playable_card_outer :413-451
  o.tell_dispensable = -1;                             // tell_dispensable_factory :454-514
  o.tell_unknown = -1;                                 // :285-309 (next player, colour first)
  const bool want_tp = used & ((1u << R_TELL_PLAYABLE_CARD_OUTER) | (1u << R_TELL_ANYONE_USEFUL_CARD));
  const bool want_td = used & (1u << R_TELL_DISPENSABLE);
  const bool want_tu = used & (1u << R_TELL_UNKNOWN);
  if (want_tp || want_td || want_tu) {
    uint32_t tp_w = 0, td_w = 0, tu_w = 0;
    int tp_k = 1, td_k = 1;
    // tell_dispensable by hint state: nothing hinted -> finished colour or rank below every stack; colour
    // hinted -> rank below every stack, or below its own stack; rank hinted -> finished colour, or below its
    // own stack.  A fully hinted card is told nothing by any of the three rules: its one-hot is dropped.
    // (a scan the rule list does not need tests against empty grids: no flag inside the loop)
    const uint32_t tdm = want_td ? SH::kGridAll : 0u;
    const uint32_t t0 = (done | ltmin) & tdm, t1 = (ltmin | g.fw) & tdm, t2 = (done | g.fw) & tdm;
    const uint32_t pg_tp = want_tp ? pg : 0u, all_tu = want_tu ? SH::kGridAll : 0u;
#pragma unroll
    for (int k = P - 1; k >= 1; --k)                   // reverse order, last write = first match
#pragma unroll
      for (int i = H - 1; i >= 0; --i) {
        const uint32_t w = g.hand[k][i];
        const bool ch = (w & (1u << 25)) != 0, rh = (w & (1u << 26)) != 0;
        const uint32_t oh = (ch && rh) ? 0u : g.oh[k][i];
        if ((oh & pg_tp) != 0) { tp_w = w; tp_k = k; }
        if ((oh & (ch ? t1 : (rh ? t2 : t0))) != 0) { td_w = w; td_k = k; }
        if (k == 1 && (oh & all_tu) != 0) tu_w = w;
      }
    if (tp_w != 0) {                                   // rank first, then colour
      const int card = slot_card(tp_w);
      o.tell_playable = !slot_rhint(tp_w) ? uid_rank<SH>(tp_k, card_rank(card))
                                          : uid_color<SH>(tp_k, card_color(card));
    }
    if (td_w != 0) {
      const uint32_t oh = slot_onehot(td_w);
      const int card = slot_card(td_w), c = card_color(card), r = card_rank(card);
      const bool ch = slot_chint(td_w), rh = slot_rhint(td_w);
      const bool f5 = (oh & done) != 0, lt = (oh & ltmin) != 0;
      // :470-510 in order: colour of a finished stack, rank below every stack, then whichever of
      // colour / rank is still unknown of a card below its stack
      const bool by_color = (!ch && f5) || (!(!rh && lt) && !ch);
      o.tell_dispensable = by_color ? uid_color<SH>(td_k, c) : uid_rank<SH>(td_k, r);
    }
    if (tu_w != 0) {                                   // colour first, then rank
      const int card = slot_card(tu_w);
      o.tell_unknown = !slot_chint(tu_w) ? uid_color<SH>(1, card_color(card))
                                         : uid_rank<SH>(1, card_rank(card));
    }
  }
}

// discard_probably_useless_factory  ruleset.py:638-649 with get_probability_useless :127-170
// (intended semantics): same bookkeeping as play_probably_safe with the useless grid.
template <class SH>
__device__ __forceinline__ int discard_probably_useless(const Game<SH>& g, double threshold,
                                                        const RuleSet& rs, int slot) {
  const int n = g.hand_n(0);
  if (!(g.info < 8) || n == 0) return -1;
  uint32_t l0, l1;
  unseen_planes(g, l0, l1);
  const uint32_t ug = useless_grid(g);
  int best = 0, bu = 0, bt = 1;
#pragma unroll
  for (int i = 0; i < SH::H; ++i) {
    const uint32_t grid = slot_grid(g.hand[0][i]);
    const int total = count_planes(grid, l0, l1);
    const int useless = count_planes(grid & ug, l0, l1);
    if (i == 0 || useless * bt > bu * total) { bu = useless; bt = total; best = i; }
  }
  if (ratio_at_least(rs, slot, bu, bt, threshold)) return uid_discard<SH>(best);
  return -1;
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

// A rule that is evaluated only when a game reaches it without a move: the random rules (two 64-bit
// hashes each), play_probably_safe without a comparison table, and the rules that are dead in the
// reference.  Returns a move uid or -1.
template <class SH, bool EXT>
__device__ __forceinline__ int lazy_rule(const Game<SH>& g, const Cores& co, const RuleSet& rs, int k,
                                         unsigned long long game_id, unsigned long long step) {
  constexpr int H = SH::H;
  const unsigned long long seed = rs.seed;
  const int rule = rs.id[k];
  const int n0 = g.hand_n(0);
  if constexpr (EXT) {
    if (rule >= R_TELL_ANYONE_USELESS_CARD)
      return run_ext_rule(g, rule, rs.farg[k], rs, rs.tslot[k], seed, game_id, step);
  }
  switch (rule) {
    case R_LEGAL_RANDOM: {                             // :598-604 (intended semantics)
      unsigned long long m = legal_mask(g);
      const int n = __popcll(m);
      if (n == 0) return -1;
      int j = rand_below(rule_rand(seed, game_id, step, 2), n);
      while (j-- > 0) m &= m - 1;
      return __ffsll((long long)m) - 1;
    }
    case R_TELL_DISPENSABLE:                           // a third tell_dispensable of one list
      return (g.info < rs.iarg[k] && g.info > 0) ? co.tell_dispensable : -1;
    case R_PLAY_PROBABLY_SAFE:                         // more threshold rules than comparison tables
      if (n0 > 0 && (rs.iarg[k] == 0 || g.life > 1) &&
          ratio_at_least(rs, rs.tslot[k], co.bp, co.bt, rs.farg[k]))
        return uid_play<SH>(co.best);
      return -1;
    default:
      return -1;
  }
}

// RulebasedAgent.get_move  rule_based.py:13-25: the first rule of the list that yields a move wins.
// *hit = index of the rule that fired (rs.n = the legal_random fallback, stored by the host as
// entry [n]).
// No walk over the list: every row's verdict becomes the key (position << 8 | move), "no move" keys
// sort last, and the winner is one chain of 3-input minima.  (A `switch (rule id)` per list entry was
// 290 instructions per step and 30 % of the stall samples of the self-play kernel, a shared-memory
// table of verdicts read back per entry still 250: profiles/r01_selfplay_history.md, v5.)
template <class SH, bool EXT = false>
__device__ __forceinline__ int rules_resolve(const Game<SH>& g, const Cores& co, const RuleSet& rs,
                                             const unsigned long long* __restrict__ sthr,
                                             unsigned long long game_id, unsigned long long step,
                                             int* hit) {
  const int n0 = g.hand_n(0);
  const bool can_tell = g.info > 0, can_discard = g.info < 8;
  const int play_best = n0 > 0 ? uid_play<SH>(co.best) : -1;
  // -1 & 0x100ff = "no move"
#define HB_KEY(row, gate, mv) ((((gate) ? (unsigned int)(mv) : ~0u) & (kKeyNone | 0xffu)) | rs.key_hi[row])
  unsigned int best = kKeyNone;
  best = min(best, HB_KEY(ROW_DISCARD_OLDEST, can_discard, 0));                       // ruleset.py:206-216
  best = min(best, HB_KEY(ROW_OSAWA, g.info != 8, co.osawa));
  best = min(best, HB_KEY(ROW_TELL_UNKNOWN, can_tell, co.tell_unknown));
  best = min(best, HB_KEY(ROW_PLAY_SAFE, true, co.play_safe));
  best = min(best, HB_KEY(ROW_PLAY_CERTAIN, true, co.play_certain));
  best = min(best, HB_KEY(ROW_TELL_PLAYABLE, can_tell, co.tell_playable));            // :413-451, :518-519
  best = min(best, HB_KEY(ROW_TELL_DISPENSABLE0, can_tell && g.info < rs.td_ilt[0], co.tell_dispensable));
  best = min(best, HB_KEY(ROW_TELL_DISPENSABLE1, can_tell && g.info < rs.td_ilt[1], co.tell_dispensable));
  // hail_mary :653-655 = play_probably_safe(0.0) when the deck is empty and a life can be spared
  best = min(best, HB_KEY(ROW_HAIL_MARY, SH::DECK - g.deck_pos == 0 && g.life > 1, play_best));
  // The two cheap random rules are rows as well: with one hash per (game, step) their verdict is a dozen
  // instructions, while as lazily evaluated entries they were two divergent branches that nearly every
  // warp took anyway (some game of the 32 gets that far).
  if (rs.used & ((1u << R_TELL_RANDOMLY) | (1u << R_DISCARD_RANDOMLY))) {             // uniform
    const int n1 = g.hand_n(1);
    const int idx = rand_below(rule_rand(rs.seed, game_id, step, 0), n1);             // :318
    const int coin = rand_below(rule_rand(rs.seed, game_id, step, 1), 2);             // :321
    uint32_t w = 0;
#pragma unroll
    for (int i = 0; i < SH::H; ++i)
      if (i == idx) w = g.hand[1][i];
    const int card = slot_card(w);
    const int tr = coin == 0 ? uid_rank<SH>(1, card_rank(card)) : uid_color<SH>(1, card_color(card));
    best = min(best, HB_KEY(ROW_TELL_RANDOMLY, can_tell && n1 > 0, tr));              // :314-347
    const int dr = uid_discard<SH>(rand_below(rule_rand(rs.seed, game_id, step, 3), n0));
    best = min(best, HB_KEY(ROW_DISCARD_RANDOMLY, can_discard && n0 > 0, dr));        // :607-614
  }
  {                                                      // play_probably_safe :617-635 (:105, :622)
    const int bt = co.bt < kThrRows ? co.bt : 0;         // row 0 of a table is empty
#define HB_PPS(s)                                                                                      \
    {                                                                                                  \
      const bool ok = (rs.slot_extra[s] == 0 || g.life > 1) && ((sthr[(s) * kThrRows + bt] >> co.bp) & 1ull); \
      best = min(best, HB_KEY(ROW_PPS0 + (s), ok, play_best));                                         \
    }
    if (rs.n_slots > 0) HB_PPS(0)                        // uniform; one table is the common case
#pragma unroll 1
    for (int s = 1; s < rs.n_slots; ++s) HB_PPS(s)
#undef HB_PPS
  }
#undef HB_KEY
  // the lazily evaluated entries, in list order: one matters only to a game that gets there
  for (unsigned int lz = rs.lazy; lz != 0; lz &= lz - 1) {                            // uniform
    const int k = __ffs((int)lz) - 1;
    if ((unsigned int)(k << 8) < best) {
      const int r = lazy_rule<SH, EXT>(g, co, rs, k, game_id, step);
      if (r >= 0) best = (unsigned int)(k << 8) | (unsigned int)r;
    }
  }
  const bool fired = best < kKeyNone;
  *hit = fired ? (int)(best >> 8) : rs.n;
  return fired ? (int)(best & 0xffu) : 0;
}

// The comparison tables, copied to shared memory by the whole CTA (the row read per game depends on
// the game: an indexed read of the kernel-argument bank would replay per distinct row).  Every thread
// of the CTA must call it; ends with a barrier.
__device__ __forceinline__ void stage_thresholds(const RuleSet& rs, unsigned long long* sthr) {
  for (int i = threadIdx.x; i < rs.n_slots * kThrRows; i += blockDim.x) sthr[i] = rs.thr[0][i];
  __syncthreads();
}

template <class SH, bool EXT = false>
__device__ __forceinline__ int rules_act(const Game<SH>& g, const RuleSet& rs,
                                         const unsigned long long* __restrict__ sthr,
                                         unsigned long long game_id, unsigned long long step,
                                         int* hit) {
  Cores co;
  rule_cores(g, rs.used, co);
  return rules_resolve<SH, EXT>(g, co, rs, sthr, game_id, step, hit);
}

}  // namespace hbg
