/*
 * TEST INFRASTRUCTURE -- not part of the product.
 *
 * CPU restatement of the reference's rule-based agents:
 *   /root/reference/hanabi_agents/rule_based/ruleset.py   (rules, helpers)
 *   /root/reference/hanabi_agents/rule_based/rule_based.py:13-25 (first rule that fires)
 * over the restated observer view of hanabi_oracle.c.  No reference test pins these rules and
 * they need the absent `pyhanabi_pybind` objects to run: PARITY UNPINNED.
 *
 * Divergences from the literal reference (documented in DESIGN.md):
 *  - python `random` draws (tell_randomly :318-321, legal_random :601, discard_randomly :612)
 *    become a counter-based stream keyed by (seed, game id, step index, draw index).
 *  - legal_random / discard_randomly index the observation dict-style and raise TypeError on
 *    the pybind objects (ruleset.py:598-614); their INTENDED behaviour is restated.
 *  - tell_unknown builds its hint from the knowledge object's unknown colour/rank
 *    (ruleset.py:291-307), an illegal move; the intended value (the real card's) is used.
 *  - osawa_discard :229-230 evaluates fireworks[card.color] for un-hinted cards; with
 *    kUnknownColor = -1 that is Python's negative index = the LAST colour's stack.  Kept.
 *  - tell_anyone_useless_card :522-535, tell_playable_card :570-594 and
 *    discard_probably_useless_factory :638-649 (+ get_probability_useless :127-170, useless_card
 *    :35-42) index the observation dict-style and raise TypeError on the pybind objects; their
 *    INTENDED behaviour is restated, the coin of tell_playable_card :584 is draw 4.
 */
#include "hanabi_oracle.h"
#include <string.h>

enum {
  RO_DISCARD_OLDEST_FIRST = 0, RO_OSAWA_DISCARD = 1, RO_TELL_UNKNOWN = 2, RO_TELL_RANDOMLY = 3,
  RO_PLAY_SAFE_CARD = 4, RO_PLAY_IF_CERTAIN = 5, RO_TELL_PLAYABLE_CARD_OUTER = 6,
  RO_TELL_DISPENSABLE = 7, RO_TELL_ANYONE_USEFUL_CARD = 8, RO_PLAY_PROBABLY_SAFE = 9,
  RO_HAIL_MARY = 10, RO_LEGAL_RANDOM = 11, RO_DISCARD_RANDOMLY = 12, RO_TELL_MOST_INFORMATION = 13,
  RO_TELL_ANYONE_USELESS_CARD = 14, RO_TELL_PLAYABLE_CARD = 15, RO_DISCARD_PROBABLY_USELESS = 16,
  RO_NUM_RULES = 17
};

static const int kNumInDeckByRank[5] = {3, 2, 2, 2, 1}; /* ruleset.py:9 */

static uint32_t ro_rand(uint64_t seed, uint64_t game, uint64_t step, uint32_t k) {
  /* 64-bit key per (seed, game), one 32-bit multiply-xorshift hash per (game, step); index draws
   * (k = 0, 2, 3) read it from the top, coin draws (k = 1, 4) from bit 15 down: at most one random
   * rule decides a move, so only the draws of one rule need to be independent of each other */
  uint64_t x = seed + 0x9E3779B97F4A7C15ull * (game + 1);
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  uint32_t h = (uint32_t)x + (uint32_t)step * 0x9E3779B1u;
  h ^= h >> 16; h *= 0x7FEB352Du;
  h ^= h >> 15; h ^= (uint32_t)(x >> 32); h *= 0x846CA68Bu;
  h ^= h >> 16;
  return (k == 1 || k == 4) ? h << 16 : h;
}
static int ro_below(uint32_t r, int n) { return (int)(((uint64_t)r * (uint64_t)n) >> 32); }

typedef struct {
  const ho_game* g;
  const ho_state* s;
  int me;
  uint64_t seed, game_id, step;
} view;

static int pl(const view* v, int offset) { return (v->me + offset) % v->g->players; }
static int mk(const view* v, int type, int idx, int off, int color, int rank) {
  ho_move m = {type, idx, off, color, rank};
  return ho_get_move_uid(v->g, m);
}
static int min_fireworks(const view* v) {
  int m = v->s->fireworks[0];
  for (int c = 1; c < v->g->colors; ++c) if (v->s->fireworks[c] < m) m = v->s->fireworks[c];
  return m;
}

/* get_max_fireworks  ruleset.py:174-193 : 5 unless every copy of a rank is discarded */
static void max_fireworks(const view* v, int* out) {
  for (int c = 0; c < v->g->colors; ++c) out[c] = 5;
  for (int c = 0; c < v->g->colors; ++c)
    for (int r = 0; r < v->g->ranks; ++r) {
      const int n = v->s->discard_count[c * v->g->ranks + r];
      if (n > 0 && n >= kNumInDeckByRank[r] && out[c] >= r) out[c] = r;
    }
}

/* get_card_playability  ruleset.py:82-125 (player_offset = 0) */
static void card_playability(const view* v, double* out) {
  const ho_game* g = v->g; const ho_state* s = v->s;
  int visible[HO_MAXCARDS] = {0}; /* get_visible_cards :60-76 */
  for (int k = 1; k < g->players; ++k)
    for (int i = 0; i < s->hand_n[pl(v, k)]; ++i) visible[s->hand_card[pl(v, k)][i]]++;
  for (int c = 0; c < g->bits_per_card; ++c) visible[c] += s->discard_count[c];
  for (int c = 0; c < g->colors; ++c)
    for (int r = 0; r < s->fireworks[c]; ++r) visible[c * g->ranks + r]++;
  for (int i = 0; i < s->hand_n[v->me]; ++i) {
    int total = 0, playable = 0;
    for (int c = 0; c < g->colors; ++c)
      for (int r = 0; r < g->ranks; ++r)
        if ((s->color_plaus[v->me][i] >> c & 1) && (s->rank_plaus[v->me][i] >> r & 1)) {
          const int n = g->card_count[c * g->ranks + r] - visible[c * g->ranks + r];
          total += n;
          if (r == s->fireworks[c]) playable += n; /* playable_card :14-33 */
        }
    out[i] = (double)playable / (double)total; /* :105 */
  }
}

/* play_probably_safe_factory  ruleset.py:617-635 */
static int play_probably_safe(const view* v, double threshold, int require_extra_lives) {
  double p[HO_MAXH];
  const int n = v->s->hand_n[v->me];
  if (n == 0) return -1;
  card_playability(v, p);
  int best = 0;
  for (int i = 1; i < n; ++i) if (p[i] > p[best]) best = i; /* np.argmax: first maximum */
  if (!require_extra_lives || v->s->life > 1)
    if (p[best] >= threshold) return mk(v, HO_PLAY, best, -1, -1, -1);
  return -1;
}

/* useless_card  ruleset.py:35-42 */
static int useless_card(const view* v, const int* maxfw, int c, int r) {
  return r < v->s->fireworks[c] || r >= maxfw[c];
}

/* discard_probably_useless_factory  ruleset.py:638-649 with get_probability_useless :127-170
 * (player_offset = 0), intended semantics */
static int discard_probably_useless(const view* v, double threshold) {
  const ho_game* g = v->g; const ho_state* s = v->s;
  const int n = s->hand_n[v->me];
  if (!(s->info < 8) || n == 0) return -1;
  int maxfw[HO_MAXC];
  max_fireworks(v, maxfw);
  int visible[HO_MAXCARDS] = {0};
  for (int k = 1; k < g->players; ++k)
    for (int i = 0; i < s->hand_n[pl(v, k)]; ++i) visible[s->hand_card[pl(v, k)][i]]++;
  for (int c = 0; c < g->bits_per_card; ++c) visible[c] += s->discard_count[c];
  for (int c = 0; c < g->colors; ++c)
    for (int r = 0; r < s->fireworks[c]; ++r) visible[c * g->ranks + r]++;
  double p[HO_MAXH];
  for (int i = 0; i < n; ++i) {
    int total = 0, useless = 0;
    for (int c = 0; c < g->colors; ++c)
      for (int r = 0; r < g->ranks; ++r)
        if ((s->color_plaus[v->me][i] >> c & 1) && (s->rank_plaus[v->me][i] >> r & 1)) {
          const int k = g->card_count[c * g->ranks + r] - visible[c * g->ranks + r];
          total += k;
          if (useless_card(v, maxfw, c, r)) useless += k;
        }
    p[i] = (double)useless / (double)total; /* :151 */
  }
  int best = 0;
  for (int i = 1; i < n; ++i) if (p[i] > p[best]) best = i; /* np.argmax :643 */
  if (p[best] >= threshold) return mk(v, HO_DISCARD, best, -1, -1, -1);
  return -1;
}

/* tell_playable_card_outer  ruleset.py:413-451 */
static int tell_playable_outer(const view* v) {
  const ho_game* g = v->g; const ho_state* s = v->s;
  if (s->info <= 0) return -1;
  for (int k = 1; k < g->players; ++k) {
    const int t = pl(v, k);
    for (int i = 0; i < s->hand_n[t]; ++i) {
      const int c = s->hand_card[t][i] / g->ranks, r = s->hand_card[t][i] % g->ranks;
      const int playable = (r == s->fireworks[c]);
      if (playable && s->rank_hint[t][i] < 0) return mk(v, HO_REVEAL_RANK, -1, k, -1, r);
      else if (playable && s->color_hint[t][i] < 0) return mk(v, HO_REVEAL_COLOR, -1, k, c, -1);
    }
  }
  return -1;
}

/* one rule; returns move uid or -1 (None) */
static int run_rule(const view* v, int rule, double farg, int iarg) {
  const ho_game* g = v->g; const ho_state* s = v->s;
  const int me = v->me, C = g->colors, R = g->ranks;
  switch (rule) {
    case RO_DISCARD_OLDEST_FIRST: /* :206-216 (constant 8 is the reference's) */
      return s->info < 8 ? mk(v, HO_DISCARD, 0, -1, -1, -1) : -1;

    case RO_OSAWA_DISCARD: { /* :220-280 */
      if (s->info == 8) return -1;
      int mf[HO_MAXC];
      max_fireworks(v, mf);
      const int minfw = min_fireworks(v);
      for (int i = 0; i < s->hand_n[me]; ++i) {
        const int ch = s->color_hint[me][i], rh = s->rank_hint[me][i];
        /* :229-230 `if card.color is not None: if fireworks[card.color] == 5`.  iarg == 0 (default): the
         * quirk reading -- an un-hinted colour is kUnknownColor = -1, never None, so the test runs on
         * fireworks[-1], the LAST colour's stack.  iarg != 0: the intended reading, un-hinted colours are
         * skipped (what the comment two lines above it in the reference, `color = card['color']`, did). */
        if (ch >= 0 || iarg == 0) {
          const int cidx = ch >= 0 ? ch : C - 1;
          if (s->fireworks[cidx] == 5) return mk(v, HO_DISCARD, i, -1, -1, -1);
        }
        if (ch >= 0 && rh >= 0)
          if (rh < s->fireworks[ch] || rh >= mf[ch]) return mk(v, HO_DISCARD, i, -1, -1, -1);
        if (rh >= 0)
          if (rh < minfw) return mk(v, HO_DISCARD, i, -1, -1, -1);
      }
      for (int i = 0; i < s->hand_n[me]; ++i) {
        int eventually = 0;
        for (int c = 0; c < C && !eventually; ++c)
          for (int r = 0; r < R; ++r)
            if ((s->color_plaus[me][i] >> c & 1) && (s->rank_plaus[me][i] >> r & 1) && r < mf[c]) {
              eventually = 1;
              break;
            }
        if (!eventually) return mk(v, HO_DISCARD, i, -1, -1, -1);
      }
      return -1;
    }

    case RO_TELL_UNKNOWN: { /* :285-309, next player only, colour first */
      if (s->info <= 0) return -1;
      const int t = pl(v, 1);
      for (int i = 0; i < s->hand_n[t]; ++i) {
        if (s->color_hint[t][i] < 0) return mk(v, HO_REVEAL_COLOR, -1, 1, s->hand_card[t][i] / R, -1);
        if (s->rank_hint[t][i] < 0) return mk(v, HO_REVEAL_RANK, -1, 1, -1, s->hand_card[t][i] % R);
      }
      return -1;
    }

    case RO_TELL_RANDOMLY: { /* :314-347 */
      if (s->info <= 0) return -1;
      const int t = pl(v, 1);
      if (s->hand_n[t] == 0) return -1;
      const int i = ro_below(ro_rand(v->seed, v->game_id, v->step, 0), s->hand_n[t]);
      const int r = ro_below(ro_rand(v->seed, v->game_id, v->step, 1), 2);
      const int card = s->hand_card[t][i];
      if (r == 0) return mk(v, HO_REVEAL_RANK, -1, 1, -1, card % R);
      return mk(v, HO_REVEAL_COLOR, -1, 1, card / R, -1);
    }

    case RO_PLAY_SAFE_CARD: { /* :350-380: every plausible identity is playable */
      for (int i = 0; i < s->hand_n[me]; ++i) {
        int ok = 1;
        for (int c = 0; c < C && ok; ++c)
          for (int r = 0; r < R; ++r)
            if ((s->color_plaus[me][i] >> c & 1) && (s->rank_plaus[me][i] >> r & 1) &&
                r != s->fireworks[c]) {
              ok = 0;
              break;
            }
        if (ok) return mk(v, HO_PLAY, i, -1, -1, -1);
      }
      return -1;
    }

    case RO_PLAY_IF_CERTAIN: /* :383-409 */
      for (int i = 0; i < s->hand_n[me]; ++i) {
        const int ch = s->color_hint[me][i], rh = s->rank_hint[me][i];
        if (ch >= 0 && rh >= 0 && rh == s->fireworks[ch]) return mk(v, HO_PLAY, i, -1, -1, -1);
      }
      return -1;

    case RO_TELL_PLAYABLE_CARD_OUTER: /* :413-451 */
    case RO_TELL_ANYONE_USEFUL_CARD:  /* :518-519 */
      return tell_playable_outer(v);

    case RO_TELL_DISPENSABLE: { /* :454-514, iarg = min_information_tokens */
      if (!(s->info < iarg) || s->info <= 0) return -1;
      const int minfw = min_fireworks(v);
      for (int k = 1; k < g->players; ++k) {
        const int t = pl(v, k);
        for (int i = 0; i < s->hand_n[t]; ++i) {
          const int c = s->hand_card[t][i] / R, r = s->hand_card[t][i] % R;
          const int chinted = s->color_hint[t][i] >= 0, rhinted = s->rank_hint[t][i] >= 0;
          if (!chinted && s->fireworks[c] == 5) return mk(v, HO_REVEAL_COLOR, -1, k, c, -1);
          if (!rhinted && r < minfw) return mk(v, HO_REVEAL_RANK, -1, k, -1, r);
          if (r < s->fireworks[c]) {
            if (!chinted && rhinted) return mk(v, HO_REVEAL_COLOR, -1, k, c, -1);
            if (chinted && !rhinted) return mk(v, HO_REVEAL_RANK, -1, k, -1, r);
          }
        }
      }
      return -1;
    }

    case RO_PLAY_PROBABLY_SAFE: /* :617-635 */
      return play_probably_safe(v, farg, iarg);

    case RO_HAIL_MARY: /* :653-655 */
      if (g->deck_size - s->deck_pos == 0 && s->life > 1) return play_probably_safe(v, 0.0, 0);
      return -1;

    case RO_LEGAL_RANDOM: { /* :598-604, intended: uniform over legal moves in uid order */
      int legal[64], n = 0;
      for (int u = 0; u < g->num_moves; ++u)
        if (ho_move_is_legal(g, s, ho_get_move(g, u))) legal[n++] = u;
      if (n == 0) return -1;
      return legal[ro_below(ro_rand(v->seed, v->game_id, v->step, 2), n)];
    }

    case RO_DISCARD_RANDOMLY: { /* :607-614, intended */
      if (!(s->info < 8) || s->hand_n[me] == 0) return -1;
      return mk(v, HO_DISCARD, ro_below(ro_rand(v->seed, v->game_id, v->step, 3), s->hand_n[me]),
                -1, -1, -1);
    }

    case RO_TELL_ANYONE_USELESS_CARD: { /* :522-535, intended */
      if (!(s->info > 1)) return -1;
      int maxfw[HO_MAXC];
      max_fireworks(v, maxfw);
      for (int k = 1; k < g->players; ++k) {
        const int t = pl(v, k);
        for (int i = 0; i < s->hand_n[t]; ++i) {
          const int c = s->hand_card[t][i] / R, r = s->hand_card[t][i] % R;
          if (useless_card(v, maxfw, c, r)) {
            if (s->color_hint[t][i] < 0) return mk(v, HO_REVEAL_COLOR, -1, k, c, -1);
            if (s->rank_hint[t][i] < 0) return mk(v, HO_REVEAL_RANK, -1, k, -1, r);
          }
        }
      }
      return -1;
    }

    case RO_TELL_PLAYABLE_CARD: { /* :570-594, intended; the coin :584 is draw 4 */
      if (!(s->info > 0)) return -1;
      for (int k = 1; k < g->players; ++k) {
        const int t = pl(v, k);
        for (int i = 0; i < s->hand_n[t]; ++i) {
          const int c = s->hand_card[t][i] / R, r = s->hand_card[t][i] % R;
          if (r == s->fireworks[c]) {
            const int coin = ro_below(ro_rand(v->seed, v->game_id, v->step, 4), 2);
            return coin == 0 ? mk(v, HO_REVEAL_RANK, -1, k, -1, r) : mk(v, HO_REVEAL_COLOR, -1, k, c, -1);
          }
        }
      }
      return -1;
    }

    case RO_DISCARD_PROBABLY_USELESS: /* :638-649 */
      return discard_probably_useless(v, farg);

    case RO_TELL_MOST_INFORMATION: /* :539-562 always returns None (:562) */
    default:
      return -1;
  }
}

/* RulebasedAgent.get_move  rule_based.py:13-25: first rule that fires, else legal_random.
 * rule_hit[g] = index of the rule that fired, n_rules for the fallback, -1 if the game is
 * waiting for its auto-reset (no move taken). */
int ro_act(ho_env* e, const int32_t* rule_ids, int n_rules, const double* rule_farg,
           const int32_t* rule_iarg, uint64_t seed, uint64_t step, int64_t game_id_base,
           int64_t n_games, int32_t* actions, int32_t* rule_hit) {
  const ho_game* g = ho_env_game(e);
#pragma omp parallel for num_threads(ho_env_threads(e)) schedule(static)
  for (int64_t i = 0; i < n_games; ++i) {
    const ho_state* s = ho_env_state(e, i);
    if (s->step_type == HO_LAST) {
      actions[i] = 0;
      if (rule_hit) rule_hit[i] = -1;
      continue;
    }
    view v = {g, s, s->cur_player, seed, (uint64_t)(game_id_base + i), step};
    int a = -1, hit = n_rules;
    for (int k = 0; k < n_rules; ++k) {
      a = run_rule(&v, rule_ids[k], rule_farg[k], rule_iarg[k]);
      if (a >= 0) { hit = k; break; }
    }
    if (a < 0) a = run_rule(&v, RO_LEGAL_RANDOM, 0.0, 0);
    actions[i] = a < 0 ? 0 : a;
    if (rule_hit) rule_hit[i] = hit;
  }
  return 0;
}
