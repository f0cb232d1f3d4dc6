/*
 * TEST INFRASTRUCTURE -- not part of the product.  See hanabi_oracle.h for what
 * this restates and why parity is "unpinned" for the Hanabi engine.
 *
 * Citations of the form [HLE file:Function] name the upstream
 * hanabi_learning_environment function restated (source not on this machine;
 * SURVEY.md Appendix A is the written spec).
 */
#include "hanabi_oracle.h"
#include <omp.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------- game ---- */

/* [HLE hanabi_game.cc:NumberCardInstances] rank 0 -> 3, last rank -> 1, else 2 (A.1) */
static int card_instances(int ranks, int rank) {
  if (rank == 0) return 3;
  if (rank == ranks - 1) return 1;
  return 2;
}

int ho_game_init(ho_game* g, int players, int colors, int ranks, int hand_size, int max_info,
                 int max_life) {
  if (players < 2 || players > HO_MAXP || colors < 1 || colors > HO_MAXC || ranks < 1 ||
      ranks > HO_MAXR || hand_size < 1 || hand_size > HO_MAXH || max_info < 0 || max_life < 1)
    return -1;
  memset(g, 0, sizeof(*g));
  g->players = players; g->colors = colors; g->ranks = ranks; g->hand_size = hand_size;
  g->max_info = max_info; g->max_life = max_life;
  g->bits_per_card = colors * ranks;
  g->deck_size = 0;
  for (int c = 0; c < colors; ++c)
    for (int r = 0; r < ranks; ++r) {
      g->card_count[c * ranks + r] = card_instances(ranks, r);
      g->deck_size += g->card_count[c * ranks + r];
    }
  if (g->deck_size > HO_MAXDECK || players * hand_size > g->deck_size) return -1;
  /* [HLE hanabi_game.cc:MaxMoves] A.2 */
  g->num_moves = 2 * hand_size + (players - 1) * (colors + ranks);
  /* [HLE canonical_encoders.cc:*SectionLength] A.5 */
  g->hands_len = (players - 1) * hand_size * g->bits_per_card + players;
  g->board_len = (g->deck_size - players * hand_size) + colors * ranks + max_info + max_life;
  g->discard_len = g->deck_size;
  g->last_action_len = players + 4 + players + colors + ranks + hand_size + hand_size +
                       g->bits_per_card + 2;
  g->knowledge_len = players * hand_size * (g->bits_per_card + colors + ranks);
  g->obs_len = g->hands_len + g->board_len + g->discard_len + g->last_action_len + g->knowledge_len;
  return 0;
}

/* [HLE hanabi_game.cc:ConstructMove / GetMove] A.2 */
ho_move ho_get_move(const ho_game* g, int uid) {
  ho_move m = {HO_INVALID, -1, -1, -1, -1};
  const int H = g->hand_size, C = g->colors, R = g->ranks, P = g->players;
  if (uid < 0 || uid >= g->num_moves) return m;
  if (uid < H) { m.type = HO_DISCARD; m.card_index = uid; return m; }
  uid -= H;
  if (uid < H) { m.type = HO_PLAY; m.card_index = uid; return m; }
  uid -= H;
  if (uid < (P - 1) * C) {
    m.type = HO_REVEAL_COLOR; m.target_offset = 1 + uid / C; m.color = uid % C; return m;
  }
  uid -= (P - 1) * C;
  m.type = HO_REVEAL_RANK; m.target_offset = 1 + uid / R; m.rank = uid % R;
  return m;
}

/* [HLE hanabi_game.cc:GetMoveUid] */
int ho_get_move_uid(const ho_game* g, ho_move m) {
  const int H = g->hand_size, C = g->colors, R = g->ranks, P = g->players;
  switch (m.type) {
    case HO_DISCARD: return m.card_index;
    case HO_PLAY: return H + m.card_index;
    case HO_REVEAL_COLOR: return 2 * H + (m.target_offset - 1) * C + m.color;
    case HO_REVEAL_RANK: return 2 * H + (P - 1) * C + (m.target_offset - 1) * R + m.rank;
    default: return -1;
  }
}

/* --------------------------------------------------------------- state ---- */

static int deck_remaining(const ho_game* g, const ho_state* s) { return g->deck_size - s->deck_pos; }

/* [HLE hanabi_state.cc:PlayerToDeal] lowest-numbered player with a short hand, else -1 */
static int player_to_deal(const ho_game* g, const ho_state* s) {
  for (int p = 0; p < g->players; ++p)
    if (s->hand_n[p] < g->hand_size) return p;
  return -1;
}

/* [HLE hanabi_state.cc:ApplyMove kDeal + hanabi_hand.cc:AddCard] card to the END of the hand,
 * fresh knowledge (everything plausible, nothing hinted).  Deck order is an explicit input. */
static void deal_one(const ho_game* g, ho_state* s, int p) {
  const int k = s->hand_n[p]++;
  s->hand_card[p][k] = s->deck[s->deck_pos++];
  s->color_plaus[p][k] = (1 << g->colors) - 1;
  s->rank_plaus[p][k] = (1 << g->ranks) - 1;
  s->color_hint[p][k] = -1;
  s->rank_hint[p][k] = -1;
}

/* [HLE hanabi_state.cc:HanabiState ctor + initial chance deals] A.3 */
void ho_state_reset(const ho_game* g, ho_state* s, const uint8_t* deck) {
  const int64_t ep = s->episode;
  memset(s, 0, sizeof(*s));
  s->episode = ep;
  memcpy(s->deck, deck, (size_t)g->deck_size);
  s->info = g->max_info;
  s->life = g->max_life;
  s->turns_to_play = g->players;
  for (;;) { /* player 0 is filled completely first, then player 1, ... */
    const int p = player_to_deal(g, s);
    if (p < 0 || deck_remaining(g, s) == 0) break;
    deal_one(g, s, p);
  }
  s->cur_player = 0;
  s->last_valid = 0;
  s->step_type = HO_FIRST;
}

/* [HLE hanabi_state.cc:Score] */
int ho_score(const ho_game* g, const ho_state* s) {
  if (s->life <= 0) return 0;
  int sc = 0;
  for (int c = 0; c < g->colors; ++c) sc += s->fireworks[c];
  return sc;
}

/* [HLE hanabi_state.cc:EndOfGameStatus / IsTerminal] */
int ho_is_terminal(const ho_game* g, const ho_state* s) {
  if (s->life < 1) return 1;
  if (ho_score(g, s) >= g->colors * g->ranks) return 1;
  if (s->turns_to_play <= 0) return 1;
  return 0;
}

/* [HLE hanabi_state.cc:MoveIsLegal] A.4 */
int ho_move_is_legal(const ho_game* g, const ho_state* s, ho_move m) {
  const int p = s->cur_player;
  switch (m.type) {
    case HO_DISCARD:
      if (s->info >= g->max_info) return 0;
      if (m.card_index < 0 || m.card_index >= s->hand_n[p]) return 0;
      return 1;
    case HO_PLAY:
      if (m.card_index < 0 || m.card_index >= s->hand_n[p]) return 0;
      return 1;
    case HO_REVEAL_COLOR: {
      if (s->info <= 0) return 0;
      if (m.target_offset < 1 || m.target_offset >= g->players) return 0;
      if (m.color < 0 || m.color >= g->colors) return 0;
      const int t = (p + m.target_offset) % g->players;
      for (int i = 0; i < s->hand_n[t]; ++i)
        if (s->hand_card[t][i] / g->ranks == m.color) return 1;
      return 0;
    }
    case HO_REVEAL_RANK: {
      if (s->info <= 0) return 0;
      if (m.target_offset < 1 || m.target_offset >= g->players) return 0;
      if (m.rank < 0 || m.rank >= g->ranks) return 0;
      const int t = (p + m.target_offset) % g->players;
      for (int i = 0; i < s->hand_n[t]; ++i)
        if (s->hand_card[t][i] % g->ranks == m.rank) return 1;
      return 0;
    }
    default: return 0;
  }
}

/* [HLE hanabi_hand.cc:RemoveFromHand] erase: later cards (and their knowledge) shift down */
static void remove_from_hand(ho_state* s, int p, int idx) {
  for (int i = idx; i + 1 < s->hand_n[p]; ++i) {
    s->hand_card[p][i] = s->hand_card[p][i + 1];
    s->color_plaus[p][i] = s->color_plaus[p][i + 1];
    s->rank_plaus[p][i] = s->rank_plaus[p][i + 1];
    s->color_hint[p][i] = s->color_hint[p][i + 1];
    s->rank_hint[p][i] = s->rank_hint[p][i + 1];
  }
  s->hand_n[p]--;
}

static void to_discard_pile(ho_state* s, int card) {
  s->discard_pile[s->discard_n++] = card;
  s->discard_count[card]++;
}

/* [HLE hanabi_state.cc:IncrementInformationTokens] */
static int increment_info(const ho_game* g, ho_state* s) {
  if (s->info < g->max_info) { s->info++; return 1; }
  return 0;
}

/* [HLE hanabi_state.cc:ApplyMove + AdvanceToNextPlayer, rl_env.py:step] A.3 */
int ho_apply_action(const ho_game* g, ho_state* s, int uid) {
  const ho_move m = ho_get_move(g, uid);
  if (!ho_move_is_legal(g, s, m)) return -1000;
  const int score_before = ho_score(g, s);
  if (deck_remaining(g, s) == 0) --s->turns_to_play; /* evaluated BEFORE the move */
  const int p = s->cur_player;
  s->last_valid = 1;
  s->last_player = p;
  s->last_move = m;
  s->last_card_color = -1; s->last_card_rank = -1;
  s->last_scored = 0; s->last_info_token = 0;
  s->last_reveal_mask = 0; s->last_newly_revealed = 0;
  switch (m.type) {
    case HO_DISCARD: {
      const int card = s->hand_card[p][m.card_index];
      s->last_info_token = increment_info(g, s);
      s->last_card_color = card / g->ranks;
      s->last_card_rank = card % g->ranks;
      remove_from_hand(s, p, m.card_index);
      to_discard_pile(s, card);
      break;
    }
    case HO_PLAY: {
      const int card = s->hand_card[p][m.card_index];
      const int c = card / g->ranks, r = card % g->ranks;
      s->last_card_color = c;
      s->last_card_rank = r;
      /* [HLE hanabi_state.cc:AddToFireworks] */
      if (r == s->fireworks[c]) {
        s->fireworks[c]++;
        s->last_scored = 1;
        if (s->fireworks[c] == g->ranks) s->last_info_token = increment_info(g, s);
        remove_from_hand(s, p, m.card_index);
      } else {
        s->life--;
        remove_from_hand(s, p, m.card_index);
        to_discard_pile(s, card);
      }
      break;
    }
    case HO_REVEAL_COLOR: {
      s->info--;
      const int t = (p + m.target_offset) % g->players;
      /* [HLE hanabi_hand.cc:RevealColor] */
      for (int i = 0; i < s->hand_n[t]; ++i) {
        if (s->hand_card[t][i] / g->ranks == m.color) {
          s->last_reveal_mask |= 1 << i;
          if (s->color_hint[t][i] < 0) s->last_newly_revealed |= 1 << i;
          s->color_hint[t][i] = m.color;           /* ApplyIsValueHint */
          s->color_plaus[t][i] = 1 << m.color;
        } else {
          s->color_plaus[t][i] &= ~(1 << m.color); /* ApplyIsNotValueHint */
        }
      }
      break;
    }
    case HO_REVEAL_RANK: {
      s->info--;
      const int t = (p + m.target_offset) % g->players;
      for (int i = 0; i < s->hand_n[t]; ++i) {
        if (s->hand_card[t][i] % g->ranks == m.rank) {
          s->last_reveal_mask |= 1 << i;
          if (s->rank_hint[t][i] < 0) s->last_newly_revealed |= 1 << i;
          s->rank_hint[t][i] = m.rank;
          s->rank_plaus[t][i] = 1 << m.rank;
        } else {
          s->rank_plaus[t][i] &= ~(1 << m.rank);
        }
      }
      break;
    }
    default: break;
  }
  /* chance node: deal while the deck has cards and a hand is short (rl_env.py:step loop) */
  while (deck_remaining(g, s) > 0) {
    const int d = player_to_deal(g, s);
    if (d < 0) break;
    deal_one(g, s, d);
  }
  s->cur_player = (p + 1) % g->players;
  return ho_score(g, s) - score_before;
}

/* [HLE hanabi_state.cc:LegalMoves] empty unless it is `player`'s turn */
void ho_legal_mask(const ho_game* g, const ho_state* s, int player, uint8_t* out) {
  for (int u = 0; u < g->num_moves; ++u)
    out[u] = (player == s->cur_player) ? (uint8_t)ho_move_is_legal(g, s, ho_get_move(g, u)) : 0;
}

/* ------------------------------------------------------------- encoder ---- */
/* [HLE canonical_encoders.cc:CanonicalObservationEncoder::Encode] A.5.  Observer-relative:
 * offset k means player (observer + k) mod P  [HLE hanabi_observation.cc ctor]. */
void ho_encode(const ho_game* g, const ho_state* s, int observer, uint8_t* out) {
  const int P = g->players, H = g->hand_size, C = g->colors, R = g->ranks, bpc = g->bits_per_card;
  memset(out, 0, (size_t)g->obs_len);
  int off = 0;

  /* 1. hands [EncodeHands]: other players' cards, then one "hand is short" bit per player */
  for (int k = 1; k < P; ++k) {
    const int pl = (observer + k) % P;
    for (int i = 0; i < s->hand_n[pl]; ++i) out[off + i * bpc + s->hand_card[pl][i]] = 1;
    off += H * bpc;
  }
  for (int k = 0; k < P; ++k)
    if (s->hand_n[(observer + k) % P] < H) out[off + k] = 1;
  off += P;

  /* 2. board [EncodeBoard]: deck thermometer, fireworks one-hot, info + life thermometers */
  for (int i = 0; i < deck_remaining(g, s); ++i) out[off + i] = 1;
  off += g->deck_size - P * H;
  for (int c = 0; c < C; ++c) {
    if (s->fireworks[c] > 0) out[off + s->fireworks[c] - 1] = 1;
    off += R;
  }
  for (int i = 0; i < s->info; ++i) out[off + i] = 1;
  off += g->max_info;
  for (int i = 0; i < s->life; ++i) out[off + i] = 1;
  off += g->max_life;

  /* 3. discards [EncodeDiscards]: per card type a thermometer of its instance count */
  for (int c = 0; c < C; ++c)
    for (int r = 0; r < R; ++r) {
      const int card = c * R + r;
      for (int i = 0; i < s->discard_count[card]; ++i) out[off + i] = 1;
      off += g->card_count[card];
    }

  /* 4. last action [EncodeLastAction]: most recent non-deal move; zeros at game start */
  if (s->last_valid) {
    const ho_move m = s->last_move;
    const int rel = (s->last_player - observer + P) % P;
    const int reveal = (m.type == HO_REVEAL_COLOR || m.type == HO_REVEAL_RANK);
    const int card_move = (m.type == HO_PLAY || m.type == HO_DISCARD);
    int o = off;
    out[o + rel] = 1;
    o += P;
    switch (m.type) { /* order: Play, Discard, RevealColor, RevealRank */
      case HO_PLAY: out[o] = 1; break;
      case HO_DISCARD: out[o + 1] = 1; break;
      case HO_REVEAL_COLOR: out[o + 2] = 1; break;
      case HO_REVEAL_RANK: out[o + 3] = 1; break;
      default: break;
    }
    o += 4;
    if (reveal) out[o + (rel + m.target_offset) % P] = 1;
    o += P;
    if (m.type == HO_REVEAL_COLOR) out[o + m.color] = 1;
    o += C;
    if (m.type == HO_REVEAL_RANK) out[o + m.rank] = 1;
    o += R;
    if (reveal)
      for (int i = 0; i < H; ++i)
        if (s->last_reveal_mask & (1 << i)) out[o + i] = 1;
    o += H;
    if (card_move) out[o + m.card_index] = 1;
    o += H;
    if (card_move) out[o + s->last_card_color * R + s->last_card_rank] = 1;
    o += bpc;
    if (m.type == HO_PLAY) {
      if (s->last_scored) out[o] = 1;
      if (s->last_info_token) out[o + 1] = 1;
    }
  }
  off += g->last_action_len;

  /* 5. card knowledge [EncodeCardKnowledge]: every player incl. the observer */
  for (int k = 0; k < P; ++k) {
    const int pl = (observer + k) % P;
    for (int i = 0; i < s->hand_n[pl]; ++i) {
      const int o = off + i * (bpc + C + R);
      for (int c = 0; c < C; ++c)
        if (s->color_plaus[pl][i] & (1 << c))
          for (int r = 0; r < R; ++r)
            if (s->rank_plaus[pl][i] & (1 << r)) out[o + c * R + r] = 1;
      if (s->color_hint[pl][i] >= 0) out[o + bpc + s->color_hint[pl][i]] = 1;
      if (s->rank_hint[pl][i] >= 0) out[o + bpc + C + s->rank_hint[pl][i]] = 1;
    }
    off += H * (bpc + C + R);
  }
}

/* The other two HLE observation types (HanabiGame::AgentObservationType: kMinimal = 0, kCardKnowledge = 1,
 * kSeer = 2; hanabi_game.h / canonical_encoders.cc, [RECALL]: HLE is not on the box).
 *   kMinimal        the card-knowledge section is not encoded (Encode skips EncodeCardKnowledge and
 *                   Shape() leaves its length out);
 *   kCardKnowledge  ho_encode (what every environment of the reference uses, run_experiment.py:114-128);
 *   kSeer           the observer also sees its own cards: EncodeHands runs over every player, offset 0
 *                   first (show_own_cards), so the hands section grows by hand_size * bits_per_card.
 * The hands section is written out here; sections 2-5 are ho_encode's, moved to their new offset. */
int ho_obs_len_typed(const ho_game* g, int obs_type) {
  if (obs_type == 0) return g->obs_len - g->knowledge_len;
  if (obs_type == 2) return g->obs_len + g->hand_size * g->bits_per_card;
  return g->obs_len;
}

void ho_encode_typed(const ho_game* g, const ho_state* s, int observer, int obs_type, uint8_t* out) {
  const int P = g->players, H = g->hand_size, bpc = g->bits_per_card;
  const int len = ho_obs_len_typed(g, obs_type);
  uint8_t full[HO_MAX_OBS];
  ho_encode(g, s, observer, full);            /* sections 2..5 are shared; only their offset moves */
  memset(out, 0, (size_t)len);
  int off = 0;
  /* 1. hands: offset 0 (own cards) first when the observer is a seer */
  for (int k = obs_type == 2 ? 0 : 1; k < P; ++k) {
    const int pl = (observer + k) % P;
    for (int i = 0; i < s->hand_n[pl]; ++i) out[off + i * bpc + s->hand_card[pl][i]] = 1;
    off += H * bpc;
  }
  for (int k = 0; k < P; ++k)
    if (s->hand_n[(observer + k) % P] < H) out[off + k] = 1;
  off += P;
  /* 2-4. board, discards, last action; 5. card knowledge unless minimal */
  const int rest_from = g->hands_len;
  const int rest_len = (obs_type == 0 ? g->obs_len - g->knowledge_len : g->obs_len) - rest_from;
  memcpy(out + off, full + rest_from, (size_t)rest_len);
}

/* --------------------------------------------------------- batched env ---- */
struct ho_env {
  ho_game game;
  int64_t n;
  ho_state* st;
  uint8_t* decks;
  int64_t n_decks;
  int threads;
};

ho_env* ho_env_create(int players, int colors, int ranks, int hand_size, int max_info,
                      int max_life, int64_t n_games) {
  ho_env* e = (ho_env*)calloc(1, sizeof(ho_env));
  if (!e) return NULL;
  if (n_games < 1 || ho_game_init(&e->game, players, colors, ranks, hand_size, max_info, max_life)) {
    free(e);
    return NULL;
  }
  e->n = n_games;
  e->st = (ho_state*)calloc((size_t)n_games, sizeof(ho_state));
  e->threads = 1;
  if (!e->st) { free(e); return NULL; }
  return e;
}

void ho_env_destroy(ho_env* e) {
  if (!e) return;
  free(e->st);
  free(e->decks);
  free(e);
}

const ho_game* ho_env_game(const ho_env* e) { return &e->game; }
ho_state* ho_env_state(ho_env* e, int64_t g) { return &e->st[g]; }
void ho_env_set_threads(ho_env* e, int threads) { e->threads = threads < 1 ? 1 : threads; }
int ho_env_threads(const ho_env* e) { return e->threads; }

/* decks: [n_decks, deck_size] card indices; game g, episode k plays deck (g + k*N) mod n_decks */
int ho_env_set_decks(ho_env* e, const uint8_t* decks, int64_t n_decks) {
  if (n_decks < 1) return -1;
  const size_t bytes = (size_t)n_decks * (size_t)e->game.deck_size;
  /* every deck must be a permutation of the card multiset */
  for (int64_t d = 0; d < n_decks; ++d) {
    int cnt[HO_MAXCARDS] = {0};
    for (int i = 0; i < e->game.deck_size; ++i) {
      const int c = decks[d * e->game.deck_size + i];
      if (c >= e->game.bits_per_card) return -1;
      cnt[c]++;
    }
    for (int c = 0; c < e->game.bits_per_card; ++c)
      if (cnt[c] != e->game.card_count[c]) return -1;
  }
  free(e->decks);
  e->decks = (uint8_t*)malloc(bytes);
  if (!e->decks) return -1;
  memcpy(e->decks, decks, bytes);
  e->n_decks = n_decks;
  return 0;
}

static const uint8_t* deck_for(const ho_env* e, int64_t g, int64_t episode) {
  const int64_t d = (g + episode * e->n) % e->n_decks;
  return e->decks + d * e->game.deck_size;
}

static void emit(const ho_env* e, int64_t g, uint8_t* obs, uint8_t* legal, uint8_t* step_type) {
  const ho_state* s = &e->st[g];
  if (obs) ho_encode(&e->game, s, s->cur_player, obs + g * e->game.obs_len);
  if (legal) ho_legal_mask(&e->game, s, s->cur_player, legal + g * e->game.num_moves);
  if (step_type) step_type[g] = (uint8_t)s->step_type;
}

int ho_env_reset(ho_env* e, uint8_t* obs, uint8_t* legal, uint8_t* step_type) {
  if (!e->decks) return -1;
#pragma omp parallel for num_threads(e->threads) schedule(static)
  for (int64_t g = 0; g < e->n; ++g) {
    e->st[g].episode = 0;
    ho_state_reset(&e->game, &e->st[g], deck_for(e, g, 0));
    emit(e, g, obs, legal, step_type);
  }
  return 0;
}

/* One env step per game: a game whose previous step was LAST is reset onto its next deck
 * (action ignored, reward 0, FIRST); otherwise the action is applied [rl_env.py:step]. */
int64_t ho_env_step(ho_env* e, const int32_t* actions, uint8_t* obs, uint8_t* legal, float* reward,
                    uint8_t* step_type) {
  int64_t illegal = 0;
#pragma omp parallel for num_threads(e->threads) schedule(static) reduction(+ : illegal)
  for (int64_t g = 0; g < e->n; ++g) {
    ho_state* s = &e->st[g];
    float r = 0.0f;
    if (s->step_type == HO_LAST) {
      s->episode++;
      ho_state_reset(&e->game, s, deck_for(e, g, s->episode));
    } else {
      const int d = ho_apply_action(&e->game, s, actions[g]);
      if (d == -1000) {
        illegal++;
      } else {
        r = (float)d;
        s->step_type = ho_is_terminal(&e->game, s) ? HO_LAST : HO_MID;
      }
    }
    if (reward) reward[g] = r;
    emit(e, g, obs, legal, step_type);
  }
  return illegal;
}

/* observation of player (cur_player + offset) mod P; legal mask is empty unless offset == 0 */
void ho_env_observe(ho_env* e, int offset, uint8_t* obs, uint8_t* legal) {
#pragma omp parallel for num_threads(e->threads) schedule(static)
  for (int64_t g = 0; g < e->n; ++g) {
    const ho_state* s = &e->st[g];
    const int pl = (s->cur_player + offset) % e->game.players;
    if (obs) ho_encode(&e->game, s, pl, obs + g * e->game.obs_len);
    if (legal) ho_legal_mask(&e->game, s, pl, legal + g * e->game.num_moves);
  }
}

void ho_env_observe_typed(ho_env* e, int offset, int obs_type, uint8_t* obs) {
  const int len = ho_obs_len_typed(&e->game, obs_type);
#pragma omp parallel for num_threads(e->threads) schedule(static)
  for (int64_t g = 0; g < e->n; ++g) {
    const ho_state* s = &e->st[g];
    ho_encode_typed(&e->game, s, (s->cur_player + offset) % e->game.players, obs_type, obs + g * len);
  }
}

/* Canonical int32 dump (shared with the CUDA env's hb_env_dump, compared in tests):
 *  [0] cur_player [1] info [2] life [3] deck_remaining [4] turns_to_play [5] step_type
 *  [6] score [7] episode(low 31 bits)  then fireworks[C], hand_n[P],
 *  P*H x {card(-1), color_plaus, rank_plaus, color_hint(-1), rank_hint(-1)},
 *  discard_count[C*R],
 *  last: valid, player, type, card_index, target_offset, move_color, move_rank,
 *        card_color, card_rank, scored, info_token, reveal_mask   (all 0 when !valid) */
int ho_state_dump_len(const ho_game* g) {
  return 8 + g->colors + g->players + 5 * g->players * g->hand_size + g->bits_per_card + 12;
}

void ho_env_dump(ho_env* e, int32_t* out) {
  const ho_game* g = &e->game;
  const int len = ho_state_dump_len(g);
  for (int64_t i = 0; i < e->n; ++i) {
    const ho_state* s = &e->st[i];
    int32_t* o = out + i * len;
    int k = 0;
    o[k++] = s->cur_player; o[k++] = s->info; o[k++] = s->life; o[k++] = deck_remaining(g, s);
    o[k++] = s->turns_to_play; o[k++] = s->step_type; o[k++] = ho_score(g, s);
    o[k++] = (int32_t)(s->episode & 0x7fffffff);
    for (int c = 0; c < g->colors; ++c) o[k++] = s->fireworks[c];
    for (int p = 0; p < g->players; ++p) o[k++] = s->hand_n[p];
    for (int p = 0; p < g->players; ++p)
      for (int h = 0; h < g->hand_size; ++h) {
        const int have = h < s->hand_n[p];
        o[k++] = have ? s->hand_card[p][h] : -1;
        o[k++] = have ? s->color_plaus[p][h] : 0;
        o[k++] = have ? s->rank_plaus[p][h] : 0;
        o[k++] = have ? s->color_hint[p][h] : -1;
        o[k++] = have ? s->rank_hint[p][h] : -1;
      }
    for (int c = 0; c < g->bits_per_card; ++c) o[k++] = s->discard_count[c];
    if (s->last_valid) {
      o[k++] = 1; o[k++] = s->last_player; o[k++] = s->last_move.type;
      o[k++] = s->last_move.card_index; o[k++] = s->last_move.target_offset;
      o[k++] = s->last_move.color; o[k++] = s->last_move.rank;
      o[k++] = s->last_card_color; o[k++] = s->last_card_rank;
      o[k++] = s->last_scored; o[k++] = s->last_info_token; o[k++] = s->last_reveal_mask;
    } else {
      for (int j = 0; j < 12; ++j) o[k++] = 0;
    }
  }
}
