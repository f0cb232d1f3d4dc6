/*
 * TEST INFRASTRUCTURE -- not part of the product.
 *
 * CPU restatement of the Hanabi engine + canonical observation encoder the
 * reference consumes.  The engine is NOT in /root/reference: it is the
 * third-party `hanabi_learning_environment` (HLE, deepmind; the reference uses
 * an un-pinned fork exposing `pyhanabi_pybind`), imported at
 *   hanabi_agents/rule_based/ruleset.py:4, rule_based.py:2,
 *   hanabi_agents/rainbow/run_experiment.py:36
 * and absent from this machine.  This file restates HLE's published algorithm
 * (hanabi_lib/{hanabi_game,hanabi_state,hanabi_hand,hanabi_observation,
 * canonical_encoders}.cc and rl_env.py `step`) as summarised in SURVEY.md
 * Appendix A.  The reference holds no golden transition / observation vector
 * for this path, so PARITY IS UNPINNED: what pins this oracle is (a) the
 * shape constants of the upstream encoder (658/956/1041/1280 bits, 20/30/38/48
 * moves, 171/11 for Hanabi-Small), (b) the Appendix A.6 invariants, and (c) an
 * independent second restatement in Python (tests/hle_py.py) diffed against it.
 */
#ifndef HANABI_ORACLE_H_
#define HANABI_ORACLE_H_
#include <stdint.h>

#define HO_MAXP 5
#define HO_MAXH 5
#define HO_MAXC 5
#define HO_MAXR 5
#define HO_MAXDECK 50
#define HO_MAXCARDS (HO_MAXC * HO_MAXR)
#define HO_MAX_OBS 2048            /* longest observation: 1280 (Full 5p) + 100 (seer) */

/* move types, HLE HanabiMove::Type order */
enum { HO_INVALID = 0, HO_PLAY = 1, HO_DISCARD = 2, HO_REVEAL_COLOR = 3, HO_REVEAL_RANK = 4, HO_DEAL = 5 };
/* dm_env StepType as the agents read it (rlax_rainbow.py:294,300,306) */
enum { HO_FIRST = 0, HO_MID = 1, HO_LAST = 2 };

typedef struct {
  int players, colors, ranks, hand_size, max_info, max_life;
  /* derived */
  int deck_size, num_moves, obs_len, bits_per_card;
  int hands_len, board_len, discard_len, last_action_len, knowledge_len;
  int card_count[HO_MAXCARDS]; /* instances per card index (color*ranks+rank) */
} ho_game;

typedef struct {
  int type, card_index, target_offset, color, rank;
} ho_move;

typedef struct {
  /* deck: explicit order, next card = deck[deck_pos] */
  uint8_t deck[HO_MAXDECK];
  int deck_pos;
  int hand_n[HO_MAXP];
  int hand_card[HO_MAXP][HO_MAXH];   /* card index, oldest first */
  int color_plaus[HO_MAXP][HO_MAXH]; /* bitmask over colors */
  int rank_plaus[HO_MAXP][HO_MAXH];
  int color_hint[HO_MAXP][HO_MAXH];  /* hinted value or -1 */
  int rank_hint[HO_MAXP][HO_MAXH];
  int discard_count[HO_MAXCARDS];
  int discard_n;
  int discard_pile[HO_MAXDECK];
  int fireworks[HO_MAXC];
  int info, life, cur_player, turns_to_play;
  /* most recent non-deal history item (HanabiHistoryItem) */
  int last_valid, last_player;
  ho_move last_move;
  int last_card_color, last_card_rank, last_scored, last_info_token, last_reveal_mask,
      last_newly_revealed;
  int step_type; /* of the most recent step output */
  int64_t episode;
} ho_state;

typedef struct ho_env ho_env;

/* game description */
int ho_game_init(ho_game* g, int players, int colors, int ranks, int hand_size, int max_info,
                 int max_life);
ho_move ho_get_move(const ho_game* g, int uid);
int ho_get_move_uid(const ho_game* g, ho_move m);

/* single-state engine (used by the rules oracle too) */
void ho_state_reset(const ho_game* g, ho_state* s, const uint8_t* deck);
int ho_move_is_legal(const ho_game* g, const ho_state* s, ho_move m);
int ho_is_terminal(const ho_game* g, const ho_state* s);
int ho_score(const ho_game* g, const ho_state* s);
/* applies a player move + the chance deal; returns reward, or -1000 if illegal */
int ho_apply_action(const ho_game* g, ho_state* s, int uid);
void ho_legal_mask(const ho_game* g, const ho_state* s, int player, uint8_t* out /*[num_moves]*/);
void ho_encode(const ho_game* g, const ho_state* s, int observer, uint8_t* out /*[obs_len]*/);

/* batched environment */
ho_env* ho_env_create(int players, int colors, int ranks, int hand_size, int max_info,
                      int max_life, int64_t n_games);
void ho_env_destroy(ho_env* e);
const ho_game* ho_env_game(const ho_env* e);
ho_state* ho_env_state(ho_env* e, int64_t g);
int ho_env_set_decks(ho_env* e, const uint8_t* decks, int64_t n_decks);
void ho_env_set_threads(ho_env* e, int threads);
int ho_env_threads(const ho_env* e);
int ho_env_reset(ho_env* e, uint8_t* obs, uint8_t* legal, uint8_t* step_type);
/* returns the number of illegal actions seen (those games are left untouched) */
int64_t ho_env_step(ho_env* e, const int32_t* actions, uint8_t* obs, uint8_t* legal, float* reward,
                    uint8_t* step_type);
void ho_env_observe(ho_env* e, int offset, uint8_t* obs, uint8_t* legal);
/* observation types of HLE: 0 kMinimal, 1 kCardKnowledge (= ho_encode), 2 kSeer */
int ho_obs_len_typed(const ho_game* g, int obs_type);
void ho_encode_typed(const ho_game* g, const ho_state* s, int observer, int obs_type, uint8_t* out);
void ho_env_observe_typed(ho_env* e, int offset, int obs_type, uint8_t* obs /*[n, ho_obs_len_typed]*/);
/* canonical int32 dump of one game's state (layout documented in hanabi_oracle.c) */
int ho_state_dump_len(const ho_game* g);
void ho_env_dump(ho_env* e, int32_t* out /*[n_games, dump_len]*/);

#endif
