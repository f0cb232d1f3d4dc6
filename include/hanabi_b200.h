/*
 * hanabi_b200.h -- C ABI of the B200-native hot path of vsois/hanabi-agents.
 *
 * Plain pointers and sizes only (no torch / pybind types).  Every entry point
 * names the reference interface it replaces (path:line under /root/reference).
 *
 * Conventions
 *   - all functions return HB_OK (0) or a negative HB_ERR_* code;
 *     hb_last_error() gives a human readable message for the calling thread.
 *   - pointers named d_* are DEVICE pointers, h_* are HOST pointers.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *     Device-pointer entry points are stream ordered and do not synchronise.
 *   - *_host entry points take host buffers, copy in, run, copy out and
 *     synchronise: they are what a pybind11 / ctypes drop-in binds.
 *   - there is NO CPU fallback: without a usable CUDA device every compute
 *     entry point fails with HB_ERR_CUDA.
 */
#ifndef HANABI_B200_H_
#define HANABI_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HB_OK 0
#define HB_ERR_INVALID (-1) /* bad argument (NULL, negative size, bad config) */
#define HB_ERR_CUDA (-2)    /* CUDA runtime / driver failure, no device        */
#define HB_ERR_INDEX (-3)   /* index out of range (maps to Python IndexError)  */
#define HB_ERR_STATE (-4)   /* call not valid in the object's current state    */

const char* hb_last_error(void);
/* Library version and the SM architecture the kernels were compiled for. */
int hb_version(void);
const char* hb_build_arch(void);

/* Stream gate (measurement plumbing; replaces nothing in the reference): hb_gate_close enqueues a
 * one-thread kernel that holds `stream` until hb_gate_open is called from the host, so that a
 * benchmark can queue its launches first and time them without the host's launch latency
 * (bench.py).  The wait is bounded (~20 s); the kernel then ends by itself. */
typedef struct hb_gate hb_gate;
int hb_gate_create(int device, hb_gate** out);
int hb_gate_destroy(hb_gate* g);
int hb_gate_close(hb_gate* g, void* stream);
int hb_gate_open(hb_gate* g);

/* ------------------------------------------------------------------------ */
/* Sum tree  (replaces sum_tree/sum_tree/include/sum_tree.h:22-130 and the   */
/* pybind module sum_tree/sum_tree/src/sum_tree_py.cc:7-27)                  */
/* ------------------------------------------------------------------------ */
typedef struct hb_sumtree hb_sumtree;

/* SumTree<float>::SumTree(capacity)  sum_tree.h:25-30 (+populateTree :107-123).
 * depth = ceil(log2(capacity)); 2^depth zero leaves. */
int hb_st_create(int64_t capacity, int device, hb_sumtree** out);
int hb_st_destroy(hb_sumtree* t);
/* getCapacity()  sum_tree.h:76  -- the rounded-up power of two. */
int64_t hb_st_capacity(const hb_sumtree* t);
int hb_st_depth(const hb_sumtree* t);
/* Device pointer to the flat 1-indexed heap (2*capacity floats; [0] unused,
 * [1] root, leaves at [capacity, 2*capacity)).  For tests and fused callers. */
float* hb_st_heap(hb_sumtree* t);

/* updateValues(indices, values)  sum_tree.h:38-44 -> updateValue :32-35 ->
 * updateValue_ :80-90, with the sequential (OMP_NUM_THREADS=1) semantics:
 * elements applied in batch order, diff = value - leaf, every node on the path
 * `+= diff` in fp32.  Out-of-range indices set the sticky error flag and are
 * skipped (the reference segfaults there). */
int hb_st_update(hb_sumtree* t, const int64_t* d_idx, const float* d_val, int64_t n, void* stream);
/* Same, for the contiguous ascending leaf range [start, start+n) -- the shape
 * PriorityBuffer.add_transitions produces (priority_buffer.py:29-31).
 * d_val == NULL means every element takes `fill`. */
int hb_st_update_range(hb_sumtree* t, int64_t start, int64_t n, const float* d_val, float fill,
                       void* stream);
/* Same with the fill value read from DEVICE memory at execution time (the replay shard keeps
 * max_priority on the device: priority_buffer.py:29-31 without a host round trip). */
int hb_st_update_range_dev(hb_sumtree* t, int64_t start, int64_t n, const float* d_fill,
                           void* stream);
/* The same with the range itself in device memory, d_rng = {start, n} with n <= n_max (n_max sizes the
 * launches; n = 0: nothing changes).  No counterpart in the reference (its tree is host code); used by
 * hb_rb_add_masked_async, whose ring cursor lives on the device. */
int hb_st_update_range_indirect(hb_sumtree* t, const int64_t* d_rng, int64_t n_max, const float* d_fill,
                                void* stream);
/* hb_st_update_range without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same kernels and the same bits as hb_st_update_range; the
 * per-element diffs live in a fixed file-scope __device__ array of HB_ST_NOALLOC_MAX_RANGE floats
 * (256 KB, held by every GPU that loads the library until the process exits) instead of a buffer grown
 * with cudaMalloc.  n > HB_ST_NOALLOC_MAX_RANGE is rejected with HB_ERR_INVALID.  The call itself uses
 * no cudaMalloc / cudaFree / allocating routine; the first call on a device looks the array's address up
 * (cudaGetSymbolAddress, once per device).  Every call on a device SHARES the array: calls that overlap
 * on the GPU (different streams, or different trees on different streams) race -- keep them on one
 * stream or order them with events.  Nothing in the array has to start at zero. */
#define HB_ST_NOALLOC_MAX_RANGE 65536
int hb_st_update_range_noalloc(hb_sumtree* t, int64_t start, int64_t n, const float* d_val, float fill,
                               void* stream);
/* getIndices(quantiles)  sum_tree.h:52-59 -> getIndex :46-50 -> getIndex_ :92-105 */
int hb_st_get_indices(hb_sumtree* t, const float* d_q, int64_t* d_out, int64_t n, void* stream);
/* getValues(indices)  sum_tree.h:65-72 (.at() bounds check -> sticky error flag) */
int hb_st_get_values(hb_sumtree* t, const int64_t* d_idx, float* d_out, int64_t n, void* stream);
/* getTotalVal()  sum_tree.h:74 */
int hb_st_total(hb_sumtree* t, float* d_out, void* stream);
/* get_indices + get_values + get_total_val in one launch: the three calls of
 * PriorityBuffer.sample_batch  priority_buffer.py:41-42.  d_total may be NULL. */
int hb_st_sample(hb_sumtree* t, const float* d_q, int64_t* d_idx_out, float* d_val_out,
                 float* d_total, int64_t n, void* stream);
/* Reads and clears the sticky out-of-range flag (synchronises `stream`).
 * Returns HB_OK or HB_ERR_INDEX. */
int hb_st_check(hb_sumtree* t, void* stream);

/* Host-buffer forms: exactly the pybind surface of sum_tree_py.cc:11-18. */
int hb_st_update_host(hb_sumtree* t, const int64_t* h_idx, const float* h_val, int64_t n);
int hb_st_get_indices_host(hb_sumtree* t, const float* h_q, int64_t* h_out, int64_t n);
int hb_st_get_values_host(hb_sumtree* t, const int64_t* h_idx, float* h_out, int64_t n);
int hb_st_total_host(hb_sumtree* t, float* h_out);
/* Copies the whole heap (2*capacity floats) to the host: tests compare every
 * node value with the oracle. */
int hb_st_heap_host(hb_sumtree* t, float* h_out);

/* ------------------------------------------------------------------------ */
/* Batched Hanabi environment.  Replaces, for N games at once, what the       */
/* reference obtains from the third-party HLE through environment.step /      */
/* observation['vectorized'] / legal_moves_as_int                             */
/* (hanabi_agents/rainbow/run_experiment.py:266-273,293,308) and consumes as   */
/* observations[1] = (obs[N,L], legal[N,A])  (rlax_dqn/rlax_rainbow.py:278,285).*/
/* Outputs are bit-packed: obs row = obs_stride 32-bit words (bit b -> word    */
/* b>>5, bit b&31), legal row = legal_words words over move uids; reward f32;  */
/* step_type u8 in dm_env coding 0 FIRST / 1 MID / 2 LAST (rlax_rainbow.py:     */
/* 294,300,306).  A game whose last step was LAST is reset onto its next deck   */
/* by the following step (action ignored, reward 0, FIRST).                     */
/* ------------------------------------------------------------------------ */
typedef struct hb_env hb_env;

/* HanabiGame parameters (players, colors, ranks, hand_size, max_information_tokens,
 * max_life_tokens) as HLE's make('Hanabi-Full', num_players) etc. resolve them. */
int hb_env_create(int players, int colors, int ranks, int hand_size, int max_info, int max_life,
                  int64_t n_games, int device, hb_env** out);
int hb_env_destroy(hb_env* e);
/* what: 0 obs_len(bits) 1 obs_stride(words) 2 num_moves 3 legal_words 4 deck_size
 *       5 state_words 6 n_games 7 dump_len 8 players
 *       9 rule lists whose comparison tables were built so far (they are cached per env) */
int64_t hb_env_info(const hb_env* e, int what);
/* This device holds games [game_id_base, game_id_base+n_games) of an n_total-game job
 * (deck selection and the rule agents' random stream use the global game id). */
int hb_env_set_shard(hb_env* e, int64_t game_id_base, int64_t n_total);
/* Deck pool [n_decks, deck_size] of card indices color*ranks+rank: the order in which HLE's
 * chance node (HanabiState::ApplyMove kDeal / deal_random_card) deals, as an explicit input.
 * Game g, episode k plays deck (g + k*n_total) mod n_decks. */
int hb_env_set_decks(hb_env* e, const uint8_t* d_decks, int64_t n_decks, void* stream);
int hb_env_set_decks_host(hb_env* e, const uint8_t* h_decks, int64_t n_decks); /* validates */
/* environment.reset(): every game dealt from its first deck (run_experiment.py:293) */
int hb_env_reset(hb_env* e, uint32_t* d_obs, uint32_t* d_legal, uint8_t* d_step_type, void* stream);
/* environment.step(action) for every game (run_experiment.py:308): HanabiState::ApplyMove,
 * chance deal, reward = score delta, then observation + legal moves of the player to move.
 * Any output pointer may be NULL. */
int hb_env_step(hb_env* e, const int32_t* d_action, uint32_t* d_obs, uint32_t* d_legal,
                float* d_reward, uint8_t* d_step_type, void* stream);
/* HanabiObservation of player (current + offset) mod players, without stepping. */
int hb_env_observe(hb_env* e, int offset, uint32_t* d_obs, uint32_t* d_legal, void* stream);
/* The observation of player (current + offset) in another of HLE's observation types
 * (HanabiGame::AgentObservationType, chosen by `observation_type` in HLE's game config; the reference's
 * environments all use 1, run_experiment.py:114-128): 0 kMinimal = the canonical vector without the card
 * knowledge section, 1 kCardKnowledge = hb_env_observe, 2 kSeer = own cards encoded as the first block of
 * the hands section.  Rows are bit-packed, stride_words 32-bit words each (>= hb_env_obs_len_as bits). */
int64_t hb_env_obs_len_as(const hb_env* e, int obs_type);
int hb_env_observe_as(hb_env* e, int obs_type, int offset, uint32_t* d_obs, int stride_words, void* stream);
/* Number of illegal actions seen so far (HLE aborts on them; here the game is left as is). */
int hb_env_illegal_count(hb_env* e, int64_t* out, void* stream);
/* Canonical int32 state dump [n_games, dump_len] (tests compare it with the oracle's). */
int hb_env_dump(hb_env* e, int32_t* d_out, void* stream);
/* Bit-packed rows -> one int8 per bit, the layout the reference stores
 * (rlax_dqn/experience_buffer.py:9-12): [n, obs_len] and [n, num_moves]. */
int hb_env_unpack_obs(hb_env* e, const uint32_t* d_packed, int8_t* d_out, int64_t n, void* stream);
int hb_env_unpack_legal(hb_env* e, const uint32_t* d_legal, int8_t* d_out, int64_t n, void* stream);
/* Host-buffer step (copies inside, synchronises): the call a numpy binding makes. */
int hb_env_step_host(hb_env* e, const int32_t* h_action, uint32_t* h_obs, uint32_t* h_legal,
                     float* h_reward, uint8_t* h_step_type);
/* The same loop the reference drives with numpy arrays -- observations = env.step(actions);
 * actions = agent.explore(observations) (hanabi_agents/rainbow/run_experiment.py:293-308) -- as a
 * chunked pipeline over PCIe: the games are cut into `chunks` (1..16) pieces; piece c's actions go up
 * and it steps while piece c-1's outputs travel down and the agent picks piece c-1's next move.
 * hb_env_step_host_async and hb_rules_act_host_async only enqueue; hb_env_host_wait returns when every
 * host buffer is filled.  Host buffers should be pinned.  h_obs rows are ceil(obs_len / 32) words
 * (no alignment padding); any output pointer may be NULL. */
int hb_env_step_host_async(hb_env* e, const int32_t* h_action, uint32_t* h_obs, uint32_t* h_legal,
                           float* h_reward, uint8_t* h_step_type, int chunks);
int hb_rules_act_host_async(hb_env* e, const int32_t* rule_ids, const double* rule_farg,
                            const int32_t* rule_iarg, int n_rules, uint64_t seed, uint64_t step,
                            int32_t* h_action, uint64_t* d_hist);
int hb_env_host_wait(hb_env* e);
/* The same loop double-buffered: the next step's small dependent work (actions up, step kernel, agent
 * kernel, next actions down) runs while this step's outputs are still crossing PCIe.  Per step:
 *   hb_env_step_host_begin(actions)  ->  hb_rules_act_host_async(next actions)  ->
 *   hb_env_step_host_outputs(host buffers; the caller alternates two sets)  ->  hb_env_host_wait_actions()
 * hb_env_host_wait_actions returns when the next actions are on the host (the outputs may still be in
 * flight); hb_env_host_wait_outputs(slot) when the outputs of staging slot `slot` (hb_env_host_slot: the
 * slot of the step begun last) are; hb_env_host_wait when everything is. */
int hb_env_step_host_begin(hb_env* e, const int32_t* h_action);
int hb_env_step_host_outputs(hb_env* e, uint32_t* h_obs, uint32_t* h_legal, float* h_reward,
                             uint8_t* h_step_type);
int hb_env_host_wait_actions(hb_env* e);
int hb_env_host_wait_outputs(hb_env* e, int slot);
int hb_env_host_slot(const hb_env* e);

/* ------------------------------------------------------------------------ */
/* Rule-based agents.  Replaces RulebasedAgent.explore/exploit/get_move        */
/* (hanabi_agents/rule_based/rule_based.py:13-34) over the rules of             */
/* hanabi_agents/rule_based/ruleset.py.  Rule ids:                              */
/*  0 discard_oldest_first :206  1 osawa_discard :220  2 tell_unknown :285      */
/*  3 tell_randomly :314  4 play_safe_card :350  5 play_if_certain :383         */
/*  6 tell_playable_card_outer :413  7 tell_dispensable_factory(iarg) :454      */
/*  8 tell_anyone_useful_card :518  9 play_probably_safe_factory(farg,iarg) :617*/
/* 10 hail_mary :653  11 legal_random :598  12 discard_randomly :607            */
/* 13 tell_most_information :539                                               */
/* ------------------------------------------------------------------------ */
/* d_action[g] = move uid of the first rule that fires (else legal_random); d_rule_hit[g] =
 * its index in the list (n_rules = fallback, -1 = game awaiting reset); d_hist[n_rules+1]
 * accumulates the agent's histogram (rule_based.py:9-10,19-23).  Rule arrays are HOST. */
int hb_rules_act(hb_env* e, const int32_t* rule_ids, const double* rule_farg,
                 const int32_t* rule_iarg, int n_rules, uint64_t seed, uint64_t step,
                 int32_t* d_action, int32_t* d_rule_hit, uint64_t* d_hist, void* stream);
/* Fused self-play: n_steps x {rules act, env step, observe} in one launch with the game state
 * in registers.  Outputs are [n_steps, n_games, ...]; any may be NULL.  Calls that follow each other
 * directly on one stream overlap on the device (the next launch takes over groups of games as this one
 * finishes them: programmatic dependent launch + a FIFO ring of group ids); the results do not depend on
 * it, and any other call or kernel of the stream is ordered as usual. */
int hb_env_selfplay(hb_env* e, const int32_t* rule_ids, const double* rule_farg,
                    const int32_t* rule_iarg, int n_rules, uint64_t seed, uint64_t step0,
                    int n_steps, uint32_t* d_obs, uint32_t* d_legal, float* d_reward,
                    uint8_t* d_step_type, int32_t* d_action, uint64_t* d_hist, void* stream);

/* ------------------------------------------------------------------------ */
/* Replay shard (device resident).  Replaces ExperienceBuffer                  */
/* (hanabi_agents/rlax_dqn/experience_buffer.py:5-97), PriorityBuffer           */
/* (rlax_dqn/priority_buffer.py:13-52) and the transition bookkeeping of        */
/* DQNAgent.add_experience_first / add_experience (rlax_rainbow.py:292-308).    */
/* Observations / legal moves are bit-packed rows as the env emits them.        */
/* ------------------------------------------------------------------------ */
typedef struct hb_replay hb_replay;

/* ExperienceBuffer.__init__ :8-19 (prioritized == 0) / PriorityBuffer.__init__ :14-20 (alpha = 0.6:
 * max_priority = min_priority = alpha, SumTree(capacity)). */
int hb_rb_create(int obs_len, int obs_stride_words, int num_moves, int64_t capacity, int prioritized,
                 float alpha, int device, hb_replay** out);
int hb_rb_destroy(hb_replay* rb);
/* what: 0 capacity, 1 size, 2 oldest_entry, 3 obs_stride, 4 legal_words, 5 prioritized,
 *       6 generation of the sampling scratch (changes when a larger n reallocates it) */
int64_t hb_rb_info(const hb_replay* rb, int what);
/* The shard's sum tree (owned by the shard; NULL when not prioritized). */
hb_sumtree* hb_rb_tree(hb_replay* rb);
/* add_transitions (experience_buffer.py:26-81; priority_buffer.py:22-34: the leaves of the slots
 * about to be written take max_priority first, in get_update_indices order :21-24). */
int hb_rb_add(hb_replay* rb, const uint32_t* d_obs_tm1, const int32_t* d_act, const float* d_rew,
              const uint32_t* d_obs_t, const uint32_t* d_lms, const uint8_t* d_term, int64_t n,
              void* stream);
/* DQNAgent.add_experience_first (rlax_rainbow.py:292-295): last_obs[step_type == 0] = obs[...] */
int hb_rb_store_first(hb_replay* rb, uint32_t* d_last_obs, const uint32_t* d_obs,
                      const uint8_t* d_step_type, int64_t n, void* stream);
/* DQNAgent.add_experience (rlax_rainbow.py:297-308): appends, in row order, the rows with
 * step_type != 0 as (last_obs, act, rew, obs, lms, step_type == 2) and refreshes last_obs for them.
 * Synchronises once to learn the count (the ring cursor is host state, as in the reference). */
int hb_rb_add_masked(hb_replay* rb, uint32_t* d_last_obs, const uint32_t* d_obs, const uint32_t* d_lms,
                     const int32_t* d_act, const float* d_rew, const uint8_t* d_step_type, int64_t n,
                     int64_t* n_added, void* stream);
/* The same in two halves (begin: count the rows and start the copy of the total to the host; finish: wait
 * for it and append), so that a caller with several shards starts every count before waiting for the
 * first.  Same stream for both halves of a shard; finish takes hb_rb_add_masked's arguments. */
int hb_rb_add_masked_begin(hb_replay* rb, const uint8_t* d_step_type, int64_t n, void* stream);
int hb_rb_add_masked_finish(hb_replay* rb, uint32_t* d_last_obs, const uint32_t* d_obs, const uint32_t* d_lms,
                            const int32_t* d_act, const float* d_rew, const uint8_t* d_step_type, int64_t n,
                            int64_t* n_added, void* stream);
/* hb_rb_add_masked with nothing coming back to the host: the ring cursor moves to device memory, the count
 * of appended rows is never read by the host, no call in it synchronises.  The host-side view of size /
 * oldest_entry (hb_rb_info 1, 2; hb_rb_add; hb_rb_add_masked) is refreshed on demand.  Bit-identical shard
 * and tree.  No counterpart in the reference (its buffer is host code): what the self-play session uses for
 * its per-tick inserts. */
int hb_rb_add_masked_async(hb_replay* rb, uint32_t* d_last_obs, const uint32_t* d_obs, const uint32_t* d_lms,
                           const int32_t* d_act, const float* d_rew, const uint8_t* d_step_type, int64_t n,
                           void* stream);
/* __getitem__(indices) (experience_buffer.py:83-87); any output may be NULL. */
int hb_rb_gather(hb_replay* rb, const int64_t* d_idx, int64_t n, uint32_t* d_obs_tm1, int32_t* d_act,
                 float* d_rew, uint32_t* d_obs_t, uint32_t* d_lms, uint8_t* d_term, void* stream);
int hb_rb_check(hb_replay* rb, void* stream);
/* sample_batch after the keys are drawn (priority_buffer.py:41-42): indices = get_indices(keys);
 * prios = (get_values(indices) + 1e-10) / get_total_val() in float64 (+ the float32 narrowing the
 * jitted learner sees).  Keys are explicit so that "identical uniforms" is well defined. */
int hb_rb_sample(hb_replay* rb, const float* d_keys, int64_t n, int64_t* d_idx, double* d_prios64,
                 float* d_prios32, void* stream);
/* sample_batch + the learner's importance weights in ONE launch (B <= 1024): indices = get_indices(keys)
 * (priority_buffer.py:41), prios = (get_values + 1e-10) / get_total_val in float64 (:42; d_prios32 = the
 * float32 narrowing), weights_is = (1 / prios32)^beta / max (rlax_rainbow.py:188-189) and self[indices]
 * (:46) gathered into the caller's buffers (any output but d_idx may be NULL).  d_keys may be pinned host
 * memory (read in place).  d_dyn_in (optional, 8 floats, may be pinned host memory): [0] replaces `beta`,
 * all 8 are copied to d_dyn_out (device) for the kernels that follow.  pdl != 0: programmatic dependent
 * launch behind the previous kernel of the stream.  hb_rb_check reports a leaf outside the ring. */
int hb_rb_sample_gather(hb_replay* rb, const float* d_keys, int B, const float* d_dyn_in, float* d_dyn_out,
                        float beta, int64_t* d_idx, float* d_prios32, double* d_prios64, float* d_w_is,
                        uint32_t* d_obs_tm1, int32_t* d_act, float* d_rew, uint32_t* d_obs_t,
                        uint32_t* d_lms, uint8_t* d_term, int pdl, void* stream);
/* update_priorities (priority_buffer.py:48-52): p = (|td| + 1e-10) ** alpha in float32 (power
 * evaluated in double and rounded once), running max / min, ordered sum-tree scatter. */
int hb_rb_update_priorities(hb_replay* rb, const int64_t* d_idx, const float* d_td_abs, int64_t n,
                            void* stream);
int hb_rb_priority_range(hb_replay* rb, float* h_max, float* h_min, void* stream);
/* Priority of newly inserted rows: value > 0 = that constant (DEFAULT_PRIORITY = 100 of
 * rainbow/prioritized_replay_memory.py:38,95-103); value <= 0 = the running max_priority
 * (priority_buffer.py:29-31, the default). */
int hb_rb_set_insert_priority(hb_replay* rb, float value);
/* One int8 per bit <-> bit-packed rows (the reference stores np.byte rows, experience_buffer.py:9-12). */
int hb_pack_bits(const int8_t* d_in, int bits, uint32_t* d_out, int stride_words, int64_t n, void* stream);
int hb_unpack_bits(const uint32_t* d_packed, int stride_words, int bits, int8_t* d_out, int64_t n,
                   void* stream);

/* ------------------------------------------------------------------------ */
/* Q-network: NoisyMLP / MLP forward and backward, policy, learner step.      */
/* Replaces jnp.dot in NoisyLinear (hanabi_agents/rlax_dqn/noisy_mlp.py:62,80),*/
/* DQNPolicy.policy / eval_policy (rlax_rainbow.py:93-150), DQNLearning.       */
/* update_q (rlax_rainbow.py:154-217, rlax_dqn.py:144-233) and optix.adam      */
/* (rlax_rainbow.py:257).                                                      */
/* ------------------------------------------------------------------------ */
/* One tcgen05/TMA GEMM tile launch (tests, microbenchmarks):
 * D = epilogue(alpha * A . B^T + bias); A [M, K], B [N, K] fp16 row-major DEVICE matrices.
 * mode 0: fp16 out [M, ld_out] (+ReLU when relu != 0, + transposed copy d_out_t [N, ld_out_t]);
 * mode 1: fp32 out; mode 2: categorical head -- columns are [action][64] (n_atoms used),
 * d_out = fp32 logits (may be NULL), d_q[M, N/64] = q_scale * sum_j softmax_j * (z0 + j*dz). */
int hb_gemm_f16(const void* d_a, int64_t lda, const void* d_b, int64_t ldb, int M, int N, int K,
                int mode, float alpha, const float* d_bias, int relu, void* d_out, int64_t ld_out,
                void* d_out_t, int64_t ld_out_t, float* d_q, int n_atoms, float z0, float dz,
                float q_scale, void* stream);

typedef struct hb_qnet hb_qnet;

/* Network of DQNAgent.__init__ (rlax_rainbow.py:237-246): NoisyMLP(layers + [A * n_atoms]) reshaped
 * to [A, n_atoms] (noisy != 0), or hk.nets.MLP(layers + [A]) of rlax_dqn.py:26-33 (noisy == 0,
 * n_atoms == 1).  Input = bit-packed observation rows of obs_stride_words 32-bit words. */
int hb_net_create(int obs_len, int obs_stride_words, const int32_t* hidden, int n_hidden,
                  int num_actions, int n_atoms, float atom_vmax, int noisy, int device,
                  hb_qnet** out);
int hb_net_destroy(hb_qnet* q);
/* what: 0 n_layers, 1 length of the flat gradient bucket, 2 num_actions, 3 n_atoms, 4 noisy,
 *       5 Adam step count, 10+l / 20+l: in / out of layer l, 30+l / 40+l: padded K / N */
int64_t hb_net_info(const hb_qnet* q, int what);
/* Device pointer of the flat fp32 gradient bucket (dW_eff, db_eff per layer): the one buffer a
 * data-parallel job all-reduces between hb_net_grads and hb_net_adam. */
float* hb_net_grad(hb_qnet* q);
/* Parameter exchange in the reference's layout (haiku: w [in, out] row-major, b [out];
 * noisy_mlp.py:61-84).  kind: 0 w, 1 b, 2 w_mu, 3 b_mu, 4 w_sigma, 5 b_sigma, 6 eps_w, 7 eps_b
 * (the trace-time noise constants, noisy_mlp.py:79,87), 8 dW_eff, 9 db_eff (get only). HOST arrays. */
int hb_net_set_param(hb_qnet* q, int layer, int kind, const float* h_data, int64_t n);
int hb_net_get_param(hb_qnet* q, int layer, int kind, float* h_out, int64_t n);
/* The same for trg_params (the snapshot taken by hb_net_sync_target); kind 0..5. */
int hb_net_get_target_param(hb_qnet* q, int layer, int kind, float* h_out, int64_t n);
/* Recompute the fp16 effective weights after hb_net_set_param. */
int hb_net_refresh(hb_qnet* q, void* stream);
/* trg_params = online_params  (rlax_rainbow.py:336-337) */
int hb_net_sync_target(hb_qnet* q, void* stream);
/* DQNAgent.explore (greedy == 0, DQNPolicy.policy rlax_rainbow.py:93-123) / exploit (greedy != 0,
 * eval_policy :125-150) on n bit-packed observation rows + legal masks.  d_uniform: one draw per row
 * (explore: U[0, 0.999)), or NULL = counter-based stream keyed (seed, counter, row0 + row).
 * d_q (optional) [n, num_actions]: q-values with illegal moves at -inf. */
/* Scheduling knob, no counterpart in the reference: the acting pass's persistent GEMMs (hb_net_act) leave
 * `sms` SMs unoccupied for work the caller overlaps on another stream (the self-play session's replay
 * inserts); 0 restores one CTA per SM. */
int hb_net_set_act_reserve(hb_qnet* q, int sms);
int hb_net_act(hb_qnet* q, const uint32_t* d_obs, const uint32_t* d_legal, int legal_words, int64_t n,
               int greedy, float epsilon, const float* d_uniform, uint64_t seed, uint64_t counter,
               int64_t row0, int use_target, int32_t* d_action, float* d_q, void* stream);
/* network.apply on n <= 131072 packed rows: d_q [n, num_actions] (unmasked), d_logits (optional)
 * [n, A*64] with atom j of action a at column a*64 + j. */
int hb_net_forward(hb_qnet* q, const uint32_t* d_obs, int64_t n, int use_target, float* d_q,
                   float* d_logits, void* stream);
/* jax.grad(loss) of DQNLearning.update_q: categorical double-Q (rlax_rainbow.py:172-200) when
 * n_atoms > 1, double-Q with L2 loss (rlax_dqn.py:170-205) when n_atoms == 1.  Gradients are left
 * in the flat bucket; d_loss_out[B] = per-sample loss (C51) / TD error (double-Q) -- what the agent
 * feeds to update_priorities; d_mean_loss[1] = mean(loss * w).  d_prios: sampled probabilities
 * (w = (1/prios)^beta / max); d_weights: importance weights given directly; both NULL = ones. */
int hb_net_grads(hb_qnet* q, const uint32_t* d_obs_tm1, const int32_t* d_a_tm1, const float* d_r_t,
                 const uint32_t* d_obs_t, const uint8_t* d_term_t, const float* d_prios,
                 const float* d_weights, int B, float beta, float discount, int mask_terminal,
                 float* d_loss_out, float* d_mean_loss, void* stream);
/* The learner's form of hb_net_grads.  What a captured, replayed update needs on top of it:
 *   d_w_is         importance weights already on the device (hb_rb_sample_gather computed them with the
 *                  sample; hb_net_is_weights(q, B) is the network's own buffer: pass it to skip a copy)
 *   d_dyn          the step's scalars in DEVICE memory -- [0] beta [1] discount [2] lr [3] b1 [4] b2
 *                  [5] eps [6] grad_scale -- read by the kernels instead of the by-value arguments, which
 *                  a replayed CUDA graph freezes (beta_is / learning_rate are schedules, params.py:5-20)
 *   pdl            chain the kernels of the critical path with programmatic dependent launch
 *   adam_per_layer apply optix.adam (rlax_rainbow.py:215-216) to every layer as soon as its gradient is
 *                  complete, beside the rest of the backward pass (single-process learner: no all-reduce
 *                  between gradient and update); lr, b1, b2, eps, grad_scale as in hb_net_adam */
struct hb_comm;
typedef struct hb_net_step_opts {
  const float* d_w_is;
  const float* d_dyn;
  int pdl;
  int adam_per_layer;
  float lr, b1, b2, eps, grad_scale;
  /* data parallel (with adam_per_layer): every layer's slice of the gradient bucket is all-reduced over
   * this communicator (hb_comm_allreduce_range, channel = layer) between its gradient and its Adam
   * update, beside the rest of the backward pass; grad_scale = 1 / world.  NULL: single process. */
  struct hb_comm* comm;
} hb_net_step_opts;
int hb_net_grads_ex(hb_qnet* q, const uint32_t* d_obs_tm1, const int32_t* d_a_tm1, const float* d_r_t,
                    const uint32_t* d_obs_t, const uint8_t* d_term_t, const float* d_prios,
                    const float* d_weights, int B, float beta, float discount, int mask_terminal,
                    float* d_loss_out, float* d_mean_loss, const hb_net_step_opts* opts, void* stream);
/* The network's importance-weight buffer for a batch of B (valid until a call with a larger B). */
float* hb_net_is_weights(hb_qnet* q, int B);
/* hb_net_adam with the hyper-parameters read from d_dyn ([2] lr [3] b1 [4] b2 [5] eps [6] grad_scale;
 * device or pinned host memory) at execution time. */
int hb_net_adam_dyn(hb_qnet* q, const float* d_dyn, void* stream);
/* cudaEvent_t recorded on the caller's stream by the last hb_net_grads / hb_net_grads_rainbow call at the
 * point where the per-sample losses (d_loss_out) are complete: work that needs nothing but the losses --
 * PriorityBuffer.update_priorities, hanabi_agents/rlax_dqn/rlax_rainbow.py:331-334 -- can wait on it and
 * run beside the backward pass. */
/* The network's own staging for a batch of B: packed observation rows [2B, obs_stride] (obs_tm1 rows, then
 * obs_t rows from row B on), per-sample loss [B], mean loss [1].  hb_net_grads called with exactly these
 * pointers (d_obs_tm1 = staging, d_obs_t = staging + B * obs_stride, d_loss_out, d_mean_loss) skips its
 * device-to-device copies.  The pointers stay valid until a call with a larger B. */
int hb_net_learner_buffers(hb_qnet* q, int B, uint32_t** obs_staging, float** loss, float** mean_loss);
void* hb_net_loss_event(hb_qnet* q);
/* Diagnostics: the host-side handle (what cudaGraphKernelNodeGetParams reports as `func`) of the update's
 * programmatically launched kernels -- 0 = the chained-GEMM kernel, 1 / 2 = the C51 loss with / without the
 * head's backward product; NULL otherwise.  Used by the learner's HB_GRAPH_DOT dump to name graph nodes. */
const void* hb_net_kernel_handle(int which);
/* The same with the TF-Rainbow target of rainbow/rainbow_agent.py:183-236 (the `rainbow/` package's
 * learner): next action = argmax over the LEGAL next moves (d_legal_t, packed) of the
 * TARGET network's q; gamma^n masked by d_term_t (:191-197); d_weights = the caller's importance weights
 * (1/sqrt(priority + 1e-10) / max, :237-240).  Categorical networks only. */
int hb_net_grads_rainbow(hb_qnet* q, const uint32_t* d_obs_tm1, const int32_t* d_a_tm1, const float* d_r_t,
                         const uint32_t* d_obs_t, const uint32_t* d_legal_t, const uint8_t* d_term_t,
                         const float* d_weights, int B, float cumulative_gamma, float* d_loss_out,
                         float* d_mean_loss, void* stream);
/* tf.train.AdamOptimizer(learning_rate, epsilon) (rainbow_agent.py:228-230): lr_t = lr sqrt(1-b2^t)/(1-b1^t),
 * theta -= lr_t m / (sqrt(v) + eps). */
int hb_net_adam_tf(hb_qnet* q, float lr, float b1, float b2, float eps, float grad_scale, void* stream);
/* optix.adam(lr, b1, b2, eps) on every master tensor (rlax_rainbow.py:215-216,257) from
 * grad_scale * bucket, then refresh of the effective weights. */
int hb_net_adam(hb_qnet* q, float lr, float b1, float b2, float eps, float grad_scale, void* stream);

/* ------------------------------------------------------------------------ */
/* One learner step = DQNAgent.update (rlax_rainbow.py:310-339): one C call for   */
/* a single process (hb_learner_step), or two around the caller's gradient       */
/* all-reduce (hb_learner_grads / hb_learner_apply); replayed CUDA graphs.  The   */
/* scalars that are schedules in the reference (beta_is, learning_rate) are read  */
/* from staging memory at run time: changing them does not capture a new graph.   */
/* ------------------------------------------------------------------------ */
typedef struct hb_learner hb_learner;
int hb_learner_create(hb_qnet* net, hb_replay* rb, int batch, int device, hb_learner** out);
int hb_learner_destroy(hb_learner* L);
int hb_learner_set_graph(hb_learner* L, int on);
/* what: 0 batch, 1 graph captures so far, 2 eager steps run */
int64_t hb_learner_info(const hb_learner* L, int what);
/* Device views of the last step: per-sample loss / TD [batch], mean loss [1], sampled rows [batch]. */
float* hb_learner_loss(hb_learner* L);
float* hb_learner_mean_loss(hb_learner* L);
int64_t* hb_learner_indices(hb_learner* L);
/* Phase A: sample_batch (priority_buffer.py:36-46; h_keys = the B stratified keys, float32 as
 * pybind narrows them) or uniform rows (h_idx) -> __getitem__ -> jax.grad(loss); update_priorities(indices, |loss|)
 * (rlax_rainbow.py:331-334) runs here as well, beside the backward pass, as soon as the losses exist
 * (HB_LEARNER_PRIO_OVERLAP=0 in the environment: in phase B, after Adam).  HOST arrays. */
int hb_learner_grads(hb_learner* L, const float* h_keys, const int64_t* h_idx, float beta, float discount,
                     int mask_terminal, void* stream);
/* Phase B: optix.adam on grad_scale * bucket. */
/* Both phases in ONE call and one graph: optix.adam is applied layer by layer beside the backward pass.
 * With a communicator attached (hb_learner_set_comm; the learner's one exchange step) every
 * layer's gradient is all-reduced over NVLink on its own channel between its GEMM and its Adam update,
 * overlapping the rest of the backward pass, and scaled by 1 / world.  Ranks call it in lockstep. */
int hb_learner_set_comm(hb_learner* L, struct hb_comm* comm, int world);
int hb_learner_step(hb_learner* L, const float* h_keys, const int64_t* h_idx, float beta, float discount,
                    int mask_terminal, float lr, float b1, float b2, float eps, void* stream);
int hb_learner_apply(hb_learner* L, float lr, float b1, float b2, float eps, float grad_scale,
                     void* stream);

/* ------------------------------------------------------------------------ */
/* Self-play session bookkeeping (the per-player accounting of the driver loop  */
/* hanabi_agents/rainbow/run_experiment.py:280-336, over N lockstep games).     */
/* Seat k = tick mod players moves in every game; a finished game spends one    */
/* tick on the env's reset step, so all games stay on the same rotation.        */
/* ------------------------------------------------------------------------ */
typedef struct hb_session hb_session;
int hb_session_create(int players, int64_t n_games, int device, hb_session** out);
int hb_session_destroy(hb_session* s);
/* obs_stacker.reset_stack / has_played = {} / reward_since_last_action = 0  (:292-299);
 * clear_stats also zeroes the episode counters. */
int hb_session_reset(hb_session* s, int clear_stats, void* stream);
/* Before seat `agent` acts on the env's current outputs (d_game_step_type [N], dm_env coding):
 * fills the step type every seat is shown this tick --
 *   game LAST            -> LAST for every seat that has played this episode (end_episode :335)
 *   seat to move         -> MID if it has played (agent.step :319), else FIRST (begin_episode :324)
 *   anything else        -> FIRST (= nothing to record)
 * The matching actions / rewards are the seat's rows of buffers 1 / 0 below. */
int hb_session_begin_turn(hb_session* s, int agent, const uint8_t* d_game_step_type, void* stream);
/* After env.step(d_action) returned d_reward: the mover's accumulator restarts at this reward,
 * every other seat's grows by it (:305-310,331); episode return / length advance (:307,312).
 * Games the env reset in this tick close their books instead (statistics, all seats cleared).
 * lenient != 0: reward = max(reward, 0)  (LENIENT_SCORE :306). */
int hb_session_end_turn(hb_session* s, int agent, const int32_t* d_action, const float* d_reward,
                        int lenient, void* stream);
/* Device buffers.  Per seat (row `agent`): 0 reward since last action f32[N], 1 last action i32[N],
 * 2 step type shown u8[N], 3 has-played u8[N].  Per game: 4 running return f32, 5 running length
 * i32, 6 finished episodes i32, 7 / 8 return f32 / length i32 of the first finished episode,
 * 10 game step type before this tick's env step u8.  9: f64[4] sum of returns, episodes, sum of
 * lengths, sum of squared returns. */
void* hb_session_buffer(hb_session* s, int which, int agent);
/* Synchronising copy of buffer 9 to the host. */
int hb_session_stats(hb_session* s, double* h_out4, void* stream);

/* ------------------------------------------------------------------------ */
/* The learner's one exchange step: in-place sum of the flat fp32   */
/* gradient bucket over the GPUs of one node, as a two-shot all-reduce kernel   */
/* over NVLink peer memory (CUDA IPC, one process per GPU).  Replaces nothing   */
/* in the reference (it is single-process); stands where a data-parallel JAX     */
/* learner would call lax.psum on the gradients of rlax_rainbow.py:203-213.      */
/* ------------------------------------------------------------------------ */
typedef struct hb_comm hb_comm;
int hb_comm_create(int rank, int world, int device, float* d_bucket, int64_t n, hb_comm** out);
int hb_comm_destroy(hb_comm* c);
/* 128 bytes of IPC handles (bucket, flags); the caller all-gathers them in rank order ... */
int hb_comm_export(hb_comm* c, void* h_out128);
/* ... and maps every peer from the gathered world x 128 bytes. */
int hb_comm_open(hb_comm* c, const void* h_all);
/* Stream-ordered all-reduce; every rank calls it once per step, in lockstep (as with any
 * collective).  The sum runs over the ranks in fixed order, so every rank ends with the same bits.
 * A rank that waits for a late peer keeps polling (minutes, never a trap); should the bound pass, the
 * kernel ends, hb_comm_status and later hb_comm_allreduce calls return HB_ERR_STATE. */
int hb_comm_allreduce(hb_comm* c, void* stream);
int hb_comm_status(const hb_comm* c);
/* The same on bucket[offset, offset + n) (multiples of 4 elements) over `channel` 0..7.  Calls of one channel
 * must be stream-ordered on a rank and made in the same order on every rank; different channels have their
 * own barrier flags and may overlap -- the learner reduces every layer's gradient on that layer's stream
 * beside the backward pass.  Nothing that changes from call to call is a kernel argument (the call counter
 * lives in device memory), so the call can be captured into a CUDA graph and replayed. */
int hb_comm_allreduce_range(hb_comm* c, int channel, int64_t offset, int64_t n, void* stream);
/* what: 0 rank, 1 world, 2 peers mapped, 3 calls so far */
int hb_comm_info(const hb_comm* c, int what);

#ifdef __cplusplus
}
#endif
#endif /* HANABI_B200_H_ */
