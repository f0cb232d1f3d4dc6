"""ctypes wrapper of oracle/hanabi_oracle.c (+ rules_oracle.c) -- TEST INFRASTRUCTURE."""
import ctypes as C

import numpy as np

from . import oracle_lib

vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int

# env ids of the reference (run_experiment.py:115-128, hanabi_rainbow.gin:35) -> parameters (A.1)
CONFIGS = {
    "Hanabi-Full": dict(colors=5, ranks=5, max_info=8, max_life=3),
    "Hanabi-Small": dict(colors=2, ranks=5, hand_size=2, max_info=3, max_life=1),
    "Hanabi-Very-Small": dict(colors=1, ranks=5, hand_size=2, max_info=3, max_life=1),
}


def game_params(name="Hanabi-Full", players=2):
    p = dict(CONFIGS[name])
    p["players"] = players
    if "hand_size" not in p:
        p["hand_size"] = 5 if players <= 3 else 4
    return p


def card_multiset(colors, ranks):
    out = []
    for c in range(colors):
        for r in range(ranks):
            n = 3 if r == 0 else (1 if r == ranks - 1 else 2)
            out += [c * ranks + r] * n
    return np.array(out, dtype=np.uint8)


def make_decks(colors, ranks, n_decks, seed=1234):
    """[n_decks, deck_size] uint8 permutations of the card multiset (SURVEY 8d cfg 2)."""
    base = np.tile(card_multiset(colors, ranks), (n_decks, 1))
    return np.random.default_rng(seed).permuted(base, axis=1)


_decl = False


def _lib():
    global _decl
    L = oracle_lib()
    if not _decl:
        L.ho_env_create.restype = vp
        L.ho_env_create.argtypes = [i32] * 6 + [i64]
        L.ho_env_destroy.argtypes = [vp]
        L.ho_env_game.restype = vp
        L.ho_env_game.argtypes = [vp]
        L.ho_env_set_decks.restype = i32
        L.ho_env_set_decks.argtypes = [vp, vp, i64]
        L.ho_env_set_threads.argtypes = [vp, i32]
        L.ho_env_reset.restype = i32
        L.ho_env_reset.argtypes = [vp, vp, vp, vp]
        L.ho_env_step.restype = i64
        L.ho_env_step.argtypes = [vp, vp, vp, vp, vp, vp]
        L.ho_env_observe.argtypes = [vp, i32, vp, vp]
        L.ho_env_observe_typed.argtypes = [vp, i32, i32, vp]
        L.ho_obs_len_typed.restype = i32
        L.ho_obs_len_typed.argtypes = [vp, i32]
        L.ho_state_dump_len.restype = i32
        L.ho_state_dump_len.argtypes = [vp]
        L.ho_env_dump.argtypes = [vp, vp]
        if hasattr(L, "ro_act"):
            L.ro_act.restype = i32
            L.ro_act.argtypes = [vp, vp, i32, vp, vp, C.c_uint64, C.c_uint64, i64, i64, vp, vp]
        _decl = True
    return L


class _Game(C.Structure):
    _fields_ = [(n, i32) for n in (
        "players", "colors", "ranks", "hand_size", "max_info", "max_life", "deck_size",
        "num_moves", "obs_len", "bits_per_card", "hands_len", "board_len", "discard_len",
        "last_action_len", "knowledge_len")] + [("card_count", i32 * 25)]


class HanabiOracleEnv:
    def __init__(self, n_games, players=2, colors=5, ranks=5, hand_size=5, max_info=8, max_life=3,
                 threads=1):
        self.L = _lib()
        self.h = self.L.ho_env_create(players, colors, ranks, hand_size, max_info, max_life, n_games)
        if not self.h:
            raise ValueError("bad game configuration")
        self.n = n_games
        self.game = _Game.from_address(self.L.ho_env_game(self.h))
        self.obs_len, self.num_moves = self.game.obs_len, self.game.num_moves
        self.deck_size = self.game.deck_size
        self.L.ho_env_set_threads(self.h, threads)
        self.obs = np.zeros((n_games, self.obs_len), np.uint8)
        self.legal = np.zeros((n_games, self.num_moves), np.uint8)
        self.reward = np.zeros(n_games, np.float32)
        self.step_type = np.zeros(n_games, np.uint8)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ho_env_destroy(self.h)
            self.h = None

    def set_decks(self, decks):
        d = np.ascontiguousarray(decks, dtype=np.uint8)
        assert d.ndim == 2 and d.shape[1] == self.deck_size
        if self.L.ho_env_set_decks(self.h, d.ctypes.data, d.shape[0]) != 0:
            raise ValueError("decks are not permutations of the card multiset")

    def reset(self):
        rc = self.L.ho_env_reset(self.h, self.obs.ctypes.data, self.legal.ctypes.data,
                                 self.step_type.ctypes.data)
        if rc != 0:
            raise RuntimeError("set_decks first")
        return self.obs, self.legal, self.step_type

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.int32)
        self.illegal = int(self.L.ho_env_step(self.h, a.ctypes.data, self.obs.ctypes.data,
                                              self.legal.ctypes.data, self.reward.ctypes.data,
                                              self.step_type.ctypes.data))
        return self.obs, self.legal, self.reward, self.step_type

    def observe(self, offset):
        obs = np.zeros_like(self.obs)
        legal = np.zeros_like(self.legal)
        self.L.ho_env_observe(self.h, offset, obs.ctypes.data, legal.ctypes.data)
        return obs, legal

    def observe_typed(self, obs_type, offset=0):
        """HLE observation types: 0 kMinimal, 1 kCardKnowledge, 2 kSeer -> uint8 [N, length of that type]."""
        n = int(self.L.ho_obs_len_typed(self.L.ho_env_game(self.h), obs_type))
        obs = np.zeros((self.n, n), np.uint8)
        self.L.ho_env_observe_typed(self.h, offset, obs_type, obs.ctypes.data)
        return obs

    def dump(self):
        n = self.L.ho_state_dump_len(self.L.ho_env_game(self.h))
        out = np.zeros((self.n, n), np.int32)
        self.L.ho_env_dump(self.h, out.ctypes.data)
        return out

    def rules_act(self, rules, seed, step, game_id_base=0):
        """rules: list of (rule_id, farg, iarg).  Returns (actions int32[N], rule_hit int32[N])."""
        ids = np.array([r[0] for r in rules], np.int32)
        fa = np.array([r[1] for r in rules], np.float64)
        ia = np.array([r[2] for r in rules], np.int32)
        act = np.zeros(self.n, np.int32)
        hit = np.zeros(self.n, np.int32)
        self.L.ro_act(self.h, ids.ctypes.data, len(rules), fa.ctypes.data, ia.ctypes.data,
                      int(seed), int(step), int(game_id_base), self.n, act.ctypes.data,
                      hit.ctypes.data)
        return act, hit


# rule ids (oracle/rules_oracle.c enum); (id, farg, iarg) triples
R_DISCARD_OLDEST_FIRST, R_OSAWA_DISCARD, R_TELL_UNKNOWN, R_TELL_RANDOMLY, R_PLAY_SAFE_CARD, \
    R_PLAY_IF_CERTAIN, R_TELL_PLAYABLE_CARD_OUTER, R_TELL_DISPENSABLE, R_TELL_ANYONE_USEFUL_CARD, \
    R_PLAY_PROBABLY_SAFE, R_HAIL_MARY, R_LEGAL_RANDOM, R_DISCARD_RANDOMLY, \
    R_TELL_MOST_INFORMATION, R_TELL_ANYONE_USELESS_CARD, R_TELL_PLAYABLE_CARD, \
    R_DISCARD_PROBABLY_USELESS = range(17)

# predefined_rules.py:3-39
PRESETS = {
    "flawed": [(R_PLAY_SAFE_CARD, 0, 0), (R_PLAY_PROBABLY_SAFE, 0.25, 0), (R_TELL_RANDOMLY, 0, 0),
               (R_OSAWA_DISCARD, 0, 0), (R_DISCARD_OLDEST_FIRST, 0, 0), (R_DISCARD_RANDOMLY, 0, 0)],
    "iggi": [(R_PLAY_IF_CERTAIN, 0, 0), (R_PLAY_SAFE_CARD, 0, 0), (R_TELL_PLAYABLE_CARD_OUTER, 0, 0),
             (R_OSAWA_DISCARD, 0, 0), (R_DISCARD_OLDEST_FIRST, 0, 0), (R_LEGAL_RANDOM, 0, 0)],
    "outer": [(R_PLAY_SAFE_CARD, 0, 0), (R_OSAWA_DISCARD, 0, 0), (R_TELL_PLAYABLE_CARD_OUTER, 0, 0),
              (R_TELL_UNKNOWN, 0, 0), (R_DISCARD_RANDOMLY, 0, 0)],
    "piers": [(R_HAIL_MARY, 0, 0), (R_PLAY_SAFE_CARD, 0, 0), (R_PLAY_PROBABLY_SAFE, 0.6, 1),
              (R_TELL_ANYONE_USEFUL_CARD, 0, 0), (R_TELL_DISPENSABLE, 0, 3), (R_OSAWA_DISCARD, 0, 0),
              (R_DISCARD_OLDEST_FIRST, 0, 0), (R_TELL_RANDOMLY, 0, 0), (R_DISCARD_RANDOMLY, 0, 0)],
}
