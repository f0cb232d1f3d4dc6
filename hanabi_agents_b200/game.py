"""Game descriptions (HLE environment ids -> parameters) and deck-pool helpers."""
import numpy as np

# env ids used by the reference: hanabi_agents/rainbow/run_experiment.py:115-128,
# configs/hanabi_rainbow.gin:35.  Parameters as HLE's rl_env.make resolves them.
CONFIGS = {
    "Hanabi-Full": dict(colors=5, ranks=5, max_info=8, max_life=3),
    "Hanabi-Full-CardKnowledge": dict(colors=5, ranks=5, max_info=8, max_life=3),
    "Hanabi-Small": dict(colors=2, ranks=5, hand_size=2, max_info=3, max_life=1),
    "Hanabi-Very-Small": dict(colors=1, ranks=5, hand_size=2, max_info=3, max_life=1),
}


def game_params(name="Hanabi-Full", players=2):
    if name not in CONFIGS:
        raise ValueError("unknown environment %r" % (name,))
    if not 2 <= players <= 5:
        raise ValueError("players must be 2..5")
    p = dict(CONFIGS[name])
    p["players"] = players
    if "hand_size" not in p:
        p["hand_size"] = 5 if players <= 3 else 4
    return p


def card_multiset(colors, ranks):
    """Card indices color*ranks+rank with HLE's instance counts (3 / 2 ... 2 / 1 per rank)."""
    out = []
    for c in range(colors):
        for r in range(ranks):
            n = 3 if r == 0 else (1 if r == ranks - 1 else 2)
            out += [c * ranks + r] * n
    return np.array(out, dtype=np.uint8)


def make_decks(colors, ranks, n_decks, seed=1234):
    """[n_decks, deck_size] uint8 permutations of the card multiset."""
    base = np.tile(card_multiset(colors, ranks), (n_decks, 1))
    return np.random.default_rng(seed).permuted(base, axis=1)
