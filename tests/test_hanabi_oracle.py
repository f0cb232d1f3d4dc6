"""CPU tests pinning oracle/hanabi_oracle.c as far as it can be pinned here.

The reference holds no golden vector for the Hanabi engine (parity unpinned), so:
  * upstream shape constants (obs 658/956/1041/1280, moves 20/30/38/48, Small 171/11),
  * a hand-worked opening (every bit position derived from SURVEY Appendix A.5 by hand),
  * the A.6 invariants on random play-outs,
  * a bit-for-bit diff against the independent Python restatement tests/hle_py.py.
"""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(__file__))
import hle_py  # noqa: E402
from oracle.hanabi import HanabiOracleEnv, game_params, make_decks, card_multiset  # noqa: E402

SHAPES = [
    ("Hanabi-Full", 2, 658, 20), ("Hanabi-Full", 3, 956, 30), ("Hanabi-Full", 4, 1041, 38),
    ("Hanabi-Full", 5, 1280, 48), ("Hanabi-Small", 2, 171, 11),
]


@pytest.mark.parametrize("name,players,obs_len,num_moves", SHAPES)
def test_shape_constants(name, players, obs_len, num_moves):
    e = HanabiOracleEnv(1, **game_params(name, players))
    assert (e.obs_len, e.num_moves) == (obs_len, num_moves)
    g = e.game
    assert g.hands_len + g.board_len + g.discard_len + g.last_action_len + g.knowledge_len == obs_len


def test_section_lengths_full_2p():  # 127+76+50+55+350
    g = HanabiOracleEnv(1, **game_params("Hanabi-Full", 2)).game
    assert (g.hands_len, g.board_len, g.discard_len, g.last_action_len, g.knowledge_len) == \
        (127, 76, 50, 55, 350)
    g = HanabiOracleEnv(1, **game_params("Hanabi-Full", 5)).game
    assert (g.hands_len, g.board_len, g.discard_len, g.last_action_len, g.knowledge_len) == \
        (405, 66, 50, 59, 700)
    g = HanabiOracleEnv(1, **game_params("Hanabi-Small", 2)).game
    assert (g.hands_len, g.board_len, g.discard_len, g.last_action_len, g.knowledge_len) == \
        (22, 30, 20, 31, 68)


def test_decks_must_be_permutations():
    e = HanabiOracleEnv(2, **game_params("Hanabi-Full", 2))
    bad = np.zeros((1, 50), np.uint8)
    with pytest.raises(ValueError):
        e.set_decks(bad)


def _sorted_deck():
    return card_multiset(5, 5)[None, :].copy()  # R1 R1 R1 R2 R2 R3 R3 R4 R4 R5 Y1 ...


def test_hand_worked_opening_full_2p():
    """Deck in multiset order: P0 = [R1,R1,R1,R2,R2] (cards 0,0,0,1,1), P1 = [R3,R3,R4,R4,R5]."""
    e = HanabiOracleEnv(1, **game_params("Hanabi-Full", 2))
    e.set_decks(_sorted_deck())
    obs, legal, st = e.reset()
    o = obs[0]
    assert st[0] == 0
    # hands: partner (offset 1) cards 2,2,3,3,4 one-hot in 25-bit slots; no short-hand bits
    exp = np.zeros(658, np.uint8)
    for i, c in enumerate([2, 2, 3, 3, 4]):
        exp[i * 25 + c] = 1
    # board: deck thermometer 40 of 40, no fireworks, 8 info, 3 life
    exp[127:127 + 40] = 1
    exp[127 + 40 + 25:127 + 40 + 25 + 8] = 1
    exp[127 + 40 + 25 + 8:127 + 76] = 1
    # discards, last action: empty.  knowledge: all 10 cards fully plausible, nothing hinted
    k0 = 127 + 76 + 50 + 55
    for slot in range(10):
        exp[k0 + slot * 35:k0 + slot * 35 + 25] = 1
    np.testing.assert_array_equal(o, exp)
    # legal: discards illegal at 8 tokens; plays legal; colour R only; ranks 3,4,5 (idx 2,3,4)
    exp_legal = np.zeros(20, np.uint8)
    exp_legal[5:10] = 1
    exp_legal[10 + 0] = 1
    exp_legal[15 + 2:15 + 5] = 1
    np.testing.assert_array_equal(legal[0], exp_legal)

    # P0 reveals rank 3 (uid 15+2) to P1: cards 0,1 match -> mask 0b00011
    obs, legal, rew, st = e.step([17])
    o = obs[0]
    assert rew[0] == 0 and st[0] == 1
    la = 127 + 76 + 50
    # observer is now P1; acting player relative = 1; type RevealRank = index 3; target = (1+1)%2=0
    exp_la = np.zeros(55, np.uint8)
    exp_la[1] = 1
    exp_la[2 + 3] = 1
    exp_la[6 + 0] = 1
    exp_la[8 + 5 + 2] = 1          # rank 3 (index 2) after P+4+P+C
    exp_la[18 + 0] = exp_la[18 + 1] = 1  # reveal bitmask
    np.testing.assert_array_equal(o[la:la + 55], exp_la)
    # info thermometer now 7
    assert o[127 + 40 + 25:127 + 40 + 25 + 8].tolist() == [1] * 7 + [0]
    # P1's own knowledge (offset 0): cards 0,1 rank plausible only 3 + rank-hinted bit; 2,3,4 not rank 3
    kn = o[k0:k0 + 175].reshape(5, 35)
    for i in (0, 1):
        grid = kn[i, :25].reshape(5, 5)
        assert grid[:, 2].all() and grid.sum() == 5
        assert kn[i, 25:30].sum() == 0 and kn[i, 30:35].tolist() == [0, 0, 1, 0, 0]
    for i in (2, 3, 4):
        grid = kn[i, :25].reshape(5, 5)
        assert not grid[:, 2].any() and grid.sum() == 20 and kn[i, 25:].sum() == 0
    # P1 sees P0's hand (cards 0,0,0,1,1)
    assert [int(np.argmax(o[i * 25:(i + 1) * 25])) for i in range(5)] == [0, 0, 0, 1, 1]

    # P1 plays card 0 (R3) on an empty red stack: misfire, life 2, card to discards, draws Y1? no:
    # deck continues at index 10 = Y1 (card 5)
    obs, legal, rew, st = e.step([5 + 0])
    o = obs[0]
    assert rew[0] == 0 and st[0] == 1
    assert o[127 + 40 + 25 + 8:127 + 76].tolist() == [1, 1, 0]           # life 2
    assert o[127:127 + 40].sum() == 39                                     # one card drawn
    disc = o[203:253]
    # R3 is card type 2: offset = 3 (R1) + 2 (R2) = 5, thermometer length 2
    assert disc[5] == 1 and disc.sum() == 1
    exp_la = np.zeros(55, np.uint8)
    exp_la[1] = 1                   # P1 relative to observer P0
    exp_la[2 + 0] = 1               # Play
    exp_la[23 + 0] = 1              # card position 0 (after P+4+P+C+R+H = 23)
    exp_la[28 + 2] = 1              # card identity R3 (after ... + H = 28)
    np.testing.assert_array_equal(o[la:la + 55], exp_la)   # scored=0, info_token=0
    # P0 now sees P1's hand shifted: [R3,R4,R4,R5,Y1]
    assert [int(np.argmax(o[i * 25:(i + 1) * 25])) for i in range(5)] == [2, 3, 3, 4, 5]
    d = e.dump()[0]
    assert d[:7].tolist() == [0, 7, 2, 39, 2, 1, 0]


def _to_tuple_deck(deck, ranks):
    return [(int(c) // ranks, int(c) % ranks) for c in deck]


@pytest.mark.parametrize("name,players,n_games", [
    ("Hanabi-Full", 2, 6), ("Hanabi-Full", 3, 3), ("Hanabi-Full", 4, 3), ("Hanabi-Full", 5, 4),
    ("Hanabi-Small", 2, 8), ("Hanabi-Very-Small", 2, 4),
])
def test_diff_against_python_restatement(name, players, n_games):
    params = game_params(name, players)
    rng = np.random.default_rng(hash((name, players)) % 2**32)
    e = HanabiOracleEnv(n_games, **params)
    decks = make_decks(params["colors"], params["ranks"], n_games, seed=7 + players)
    e.set_decks(decks)
    g = hle_py.Game(**params)
    states = [hle_py.State(g, _to_tuple_deck(decks[i], params["ranks"])) for i in range(n_games)]
    obs, legal, st = e.reset()
    done = [False] * n_games
    for step in range(200):
        actions = np.zeros(n_games, np.int32)
        for i in range(n_games):
            if done[i]:
                continue
            s = states[i]
            assert obs[i].tolist() == hle_py.encode(hle_py.Observation(s, s.cur_player)), (i, step)
            assert legal[i].tolist() == s.legal_mask(s.cur_player), (i, step)
            # other seats' views (incl. the player who just moved) through observe()
        for off in range(1, players):
            o2, l2 = e.observe(off)
            for i in range(n_games):
                if not done[i]:
                    s = states[i]
                    pl = (s.cur_player + off) % players
                    assert o2[i].tolist() == hle_py.encode(hle_py.Observation(s, pl))
                    assert not l2[i].any()
        for i in range(n_games):
            if not done[i]:
                actions[i] = rng.choice(np.flatnonzero(legal[i]))
        if all(done):
            break
        prev_done = list(done)
        obs, legal, rew, st = e.step(actions)
        obs, legal = obs.copy(), legal.copy()
        for i in range(n_games):
            if prev_done[i]:
                continue
            r, term = states[i].step(int(actions[i]))
            assert rew[i] == r and (st[i] == 2) == term
            if term:
                s = states[i]
                assert obs[i].tolist() == hle_py.encode(hle_py.Observation(s, s.cur_player))
                done[i] = True
    assert all(done)


@pytest.mark.parametrize("name,players", [("Hanabi-Full", 2), ("Hanabi-Full", 5), ("Hanabi-Small", 2)])
def test_invariants_random_playouts(name, players):
    params = game_params(name, players)
    n = 256
    e = HanabiOracleEnv(n, **params)
    e.set_decks(make_decks(params["colors"], params["ranks"], 4 * n, seed=3))
    rng = np.random.default_rng(1)
    obs, legal, st = e.reset()
    g = e.game
    C, R, P, H = g.colors, g.ranks, g.players, g.hand_size
    counts = np.array(g.card_count[:C * R])
    prev_score = np.zeros(n)
    moves_in_game = np.zeros(n, int)
    for step in range(400):
        d = e.dump()
        # card conservation: deck + hands + discards + fireworks == deck size
        hand_n = d[:, 8 + C:8 + C + P].sum(1)
        fw = d[:, 8:8 + C].sum(1)
        disc = d[:, 8 + C + P + 5 * P * H:8 + C + P + 5 * P * H + C * R]
        assert np.all(disc <= counts)
        assert np.all(d[:, 3] + hand_n + disc.sum(1) + fw == g.deck_size)
        assert np.all((d[:, 1] >= 0) & (d[:, 1] <= g.max_info))
        o = obs
        b0 = g.hands_len
        assert np.all(o[:, b0:b0 + g.deck_size - P * H].sum(1) == d[:, 3])      # deck thermometer
        d0 = g.hands_len + g.board_len
        assert np.all(o[:, d0:d0 + g.deck_size].sum(1) == disc.sum(1))           # discards popcount
        la = d0 + g.discard_len
        mt = o[:, la + P:la + P + 4].sum(1)
        assert np.all(mt == (d[:, -12] == 1))                                     # one move-type bit
        live = st != 2
        assert np.all(legal[live].sum(1) > 0) and np.all(legal[live][:, H:2 * H].sum(1) > 0)
        a = np.zeros(n, np.int32)
        for i in range(n):
            nz = np.flatnonzero(legal[i])
            a[i] = rng.choice(nz) if len(nz) else 0
        was_last = st == 2
        obs, legal, rew, st = e.step(a)
        assert e.illegal == 0
        assert np.all(st[was_last] == 0) and np.all(rew[was_last] == 0)
        sc = e.dump()[:, 6]
        # score is non-decreasing until life-out; reward is the score delta
        mid = ~was_last
        np.testing.assert_array_equal(rew[mid], (sc - prev_score)[mid])
        life_out = e.dump()[:, 2] < 1
        assert np.all((rew[mid & ~life_out] >= 0))
        prev_score = np.where(st == 0, 0, sc)
        moves_in_game = np.where(st == 0, 0, moves_in_game + 1)
        assert moves_in_game.max() <= 2 * g.deck_size + 40


def test_illegal_action_is_counted_and_ignored():
    e = HanabiOracleEnv(1, **game_params("Hanabi-Full", 2))
    e.set_decks(_sorted_deck())
    e.reset()
    before = e.dump().copy()
    e.step([0])  # discard at 8 info tokens is illegal
    assert e.illegal == 1
    np.testing.assert_array_equal(e.dump(), before)
    e.step([99])
    assert e.illegal == 1


# ---- the three rules that crash in the reference (dict-style access on pybind objects), restated
# ---- with their intended semantics: the C oracle against a literal Python reading of ruleset.py
def _py_rule_view(dump_row, players, colors, ranks, hand_size):
    """Unpacks one ho_env_dump row into the dict-style observation those rules index."""
    d = list(dump_row)
    cur, info = d[0], d[1]
    k = 8
    fireworks = d[k:k + colors]; k += colors
    hand_n = d[k:k + players]; k += players
    hands = []
    for p in range(players):
        cards = []
        for h in range(hand_size):
            card, cp, rp, ch, rh = d[k:k + 5]; k += 5
            if h < hand_n[p]:
                cards.append(dict(color=card // ranks, rank=card % ranks, cp=cp, rp=rp, ch=ch, rh=rh))
        hands.append(cards)
    discards = d[k:k + colors * ranks]
    # observed_hands[offset]: offset 0 = me
    observed = [hands[(cur + o) % players] for o in range(players)]
    return dict(info=info, fireworks=fireworks, observed=observed, discards=discards,
                colors=colors, ranks=ranks, hand_size=hand_size, players=players)


_NUM_IN_DECK = [3, 2, 2, 2, 1]          # ruleset.py:9


def _py_max_fireworks(v):               # ruleset.py:174-193
    out = [5] * v["colors"]
    for c in range(v["colors"]):
        for r in range(v["ranks"]):
            n = v["discards"][c * v["ranks"] + r]
            if n > 0 and n >= _NUM_IN_DECK[r] and out[c] >= r:
                out[c] = r
    return out


def _py_useless(card_c, card_r, fireworks, maxfw):   # ruleset.py:35-42
    return card_r < fireworks[card_c] or card_r >= maxfw[card_c]


def _uid(v, kind, a, b=None):
    H, C, R, P = v["hand_size"], v["colors"], v["ranks"], v["players"]
    if kind == "discard":
        return a
    if kind == "color":                 # (target offset, colour)
        return 2 * H + (a - 1) * C + b
    return 2 * H + (P - 1) * C + (a - 1) * R + b


def _py_tell_anyone_useless_card(v):    # ruleset.py:522-535
    if v["info"] > 1:
        maxfw = _py_max_fireworks(v)
        for off in range(1, v["players"]):
            for card in v["observed"][off]:
                if _py_useless(card["color"], card["rank"], v["fireworks"], maxfw):
                    if card["ch"] < 0:
                        return _uid(v, "color", off, card["color"])
                    if card["rh"] < 0:
                        return _uid(v, "rank", off, card["rank"])
    return None


def _py_tell_playable_card(v):          # ruleset.py:570-594: the two moves the coin chooses between
    if v["info"] > 0:
        for off in range(1, v["players"]):
            for card in v["observed"][off]:
                if card["rank"] == v["fireworks"][card["color"]]:
                    return (_uid(v, "rank", off, card["rank"]), _uid(v, "color", off, card["color"]))
    return None


def _py_discard_probably_useless(v, threshold):   # ruleset.py:638-649 + :127-170
    if not v["info"] < 8:
        return None
    C, R = v["colors"], v["ranks"]
    visible = list(v["discards"])
    for off in range(1, v["players"]):
        for card in v["observed"][off]:
            visible[card["color"] * R + card["rank"]] += 1
    for c in range(C):
        for r in range(v["fireworks"][c]):
            visible[c * R + r] += 1
    maxfw = _py_max_fireworks(v)
    probs = []
    for card in v["observed"][0]:
        total = useless = 0
        for c in range(C):
            for r in range(R):
                if (card["cp"] >> c) & 1 and (card["rp"] >> r) & 1:
                    n = _NUM_IN_DECK[r] - visible[c * R + r]
                    total += n
                    if _py_useless(c, r, v["fireworks"], maxfw):
                        useless += n
        probs.append(useless / total)
    if not probs:
        return None
    idx = int(np.argmax(probs))
    return _uid(v, "discard", idx) if probs[idx] >= threshold else None


@pytest.mark.parametrize("name,players", [("Hanabi-Full", 2), ("Hanabi-Full", 4), ("Hanabi-Small", 2)])
def test_intended_semantics_of_the_dead_rules(name, players):
    from oracle.hanabi import (R_DISCARD_PROBABLY_USELESS, R_TELL_ANYONE_USELESS_CARD, R_TELL_PLAYABLE_CARD,
                               PRESETS)
    p = game_params(name, players)
    n = 96
    env = HanabiOracleEnv(n, **p)
    env.set_decks(make_decks(p["colors"], p["ranks"], 4 * n, 21))
    env.reset()
    fired = [0, 0, 0]
    for step in range(150):
        dump = env.dump()
        views = [_py_rule_view(dump[g], players, p["colors"], p["ranks"], p["hand_size"]) for g in range(n)]
        live = env.step_type != 2
        for j, (rid, farg, py) in enumerate((
                (R_TELL_ANYONE_USELESS_CARD, 0.0, _py_tell_anyone_useless_card),
                (R_TELL_PLAYABLE_CARD, 0.0, _py_tell_playable_card),
                (R_DISCARD_PROBABLY_USELESS, 0.4, lambda v: _py_discard_probably_useless(v, 0.4)))):
            act, hit = env.rules_act([(rid, farg, 0)], 5, step)
            for g in np.flatnonzero(live):
                want = py(views[g])
                if want is None:
                    assert hit[g] == 1, (name, step, g, j)          # fell through to legal_random
                else:
                    assert hit[g] == 0
                    fired[j] += 1
                    if isinstance(want, tuple):
                        assert act[g] in want
                    else:
                        assert act[g] == want
        # play on so that fireworks / discards / hints all develop (piers hard-codes 8 information
        # tokens like the reference's rules, ruleset.py:207,221: on Hanabi-Small use random legal moves)
        from oracle.hanabi import R_LEGAL_RANDOM
        driver = PRESETS["piers"] if p["max_info"] == 8 else [(R_LEGAL_RANDOM, 0, 0)]
        a, _ = env.rules_act(driver, 5, step)
        env.step(a)
        assert env.illegal == 0
    assert all(f > 20 for f in fired), fired


@pytest.mark.parametrize("name,players", [("Hanabi-Full", 2), ("Hanabi-Full", 5), ("Hanabi-Small", 2)])
def test_minimal_and_seer_observation_types(name, players):
    """HLE's other observation types (oracle/hanabi_oracle.c ho_encode_typed): kMinimal drops the card
    knowledge section, kSeer prepends the observer's own cards to the hands section.  Checked against the
    canonical vector and the state dump on random play-outs, from both seats' points of view."""
    p = game_params(name, players)
    n = 64
    env = HanabiOracleEnv(n, threads=2, **p)
    env.set_decks(make_decks(p["colors"], p["ranks"], 4 * n, 3))
    env.reset()
    H, bpc, P = p["hand_size"], p["colors"] * p["ranks"], players
    know = P * H * (bpc + p["colors"] + p["ranks"])
    rng = np.random.default_rng(1)
    if (name, players) == ("Hanabi-Full", 2):
        assert env.observe_typed(0).shape[1] == 658 - 350 and env.observe_typed(2).shape[1] == 658 + 125
    for step in range(60):
        for offset in (0, players - 1):
            canon, _ = env.observe(offset)
            mini, card, seer = env.observe_typed(0, offset), env.observe_typed(1, offset), env.observe_typed(2, offset)
            np.testing.assert_array_equal(card, canon)
            np.testing.assert_array_equal(mini, canon[:, :canon.shape[1] - know])
            np.testing.assert_array_equal(seer[:, H * bpc:], canon)         # everything else moves up
            d = env.dump()
            cur = d[:, 0]
            hands = d[:, 8 + p["colors"] + P:8 + p["colors"] + P + 5 * P * H].reshape(n, P, H, 5)
            own = hands[np.arange(n), (cur + offset) % P, :, 0]             # card index or -1, oldest first
            want = np.zeros((n, H, bpc), np.uint8)
            for i in range(H):
                ok = own[:, i] >= 0
                want[np.nonzero(ok)[0], i, own[ok, i]] = 1
            np.testing.assert_array_equal(seer[:, :H * bpc], want.reshape(n, -1))
        legal = env.legal
        acts = np.array([rng.choice(np.nonzero(legal[g])[0]) for g in range(n)], np.int32)
        env.step(acts)
