"""CPU: the batched session restatement (oracle/session.py OracleSession) against ONE game driven
literally as run_experiment.py:280-336 drives it (oracle.session.run_one_episode) -- what every
player is told (first observation, reward since its last action, final reward) must agree."""
import numpy as np
import pytest

from oracle.hanabi import HanabiOracleEnv, game_params, make_decks
from oracle.session import OracleSession, run_one_episode
from session_util import FIRST, LAST, MID, DopamineRecorder, OracleRecorder


def _session(name, players, n, lenient=False, seed=3):
    p = game_params(name, players)
    decks = make_decks(p["colors"], p["ranks"], n, seed)
    env = HanabiOracleEnv(n, **p)
    env.set_decks(decks)
    holder = {}
    agents = [OracleRecorder("seat%d" % k, lambda: holder["s"].tick) for k in range(players)]
    ses = OracleSession(env, agents, lenient_score=lenient)
    holder["s"] = ses
    return ses, agents, decks, p


@pytest.mark.parametrize("name,players,lenient", [("Hanabi-Full", 2, False), ("Hanabi-Full", 3, True),
                                                  ("Hanabi-Small", 2, False), ("Hanabi-Full", 5, False)])
def test_first_episode_matches_the_literal_driver(name, players, lenient):
    n = 6
    ses, agents, decks, p = _session(name, players, n, lenient)
    for _ in range(120):
        ses.step(train=True)
    first_closed = {}
    for g, length, ret in ses.episodes:
        first_closed.setdefault(g, (length, ret))
    assert len(first_closed) == n, "every game should finish once in 120 ticks of this policy"
    for g in range(n):
        one = HanabiOracleEnv(1, **p)
        one.set_decks(decks[g:g + 1])
        rec = DopamineRecorder(g)
        steps, total = run_one_episode(one, rec, lenient=lenient)
        assert (steps, total) == first_closed[g]
        for player in range(players):
            # seat k plays player k in the first episode (all games start at tick 0)
            mine = [e for e in agents[player].events if e[1] == g]
            want = rec.per_player.get(player, [])
            got = mine[:len(want)]
            assert len(got) == len(want)
            for (tick, _, st, a, r, obs, _), (wst, wa, wr, wobs) in zip(got, want):
                assert st == wst
                if wst != "seed":
                    assert (a, r) == (wa, wr)
                if wobs is not None:
                    assert obs == wobs


def test_every_seat_gets_a_terminal_transition_and_books_close():
    ses, agents, _, _ = _session("Hanabi-Small", 2, 8)
    for _ in range(150):
        ses.step(train=True)
    assert len(ses.episodes) >= 8
    # a closed episode of `length` moves hands LAST to min(length, P) seats (a seat that never moved
    # has nothing to record -- Hanabi-Small has one life, games of one move happen)
    n_last = sum(1 for a in agents for e in a.events if e[2] == LAST)
    assert n_last == sum(min(length, 2) for _, length, _ in ses.episodes)
    for a in agents:
        # per game the stream is seed (MID)* LAST seed ...
        for g in range(8):
            seq = [e[2] for e in a.events if e[1] == g]
            state = "idle"
            for s in seq:
                if s == "seed":
                    state = "open"       # a re-seed with nothing pending is allowed
                elif s == MID:
                    assert state == "open"
                elif s == LAST:
                    assert state == "open"
                    state = "idle"


def test_eval_mode_records_nothing():
    ses, agents, _, _ = _session("Hanabi-Small", 2, 4)
    for _ in range(40):
        ses.step(train=False)
    assert all(len(a.events) == 0 for a in agents)
    assert len(ses.episodes) > 0
