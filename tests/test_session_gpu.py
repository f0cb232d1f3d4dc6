"""GPU parity of the self-play session (hanabi_agents_b200.session, csrc/session.cu) against
oracle.session.OracleSession on the same decks with the same deterministic policy: every verb
call every seat receives (step types, last actions, rewards since, observations from its own seat,
legal moves) is compared bit for bit, over several episodes per game; then the real agents."""
import numpy as np
import pytest

from session_util import FIRST, LAST, MID, DeviceRecorder, OracleRecorder

pytestmark = pytest.mark.gpu


def _pair(name, players, n, lenient=False, seed=11):
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.game import game_params, make_decks
    from hanabi_agents_b200.session import HanabiParallelSession
    from oracle.hanabi import HanabiOracleEnv
    from oracle.session import OracleSession
    p = game_params(name, players)
    decks = make_decks(p["colors"], p["ranks"], 3 * n + 1, seed)
    env = HanabiParallelEnv(n, name, players, decks=decks)
    oenv = HanabiOracleEnv(n, **p)
    oenv.set_decks(decks)
    hold = {}
    dev_agents = [DeviceRecorder("seat%d" % k, lambda: hold["d"].tick) for k in range(players)]
    ora_agents = [OracleRecorder("seat%d" % k, lambda: hold["o"].tick) for k in range(players)]
    hold["d"] = HanabiParallelSession(env, dev_agents, lenient_score=lenient)
    hold["o"] = OracleSession(oenv, ora_agents, lenient_score=lenient)
    return hold["d"], hold["o"], dev_agents, ora_agents


@pytest.mark.parametrize("name,players,lenient", [("Hanabi-Full", 2, False), ("Hanabi-Full", 3, True),
                                                  ("Hanabi-Small", 2, False), ("Hanabi-Full", 5, False),
                                                  ("Hanabi-Very-Small", 2, True)])
def test_session_matches_oracle(name, players, lenient):
    import torch
    n = 48
    dev, ora, dag, oag = _pair(name, players, n, lenient)
    dev.reset()
    for t in range(170):
        a_d = dev.step(train=True)
        a_o = ora.step(train=True)
        np.testing.assert_array_equal(a_d.cpu().numpy(), a_o)
        for k in range(players):
            np.testing.assert_array_equal(dev.reward_since_last_action[k].cpu().numpy(), ora.rsl[k])
            np.testing.assert_array_equal(dev.last_action[k].cpu().numpy(), ora.last_action[k])
        np.testing.assert_array_equal(dev.episode_return.cpu().numpy(), ora.total)
        np.testing.assert_array_equal(dev.episode_length.cpu().numpy(), ora.steps)
    torch.cuda.synchronize()
    for k in range(players):
        assert dag[k].events == oag[k].events
        assert any(e[2] == LAST for e in dag[k].events)
    st = dev.stats()
    assert st["episodes"] == len(ora.episodes) > n
    assert st["sum_returns"] == pytest.approx(sum(e[2] for e in ora.episodes))
    assert st["mean_length"] == pytest.approx(np.mean([e[1] for e in ora.episodes]))
    counts = np.bincount([e[0] for e in ora.episodes], minlength=n)
    np.testing.assert_array_equal(dev.episodes_done.cpu().numpy(), counts)
    first = {}
    for g, length, ret in ora.episodes:
        first.setdefault(g, (length, ret))
    got_r, got_l = dev.first_return.cpu().numpy(), dev.first_length.cpu().numpy()
    for g, (length, ret) in first.items():
        assert (got_l[g], got_r[g]) == (length, ret)


def test_eval_mode_and_run_eval():
    dev, ora, dag, oag = _pair("Hanabi-Small", 2, 32)
    out = dev.run_eval(check_every=8)
    assert out["episodes"] == 32
    assert all(len(a.events) == 0 for a in dag)
    # the same policy on the oracle until every game closed once
    while len({e[0] for e in ora.episodes}) < 32:
        ora.step(train=False)
    first = {}
    for g, length, ret in ora.episodes:
        first.setdefault(g, (length, ret))
    assert out["eval_episode_returns"] == pytest.approx(np.mean([v[1] for v in first.values()]))
    assert out["eval_episode_lengths"] == pytest.approx(np.mean([v[0] for v in first.values()]))


def test_end_turn_must_follow_begin_turn():
    import ctypes as C
    from hanabi_agents_b200 import _lib
    dev, _, _, _ = _pair("Hanabi-Small", 2, 4)
    dev.reset()
    a = dev.env.reward.to(dtype=__import__("torch").int32)
    with pytest.raises(_lib.HanabiB200Error):
        _lib.check(_lib.lib().hb_session_end_turn(dev._h, 0, a.data_ptr(), dev.env.reward.data_ptr(), 0, None))
    with pytest.raises(ValueError):
        _lib.check(_lib.lib().hb_session_begin_turn(dev._h, 7, dev.env.step_type.data_ptr(), None))
    h = C.c_void_p()
    with pytest.raises(ValueError):
        _lib.check(_lib.lib().hb_session_create(1, 4, 0, C.byref(h)))


def _dqn(env, seed, **kw):
    from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams
    params = RlaxRainbowParams(train_batch_size=32, layers=[64], experience_buffer_size=4096,
                               target_update_period=5, seed=seed, **kw)
    return DQNAgent(env.observation_spec(), env.action_spec(), params)


def test_dqn_agents_train_through_the_session():
    """Two learning agents: transitions recorded == non-FIRST step types shown; the replay rows are
    the seat's own previous observation -> its own next observation; updates run."""
    import torch
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.session import HanabiParallelSession
    n = 256
    env = HanabiParallelEnv(n, "Hanabi-Small", 2, n_decks=1024, seed=3)
    agents = [_dqn(env, 1), _dqn(env, 2)]
    ses = HanabiParallelSession(env, agents)
    ses.reset()
    shown = [0, 0]
    for t in range(60):
        k = t % 2
        ses.step(train=True)
        torch.cuda.synchronize()
        for j in range(2):
            shown[j] += int((ses.step_types[j] != FIRST).sum().item())
    for j in range(2):
        assert agents[j].experience.size == min(shown[j], 4096) > 0
    assert env.illegal_count() == 0
    # every stored transition of seat 0: obs_tm1 and obs_t are both observations where seat 0 is to
    # move or terminal; rewards are small integers
    tr = agents[0].experience[np.arange(min(64, agents[0].experience.size))]
    assert set(np.unique(tr.reward_t)).issubset({-4.0, -3.0, -2.0, -1.0, 0.0, 1.0, 2.0})
    stats = ses.train(n_iter=2, n_sim_steps=4, n_train_steps=3)
    assert agents[0].train_step == agents[1].train_step == 6
    assert np.isfinite(float(agents[0].last_loss.item()))
    assert stats["episodes"] > 0
    ev = ses.run_eval()
    assert ev["episodes"] == n and np.isfinite(ev["eval_episode_returns"])


def test_one_agent_on_both_seats_keeps_a_last_obs_per_seat():
    """Self-play with ONE DQNAgent object: the session swaps in a per-seat last_obs, so the stored
    obs_tm1 is the observation that seat acted on, not the other seat's."""
    import torch
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.session import HanabiParallelSession
    n = 64
    env = HanabiParallelEnv(n, "Hanabi-Small", 2, n_decks=512, seed=5)
    agent = _dqn(env, 4)
    ses = HanabiParallelSession(env, [agent, agent])
    ses.reset()
    seen = {}                       # (seat, game) -> packed obs it last acted on
    checked = 0
    for t in range(24):
        k = t % 2
        size0 = agent.experience.size
        obs_before = env.obs_packed.clone()
        st_before = env.step_type.clone()
        ses.step(train=True)
        torch.cuda.synchronize()
        types = ses.step_types[k].cpu().numpy()
        other = ses.step_types[1 - k].cpu().numpy()
        n_new = agent.experience.size - size0
        assert n_new == int((types != FIRST).sum() + (other != FIRST).sum())
        # rows were appended seat k first (games in order), then the other seat's terminal rows
        if n_new:
            rows = agent.experience.gather_packed(np.arange(size0, size0 + n_new))
            o1 = rows[0].cpu().numpy()
            r = 0
            for g in np.flatnonzero(types != FIRST):
                np.testing.assert_array_equal(o1[r], seen[(k, int(g))])
                r += 1
                checked += 1
            for g in np.flatnonzero(other != FIRST):
                np.testing.assert_array_equal(o1[r], seen[(1 - k, int(g))])
                r += 1
                checked += 1
        ob = obs_before.cpu().numpy()
        for g in np.flatnonzero(st_before.cpu().numpy() != LAST):
            seen[(k, int(g))] = ob[g]
    assert checked > 100
