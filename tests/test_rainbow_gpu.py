"""GPU: TF-Rainbow semantics (hanabi_agents/rainbow/) as an alternate mode of the same kernels --
the learner variant against oracle/learner.py (tf_variant), TF's Adam, and the Dopamine-style verbs of
RainbowAgent against a literal restatement of dqn_agent.py:284-383 (what ends up in the replay)."""
import random

import numpy as np
import pytest

from test_qnet_gpu import _obs, _setup, pack_bits

pytestmark = pytest.mark.gpu


def test_rainbow_target_and_tf_adam_parity():
    import ctypes as C
    import torch
    from hanabi_agents_b200 import _lib
    obs_len, hidden, A, n_atoms, B = 658, [512, 512], 20, 51, 64
    net, ora, stride = _setup(obs_len, hidden, A, n_atoms, noisy=False, seed=9)
    rng = np.random.default_rng(4)
    L = _lib.lib()
    lr, eps, gamma = 2.5e-3, 3.125e-5, 0.99
    for it in range(3):
        x1, x1p = _obs(rng, B, obs_len, stride)
        x2, x2p = _obs(rng, B, obs_len, stride)
        a = rng.integers(0, A, B).astype(np.int32)
        r = rng.integers(-2, 2, B).astype(np.float32)
        term = (rng.random(B) < 0.25).astype(np.uint8)
        legal = (rng.random((B, A)) < 0.5).astype(np.uint8)
        legal[:, 3] = 1
        legal[term == 1] = 0                                   # terminal rows: no legal next move
        w = (rng.random(B) * 0.9 + 0.1).astype(np.float32)
        w /= w.max()
        dev = lambda v: torch.as_tensor(v, device="cuda")  # noqa: E731
        lp = dev(pack_bits(legal, 1).view(np.int32))
        loss = torch.empty(B, device="cuda")
        mean = torch.empty(1, device="cuda")
        da, dr, dt, dw = dev(a), dev(r), dev(term), dev(w)
        _lib.check(L.hb_net_grads_rainbow(net._h, x1p.data_ptr(), da.data_ptr(), dr.data_ptr(), x2p.data_ptr(),
                                          lp.data_ptr(), dt.data_ptr(), dw.data_ptr(), B, gamma,
                                          loss.data_ptr(), mean.data_ptr(), _lib.current_stream_ptr()))
        torch.cuda.synchronize()
        t = torch.tensor
        mean_ref, per_ref, g_ref = ora.grads(t(x1), t(a), t(r), t(x2), t(term), weights=t(w), discount=gamma,
                                             tf_variant=True, legal_t=t(legal))
        torch.testing.assert_close(loss.cpu(), per_ref, rtol=1e-3, atol=1e-3 * float(per_ref.abs().max()))
        assert abs(float(mean) - mean_ref) <= 1e-3 * abs(mean_ref) + 1e-6
        for l, (gg, gr) in enumerate(zip(net.get_params(grads=True), g_ref)):
            for k_dev, k_ref in (("dw", "w"), ("db", "b")):
                ref = gr[k_ref].numpy()
                err = gg[k_dev] - ref
                assert np.linalg.norm(err) <= 1e-2 * np.linalg.norm(ref) + 1e-12, (it, l, k_dev)
        # the double-Q (rlax) target on the same batch differs: the variants are not aliases
        if it == 0:
            loss2, _ = net.grads(x1p, da, dr, x2p, dt, prios=None, discount=gamma)
            assert not torch.allclose(loss2, loss)
            _lib.check(L.hb_net_grads_rainbow(net._h, x1p.data_ptr(), da.data_ptr(), dr.data_ptr(), x2p.data_ptr(),
                                              lp.data_ptr(), dt.data_ptr(), dw.data_ptr(), B, gamma,
                                              loss.data_ptr(), mean.data_ptr(), _lib.current_stream_ptr()))
        _lib.check(L.hb_net_adam_tf(net._h, lr, 0.9, 0.999, eps, 1.0, _lib.current_stream_ptr()))
        if it == 0:
            ora.adam_tf(g_ref, lr=lr, eps=eps)
            for l, (gp, rp) in enumerate(zip(net.get_params(), ora.params)):
                for k in rp:
                    assert np.abs(gp[k] - rp[k].numpy()).max() <= 0.35 * lr, (l, k)
        for gp, rp in zip(net.get_params(), ora.params):
            for k in rp:
                rp[k] = torch.tensor(gp[k])
        ora.t = it + 1
    with pytest.raises(ValueError):
        _lib.check(L.hb_net_grads_rainbow(net._h, x1p.data_ptr(), da.data_ptr(), dr.data_ptr(), x2p.data_ptr(),
                                          None, dt.data_ptr(), dw.data_ptr(), B, gamma, loss.data_ptr(),
                                          mean.data_ptr(), None))


def test_tf_adam_multi_step():
    import torch
    from hanabi_agents_b200 import _lib
    net, ora, _ = _setup(171, [64], 11, 51, False, seed=5)
    g = torch.Generator(device="cuda").manual_seed(0)
    for step in range(5):
        net.grad_bucket.copy_(torch.randn(net.n_grad, device="cuda", generator=g) * 1e-3)
        torch.cuda.synchronize()
        grads = [{"w": torch.tensor(ge["dw"]), "b": torch.tensor(ge["db"])} for ge in net.get_params(grads=True)]
        _lib.check(_lib.lib().hb_net_adam_tf(net._h, 1e-3, 0.9, 0.999, 3.125e-5, 1.0, _lib.current_stream_ptr()))
        ora.adam_tf(grads, lr=1e-3, eps=3.125e-5)
        for gp, rp in zip(net.get_params(), ora.params):
            for k in rp:
                np.testing.assert_allclose(gp[k], rp[k].numpy(), rtol=1e-5, atol=2e-7)


class _LiteralReplayLog:
    """dqn_agent.py:333-383 restated: what _post_transitions hands to _store_transition, in order."""

    def __init__(self, num_players):
        self.transitions = [[] for _ in range(num_players)]
        self.stored = []          # (obs, action, reward, terminal, legal)

    def record(self, player, reward, obs, legal, action, begin=False):
        self.transitions[player].append((reward, np.array(obs, np.uint8), np.array(legal, np.float32), action, begin))

    def post(self, terminal_rewards):
        for player in range(len(self.transitions)):
            n = len(self.transitions[player])
            for index, tr in enumerate(self.transitions[player]):
                final = index == n - 1
                reward = terminal_rewards[player] if final else self.transitions[player][index + 1][0]
                self.stored.append((tr[1], tr[3], reward, final, tr[2]))
            self.transitions[player] = []


def test_dopamine_verbs_fill_the_replay_like_the_reference():
    """run_one_episode (run_experiment.py:280-336) over the oracle env with epsilon = 1 (actions =
    np.random.choice over the legal moves, the reference's own draw) -- every replay row must be the
    transition the reference's _post_transitions stores, with its successor observation / legal moves."""
    import torch
    from hanabi_agents_b200.rainbow import RainbowAgent
    from oracle.hanabi import HanabiOracleEnv, game_params, make_decks
    from oracle.session import run_one_episode
    p = game_params("Hanabi-Small", 2)
    env = HanabiOracleEnv(1, **p)
    agent = RainbowAgent(num_actions=env.num_moves, observation_size=env.obs_len, num_players=2,
                         epsilon_train=1.0, min_replay_history=10 ** 9, layer_size=64, num_layers=1,
                         replay_capacity=4096, batch_size=8, seed=1)
    log = _LiteralReplayLog(2)

    class Tap:                    # the verbs as the driver calls them; 0/1 legal -> 0/-inf as parse_observations
        @staticmethod
        def _fmt(legal):
            return np.where(np.asarray(legal) != 0, 0.0, -np.inf).astype(np.float32)

        def begin_episode(self, player, legal, obs):
            a = agent.begin_episode(player, self._fmt(legal), obs)
            log.record(player, 0, obs, self._fmt(legal), a, begin=True)
            return a

        def step(self, reward, player, legal, obs):
            a = agent.step(reward, player, self._fmt(legal), obs)
            log.record(player, reward, obs, self._fmt(legal), a)
            return a

        def end_episode(self, final_rewards):
            agent.end_episode(final_rewards)
            log.post(final_rewards)

    random.seed(3)
    np.random.seed(3)
    for ep in range(12):
        env.set_decks(make_decks(p["colors"], p["ranks"], 1, 100 + ep))
        run_one_episode(env, Tap())
    rep = agent._replay
    assert rep.size == len(log.stored) == agent.add_count > 20
    tr = rep[np.arange(rep.size)]
    pri = np.asarray(rep.sum_tree.get_values(list(range(rep.size))))
    np.testing.assert_array_equal(pri, np.full(rep.size, 100.0, np.float32))      # DEFAULT_PRIORITY
    for i, (obs, action, reward, final, legal) in enumerate(log.stored):
        np.testing.assert_array_equal(tr.observation_tm1[i], obs)
        assert int(tr.action_tm1[i, 0]) == action and float(tr.reward_t[i, 0]) == float(reward)
        assert bool(tr.terminal_t[i, 0]) == final
        if not final:             # successor = the same player's next stored element
            nobs, _, _, _, nlegal = log.stored[i + 1]
            np.testing.assert_array_equal(tr.observation_t[i], nobs)
            np.testing.assert_array_equal(tr.legal_moves_t[i], (nlegal == 0).astype(np.int8))
    # eval mode stores nothing and does not train
    agent.eval_mode = True
    before = rep.size
    env.set_decks(make_decks(p["colors"], p["ranks"], 1, 999))
    run_one_episode(env, Tap())
    assert rep.size == before


def test_training_cadence_and_priorities():
    import torch
    from hanabi_agents_b200.rainbow import RainbowAgent
    from oracle.hanabi import HanabiOracleEnv, game_params, make_decks
    from oracle.session import run_one_episode
    p = game_params("Hanabi-Small", 2)
    env = HanabiOracleEnv(1, **p)
    agent = RainbowAgent(num_actions=env.num_moves, observation_size=env.obs_len, num_players=2,
                         epsilon_train=0.2, epsilon_decay_period=50, min_replay_history=40, update_period=4,
                         target_update_period=16, layer_size=64, num_layers=2, replay_capacity=2048,
                         batch_size=16, learning_rate=1e-3, seed=2)

    class Verbs:
        def begin_episode(self, player, legal, obs):
            return agent.begin_episode(player, np.where(legal != 0, 0.0, -np.inf), obs)

        def step(self, reward, player, legal, obs):
            return agent.step(reward, player, np.where(legal != 0, 0.0, -np.inf), obs)

        def end_episode(self, final_rewards):
            agent.end_episode(final_rewards)

    random.seed(5)
    np.random.seed(5)
    trained = 0
    for ep in range(40):
        env.set_decks(make_decks(p["colors"], p["ranks"], 1, 500 + ep))
        before = agent.last_loss
        run_one_episode(env, Verbs())
        assert env.illegal == 0
    assert agent.last_loss is not None and np.isfinite(float(agent.last_loss.item()))
    rep = agent._replay
    pri = np.asarray(rep.sum_tree.get_values(list(range(rep.size))))
    assert (pri == 100.0).any() and (pri != 100.0).any()            # sampled rows got sqrt(loss + 1e-10)
    idx, w = agent._last_batch
    assert float(w.max()) == 1.0 and float(w.min()) > 0.0
    # one explicit training step with known draws: priorities of the sampled rows = sqrt(loss + 1e-10)
    u = np.linspace(0.01, 0.99, 16)
    agent._train(uniforms=u)
    idx, _ = agent._last_batch
    got = torch.as_tensor(rep.sum_tree.get_values(idx)).cpu().numpy()
    loss = agent._loss.cpu().numpy().astype(np.float64)
    want = {}
    for i, l in zip(idx.cpu().numpy(), loss):                       # last duplicate wins
        want[int(i)] = np.float32((abs(l) + 1e-10) ** 0.5)
    # the tree leaf ends as old + (new - old) in fp32 (sum_tree.h:80-90), not exactly `new`: ulp(100) slack
    np.testing.assert_allclose(got, [want[int(i)] for i in idx.cpu().numpy()], rtol=0, atol=1.6e-5)
    assert agent.training_steps > 100
