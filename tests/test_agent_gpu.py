"""DQNAgent verbs (hanabi_agents_b200/rlax_dqn/agent.py) driven by the CUDA environment, in the
shape of BASELINE configs[0] (rlax_dqn on Hanabi-Small 2p, batch 32, prioritized replay) and of
the plain double-DQN variant; the learner step is re-derived with oracle/learner.py from the very
batch the agent sampled."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _oracle_from(net, n_atoms, noisy, A, vmax=25.0):
    import torch
    from oracle.learner import OracleQNet
    layers = net.get_params(with_noise=True)
    ora = OracleQNet(layers, A, n_atoms, vmax, noisy=noisy, mirror_fp16=True)
    tl = net.get_params(target=True)
    tora = OracleQNet([dict(p, **{k: v for k, v in l.items() if k.startswith("eps")})
                       for p, l in zip(tl, layers)], A, n_atoms, vmax, noisy=noisy, mirror_fp16=True)
    ora.target = tora.target
    return ora


@pytest.mark.parametrize("n_atoms,noisy,priority", [(51, True, True), (1, False, False)])
def test_agent_loop_small_2p(n_atoms, noisy, priority):
    import torch
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams
    N, B = 256, 32
    env = HanabiParallelEnv(N, "Hanabi-Small", 2, seed=3)
    params = RlaxRainbowParams(train_batch_size=B, layers=[512], experience_buffer_size=4096,
                               n_atoms=n_atoms, use_priority=priority, target_update_period=4,
                               epsilon=0.2, beta_is=0.4, seed=7)
    agent = DQNAgent(env.observation_spec(), env.action_spec(), params, noisy=noisy,
                     adam_eps=3.125e-5 if n_atoms > 1 else 1e-8)
    assert agent.requires_vectorized_observation() is True
    assert "DQNAgent" in repr(agent)
    env.reset()
    obs = env.observations_packed()
    agent.add_experience_first(obs, env.step_type)
    expected = 0
    rng = np.random.default_rng(0)
    for t in range(40):
        a = agent.explore(obs)
        assert a.is_cuda and a.dtype == torch.int32
        env.step(a)
        obs = env.observations_packed()
        st = env.step_type.cpu().numpy()
        expected += int((st != 0).sum())
        n_added = agent.add_experience(obs, a, env.reward, env.step_type)
        assert n_added == int((st != 0).sum())
        agent.add_experience_first(obs, env.step_type)
        assert agent.experience.size == min(expected, 4096)
        if t >= 8 and t % 4 == 0:
            before = agent.network.get_params()
            u = rng.uniform(0, 1.0 / B, B)
            if priority:
                # re-derive the step from the very batch the agent is about to sample
                heap0 = agent.experience.sum_tree.heap()
                idx, _, p32, (o1, ac, r, o2, _lm, term) = agent.experience.sample_device(B, u)
                ora = _oracle_from(agent.network, n_atoms, noisy, env.num_moves)
                x1 = agent.experience.unpack_obs(o1).cpu()
                x2 = agent.experience.unpack_obs(o2).cpu()
                ref_mean, ref_per, _ = ora.grads(x1, ac.cpu(), r.cpu(), x2, term.cpu(), prios=p32.cpu(),
                                                 beta=0.4, discount=0.99)
                mean = agent.update(uniforms=u)
                assert abs(float(mean) - ref_mean) <= 1e-3 * abs(ref_mean) + 1e-6
                # priorities of the sampled rows were refreshed from |loss|
                from oracle.replay import priority_pow
                vals = np.array(agent.experience.sum_tree.get_values(idx.cpu().numpy()))
                want = {}
                for i, p in zip(idx.cpu().tolist(), priority_pow(ref_per.abs().numpy(), 0.6)):
                    want[i] = p                                  # last duplicate wins
                np.testing.assert_allclose(vals, [want[i] for i in idx.cpu().tolist()], rtol=2e-3)
                assert not np.array_equal(heap0, agent.experience.sum_tree.heap())
            else:
                np.random.seed(t)
                mean = agent.update()
            assert np.isfinite(float(mean))
            after = agent.network.get_params()
            assert any(not np.array_equal(b[k], a_[k]) for b, a_ in zip(before, after) for k in b)
    assert env.illegal_count() == 0                 # the policy never picked an illegal move
    assert agent.train_step == 8
    # exploit is deterministic given the parameters (no ties in practice) and legal
    g1, g2 = agent.exploit(obs), agent.exploit(obs)
    lm = env.unpack_legal().cpu().numpy()
    live = env.step_type.cpu().numpy() != 2
    assert (lm[np.arange(N), g1.cpu().numpy()][live] == 1).all()
    assert torch.equal(g1, g2)


def test_agent_numpy_compat_and_weights_roundtrip(tmp_path):
    import torch
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams
    N = 64
    env = HanabiParallelEnv(N, "Hanabi-Small", 2, seed=5)
    params = RlaxRainbowParams(train_batch_size=16, experience_buffer_size=512, seed=11)
    agent = DQNAgent(env.observation_spec(), env.action_spec(), params)
    env.reset()
    _, (o8, l8) = env.observations()                 # int8 0/1 rows, the reference's layout
    a_np = agent.exploit((None, (o8.cpu().numpy(), l8.cpu().numpy())))
    assert isinstance(a_np, np.ndarray) and a_np.shape == (N,)
    a_dev = agent.exploit(env.observations_packed())
    np.testing.assert_array_equal(a_np, a_dev.cpu().numpy())
    # save / restore (rlax_rainbow.py:344-365): file names and a faithful round trip
    agent.save_weights(str(tmp_path), "t")
    on, tg = tmp_path / "rlax_rainbow_t_online.pkl", tmp_path / "rlax_rainbow_t_target.pkl"
    assert on.exists() and tg.exists()
    other = DQNAgent(env.observation_spec(), env.action_spec(), params._replace(seed=99))
    assert not torch.equal(other.network.forward(env.obs_packed), agent.network.forward(env.obs_packed))
    other.restore_weights(str(on), str(tg))
    # the trace-time noise constants are not part of the reference's pickles; every network regenerates the
    # reference's own (jax PRNGKey(1234) threefry stream, rlax_dqn/jax_prng.py), whatever params.seed is:
    # the restored network evaluates to the saved one without anything else being copied
    assert torch.equal(other.network.forward(env.obs_packed), agent.network.forward(env.obs_packed))
    # (setting the noise explicitly stays possible)
    noise = [{k: v for k, v in l.items() if k.startswith("eps")} for l in agent.network.get_params(with_noise=True)]
    other.network.set_params(noise, sync_target=True)
    assert torch.equal(other.network.forward(env.obs_packed), agent.network.forward(env.obs_packed))


def test_graph_replay_equals_eager_updates():
    """The learner step replayed as CUDA graphs must be bit-identical to the eager launches."""
    import torch
    from hanabi_agents_b200 import _lib
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams
    N, B = 512, 64
    params = RlaxRainbowParams(train_batch_size=B, layers=[512, 512], experience_buffer_size=8192, seed=5,
                               target_update_period=3)
    agents = []
    for use_graph in (1, 0):
        env = HanabiParallelEnv(N, "Hanabi-Full", 2, seed=9)
        agent = DQNAgent(env.observation_spec(), env.action_spec(), params)
        env.reset()
        obs = env.observations_packed()
        agent.add_experience_first(obs, env.step_type)
        for _ in range(6):
            a = agent.explore(obs)
            env.step(a)
            obs = env.observations_packed()
            agent.add_experience(obs, a, env.reward, env.step_type)
            agent.add_experience_first(obs, env.step_type)
        agent._make_learner()
        _lib.check(_lib.lib().hb_learner_set_graph(agent._learner, use_graph))
        rng = np.random.default_rng(1)
        losses = [float(agent.update(uniforms=rng.uniform(0, 1.0 / B, B))) for _ in range(8)]
        torch.cuda.synchronize()
        captures = int(_lib.lib().hb_learner_info(agent._learner, 1))
        # one graph per staging slot of the ring (2 slots)
        assert captures == (2 if use_graph else 0)
        assert int(_lib.lib().hb_net_info(agent.network._h, 5)) == 8          # Adam steps taken
        agents.append((agent, losses, agent.network.get_params(), agent.experience.sum_tree.heap()))
    (_, l0, p0, h0), (_, l1, p1, h1) = agents
    assert l0 == l1
    for a, b in zip(p0, p1):
        for k in a:
            np.testing.assert_array_equal(a[k], b[k])
    np.testing.assert_array_equal(h0, h1)


def test_learner_update_gradients_and_step_against_the_oracle():
    """DQNAgent.update on the 658-512-512 network through hb_learner_step (the path with the chained forward
    pass, the head's backward product inside the loss, the chained backward tail and per-layer Adam): the
    gradients left in the bucket and the first optix.adam step against oracle/learner.py on the very batch
    the agent samples.  Tolerances as in test_qnet_gpu.py::test_learner_step_parity."""
    import torch
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams
    N, B = 512, 64
    params = RlaxRainbowParams(train_batch_size=B, layers=[512, 512], experience_buffer_size=8192, seed=5)
    env = HanabiParallelEnv(N, "Hanabi-Full", 2, seed=9)
    agent = DQNAgent(env.observation_spec(), env.action_spec(), params)
    env.reset()
    obs = env.observations_packed()
    agent.add_experience_first(obs, env.step_type)
    for _ in range(6):
        a = agent.explore(obs)
        env.step(a)
        obs = env.observations_packed()
        agent.add_experience(obs, a, env.reward, env.step_type)
        agent.add_experience_first(obs, env.step_type)
    rng = np.random.default_rng(3)
    u = rng.uniform(0, 1.0 / B, B)
    idx, _, p32, (o1, ac, r, o2, _lm, term) = agent.experience.sample_device(B, u)
    ora = _oracle_from(agent.network, params.n_atoms, True, env.num_moves)
    x1 = agent.experience.unpack_obs(o1).cpu()
    x2 = agent.experience.unpack_obs(o2).cpu()
    beta = float(params.beta_is(agent.train_step))
    lr = params.learning_rate
    lr = float(lr(agent.train_step)) if callable(lr) else float(lr)
    ref_mean, _, g_ref = ora.grads(x1, ac.cpu(), r.cpu(), x2, term.cpu(), prios=p32.cpu(), beta=beta,
                                   discount=float(params.discount))
    mean = agent.update(uniforms=u)
    torch.cuda.synchronize()
    assert abs(float(mean) - ref_mean) <= 1e-3 * abs(ref_mean) + 1e-6
    got = agent.network.get_params(grads=True)
    assert len(got) == 3
    for l, (gg, gr) in enumerate(zip(got, g_ref)):
        for k_dev, k_ref in (("dw", "w"), ("db", "b")):
            ref = gr[k_ref].numpy()
            err = gg[k_dev] - ref
            assert np.linalg.norm(err) <= 1e-2 * np.linalg.norm(ref) + 1e-12, (l, k_dev)
            assert np.abs(err).max() <= 1e-1 * max(np.abs(ref).max(), 1e-12), (l, k_dev)
    ora.adam(g_ref, lr=lr, eps=agent.adam_eps)
    for l, (gp, rp) in enumerate(zip(agent.network.get_params(), ora.params)):
        for k in rp:
            assert np.abs(gp[k] - rp[k].numpy()).max() <= 0.35 * lr, (l, k)
