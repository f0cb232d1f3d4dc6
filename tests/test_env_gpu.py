"""GPU parity: CUDA env / encoder / legal masks / rule agents (through the C ABI) vs the CPU
oracle on the same decks and actions -- bit-exact observations, masks, rewards, step types and
full game state at every step."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CONFIGS = [("Hanabi-Full", 2), ("Hanabi-Full", 3), ("Hanabi-Full", 4), ("Hanabi-Full", 5),
           ("Hanabi-Small", 2), ("Hanabi-Very-Small", 2)]


def _pair(name, players, n, n_decks, seed=5, threads=8):
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.game import game_params, make_decks
    from oracle.hanabi import HanabiOracleEnv
    p = game_params(name, players)
    decks = make_decks(p["colors"], p["ranks"], n_decks, seed)
    env = HanabiParallelEnv(n, name, players, decks=decks)
    ora = HanabiOracleEnv(n, threads=threads, **p)
    ora.set_decks(decks)
    return env, ora


def _check(env, ora, obs, legal, reward=None, step_type=None):
    import torch
    torch.cuda.synchronize()
    np.testing.assert_array_equal(env.unpack_obs(obs).cpu().numpy().astype(np.uint8), ora.obs)
    np.testing.assert_array_equal(env.unpack_legal(legal).cpu().numpy().astype(np.uint8), ora.legal)
    if reward is not None:
        np.testing.assert_array_equal(reward.cpu().numpy(), ora.reward)
    np.testing.assert_array_equal(step_type.cpu().numpy(), ora.step_type)
    np.testing.assert_array_equal(env.dump(), ora.dump())


@pytest.mark.parametrize("name,players", CONFIGS)
def test_shapes(name, players):
    env, ora = _pair(name, players, 4, 8)
    assert (env.obs_len, env.num_moves, env.deck_size) == (ora.obs_len, ora.num_moves, ora.deck_size)
    assert env.obs_stride % 4 == 0 and env.obs_stride * 32 >= env.obs_len
    assert env.observation_spec().shape == (4, ora.obs_len)
    assert env.action_spec().num_values == ora.num_moves


@pytest.mark.parametrize("name,players", CONFIGS)
def test_random_play_parity(name, players):
    n = 512
    env, ora = _pair(name, players, n, 700)    # pool smaller than games*episodes: decks recycle
    rng = np.random.default_rng(players)
    obs, legal, st = env.reset()
    ora.reset()
    _check(env, ora, obs, legal, None, st)
    for step in range(260):
        lg = ora.legal
        pick = rng.random((n, lg.shape[1])) * lg
        a = pick.argmax(1).astype(np.int32)
        obs, legal, rew, st = env.step(a)
        ora.step(a)
        assert ora.illegal == 0
        _check(env, ora, obs, legal, rew, st)
    assert env.illegal_count() == 0
    assert ora.dump()[:, 7].max() >= 2          # several episodes were played per game
    # other seats' views
    for off in range(players):
        o, l = env.observe(off)
        oo, ol = ora.observe(off)
        np.testing.assert_array_equal(env.unpack_obs(o).cpu().numpy().astype(np.uint8), oo)
        np.testing.assert_array_equal(env.unpack_legal(l).cpu().numpy().astype(np.uint8), ol)


def test_illegal_actions_are_counted_and_ignored():
    env, ora = _pair("Hanabi-Full", 2, 64, 64)
    env.reset(); ora.reset()
    a = np.zeros(64, np.int32)                   # discard at 8 information tokens: illegal
    a[::2] = 5                                   # play card 0: legal
    a[1] = 99
    a[3] = -1
    obs, legal, rew, st = env.step(a)
    ora.step(a)
    _check(env, ora, obs, legal, rew, st)
    assert env.illegal_count() == 32 == ora.illegal


def test_step_host_matches_device_path():
    env, ora = _pair("Hanabi-Full", 2, 256, 512)
    env.reset(); ora.reset()
    rng = np.random.default_rng(0)
    for _ in range(40):
        a = (rng.random(ora.legal.shape) * ora.legal).argmax(1).astype(np.int32)
        obs, legal, rew, st = env.step_host(a)
        ora.step(a)
        bits = np.unpackbits(obs.view(np.uint8), axis=1, bitorder="little")[:, :env.obs_len]
        np.testing.assert_array_equal(bits, ora.obs)
        np.testing.assert_array_equal(rew, ora.reward)
        np.testing.assert_array_equal(st, ora.step_type)


@pytest.mark.parametrize("name,players,chunks", [("Hanabi-Full", 2, 3), ("Hanabi-Full", 5, 4),
                                                 ("Hanabi-Small", 2, 1)])
def test_host_pipeline_matches_device_path(name, players, chunks):
    """env.step_host_async + agent.explore_host_async (chunked over PCIe, unpadded observation rows)
    against env.step + agent.explore on a twin environment, and against the oracle."""
    import torch
    from hanabi_agents_b200.rule_based import RulebasedAgent, predefined_rules
    n = 1000                                  # not a multiple of the chunk size
    env, ora = _pair(name, players, n, 1500)
    twin, _ = _pair(name, players, n, 1500)
    ag = RulebasedAgent(predefined_rules.piers_rules, seed=9)
    ag2 = RulebasedAgent(predefined_rules.piers_rules, seed=9)
    env.reset(); twin.reset(); ora.reset()
    hb = env.host_buffers()
    hb["actions"].copy_(ag.explore(env).cpu())
    a2 = ag2.explore(twin)
    illegal = 0
    for _ in range(60):
        env.step_host_async(hb["actions"], hb["obs"], hb["legal"], hb["reward"], hb["step_type"], chunks=chunks)
        ag.explore_host_async(env, hb["next_actions"])
        env.host_wait()
        ora.step(hb["actions"].numpy())
        illegal += ora.illegal
        obs2, legal2, rew2, st2 = twin.step(a2)
        a2 = ag2.explore(twin)
        torch.cuda.synchronize()
        bits = np.unpackbits(hb["obs"].numpy().view(np.uint8), axis=1, bitorder="little")[:, :env.obs_len]
        np.testing.assert_array_equal(bits, ora.obs)
        np.testing.assert_array_equal(hb["obs"].numpy(), obs2.cpu().numpy()[:, :env.obs_words])
        np.testing.assert_array_equal(hb["legal"].numpy(), legal2.cpu().numpy())
        np.testing.assert_array_equal(hb["reward"].numpy(), ora.reward)
        np.testing.assert_array_equal(hb["step_type"].numpy(), ora.step_type)
        np.testing.assert_array_equal(hb["next_actions"].numpy(), a2.cpu().numpy())
        np.testing.assert_array_equal(env.dump(), ora.dump())
        hb["actions"].copy_(hb["next_actions"])
    # (piers on Hanabi-Small discards at full information tokens -- ruleset.py:207 tests `< 8` -- which
    # the environment counts and ignores; the oracle must have seen the same moves)
    assert illegal == env.illegal_count() == twin.illegal_count()
    assert ag.histogram == ag2.histogram


@pytest.mark.parametrize("name,players,n", [("Hanabi-Full", 2, 20000), ("Hanabi-Full", 5, 1000)])
def test_double_buffered_host_loop_matches_oracle(name, players, n):
    """step_host_begin / explore_host_async / step_host_outputs / host_wait_actions: the next step is begun
    while the previous step's outputs may still be on the wire; two sets of host arrays alternate.  Every
    step's outputs, checked one step late (after host_wait_outputs of their slot), must be the oracle's."""
    from hanabi_agents_b200.rule_based import RulebasedAgent, predefined_rules
    env, ora = _pair(name, players, n, n + 77)
    ag = RulebasedAgent(predefined_rules.piers_rules, seed=9)
    _, orules = _rules("piers")
    env.reset(); ora.reset()
    sets = [env.host_buffers(), env.host_buffers()]
    acts = [sets[0]["actions"], sets[0]["next_actions"], sets[1]["actions"]]   # three rotating action arrays
    acts[0].copy_(ag.explore(env).cpu())
    expected = {}
    pending = None                                   # (slot, host set, step) of the step before
    for t in range(40):
        cur, nxt = acts[t % 3], acts[(t + 1) % 3]
        env.step_host_begin(cur)
        ag.explore_host_async(env, nxt)
        hs = sets[t & 1]
        slot = env.step_host_outputs(hs["obs"], hs["legal"], hs["reward"], hs["step_type"])
        env.host_wait_actions()
        # the oracle plays the same move and must come up with the same next move
        ora.step(cur.numpy())
        expected[t] = (ora.obs.copy(), ora.legal.copy(), ora.reward.copy(), ora.step_type.copy())
        oa, _ = ora.rules_act(orules, 9, t + 1)
        np.testing.assert_array_equal(nxt.numpy(), oa)
        if pending is not None:
            pslot, ph, pt = pending
            env.host_wait_outputs(pslot)
            eo, el, er, es = expected.pop(pt)
            bits = np.unpackbits(ph["obs"].numpy().view(np.uint8), axis=1, bitorder="little")[:, :env.obs_len]
            np.testing.assert_array_equal(bits, eo)
            np.testing.assert_array_equal(ph["reward"].numpy(), er)
            np.testing.assert_array_equal(ph["step_type"].numpy(), es)
        pending = (slot, hs, t)
    env.host_wait()
    np.testing.assert_array_equal(env.dump(), ora.dump())


def _rules(preset):
    from hanabi_agents_b200.rule_based import RulebasedAgent, predefined_rules
    from oracle.hanabi import PRESETS
    return RulebasedAgent(getattr(predefined_rules, preset + "_rules"), seed=77), PRESETS[preset]


@pytest.mark.parametrize("preset", ["piers", "iggi", "outer", "flawed"])
@pytest.mark.parametrize("name,players", [("Hanabi-Full", 2), ("Hanabi-Full", 5), ("Hanabi-Full", 3)])
def test_rule_agents_parity(preset, name, players):
    n = 384
    env, ora = _pair(name, players, n, 999)
    agent, orules = _rules(preset)
    for (rid, fa, ia), r in zip(orules, agent.rules):
        assert (rid, float(fa), ia) == (r.rule_id, r.farg, r.iarg)
    env.reset(); ora.reset()
    hist = np.zeros(len(orules) + 1, np.int64)
    for step in range(220):
        a = agent.explore(env.observations())
        oa, ohit = ora.rules_act(orules, 77, step)
        np.testing.assert_array_equal(a.cpu().numpy(), oa)
        np.add.at(hist, ohit[ohit >= 0], 1)
        obs, legal, rew, st = env.step(a)
        ora.step(oa)
        _check(env, ora, obs, legal, rew, st)
    assert agent.histogram == hist.tolist() and agent.totalCalls == hist.sum()
    assert env.illegal_count() == 0


@pytest.mark.parametrize("name,players", [("Hanabi-Full", 2), ("Hanabi-Full", 4)])
def test_intended_rules_parity(name, players):
    """tell_anyone_useless_card / tell_playable_card / discard_probably_useless (dead in the
    reference, intended semantics here: SURVEY 8f row 4) against the C oracle, stepwise and fused."""
    from hanabi_agents_b200.rule_based import RulebasedAgent, Ruleset
    from oracle import hanabi as oh
    rules = [Ruleset.play_safe_card, Ruleset.tell_playable_card, Ruleset.tell_anyone_useless_card,
             Ruleset.discard_probably_useless_factory(0.4), Ruleset.osawa_discard, Ruleset.tell_randomly]
    orules = [(oh.R_PLAY_SAFE_CARD, 0, 0), (oh.R_TELL_PLAYABLE_CARD, 0, 0), (oh.R_TELL_ANYONE_USELESS_CARD, 0, 0),
              (oh.R_DISCARD_PROBABLY_USELESS, 0.4, 0), (oh.R_OSAWA_DISCARD, 0, 0), (oh.R_TELL_RANDOMLY, 0, 0)]
    n = 384
    env, ora = _pair(name, players, n, 999)
    agent = RulebasedAgent(rules, seed=77)
    env.reset(); ora.reset()
    hist = np.zeros(len(orules) + 1, np.int64)
    for step in range(200):
        a = agent.explore(env.observations())
        oa, ohit = ora.rules_act(orules, 77, step)
        np.testing.assert_array_equal(a.cpu().numpy(), oa)
        np.add.at(hist, ohit[ohit >= 0], 1)
        obs, legal, rew, st = env.step(a)
        ora.step(oa)
        _check(env, ora, obs, legal, rew, st)
    assert agent.histogram == hist.tolist()
    assert min(hist[1:4]) > 0, hist             # each of the three fired
    assert env.illegal_count() == 0
    # the same rule list through the fused self-play kernel
    env2, ora2 = _pair(name, players, n, 999)
    agent2 = RulebasedAgent(rules, seed=77)
    env2.reset(); ora2.reset()
    T = 120
    obs, legal, rew, st, act = env2.selfplay(agent2, T, obs="alloc", legal="alloc", reward="alloc",
                                             step_type="alloc", action="alloc")
    for t in range(T):
        oa, _ = ora2.rules_act(orules, 77, t)
        ora2.step(oa)
        np.testing.assert_array_equal(act[t].cpu().numpy(), oa)
    _check(env2, ora2, obs[T - 1], legal[T - 1], rew[T - 1], st[T - 1])


def test_fused_selfplay_equals_stepwise():
    n, T = 300, 150
    env, ora = _pair("Hanabi-Full", 2, n, 1000)
    agent, orules = _rules("piers")
    env.reset(); ora.reset()
    obs, legal, rew, st, act = env.selfplay(agent, T, obs="alloc", legal="alloc", reward="alloc",
                                            step_type="alloc", action="alloc")
    for t in range(T):
        oa, _ = ora.rules_act(orules, 77, t)
        ora.step(oa)
        np.testing.assert_array_equal(act[t].cpu().numpy(), oa)
        np.testing.assert_array_equal(env.unpack_obs(obs[t]).cpu().numpy().astype(np.uint8), ora.obs)
        np.testing.assert_array_equal(env.unpack_legal(legal[t]).cpu().numpy().astype(np.uint8), ora.legal)
        np.testing.assert_array_equal(rew[t].cpu().numpy(), ora.reward)
        np.testing.assert_array_equal(st[t].cpu().numpy(), ora.step_type)
    np.testing.assert_array_equal(env.dump(), ora.dump())


@pytest.mark.parametrize("name,players,n", [("Hanabi-Full", 2, 5000), ("Hanabi-Full", 5, 700),
                                            ("Hanabi-Small", 2, 64)])
def test_back_to_back_selfplay_launches_overlap_safely(name, players, n):
    """Consecutive self-play launches overlap on the device (programmatic dependent launch; game groups
    are handed from one launch to the next through a FIFO ring): many short launches in a row, with
    different step counts and separate output buffers, interleaved with stepwise calls, must produce
    exactly the oracle's games."""
    import torch
    env, ora = _pair(name, players, n, 3 * n + 7)
    agent, orules = _rules("piers")
    env.reset(); ora.reset()
    t = 0
    launches = []
    for rep in range(3):
        for T in (1, 7, 2, 33, 5, 1, 16):
            launches.append((t, T, env.selfplay(agent, T, obs="alloc", legal="alloc", reward="alloc",
                                                step_type="alloc", action="alloc")))
            t += T
        torch.cuda.synchronize()
        for t0, T, (obs, legal, rew, st, act) in launches:
            for k in range(T):
                oa, _ = ora.rules_act(orules, 77, t0 + k)
                ora.step(oa)
                np.testing.assert_array_equal(act[k].cpu().numpy(), oa)
                np.testing.assert_array_equal(rew[k].cpu().numpy(), ora.reward)
                np.testing.assert_array_equal(st[k].cpu().numpy(), ora.step_type)
            np.testing.assert_array_equal(env.unpack_obs(obs[T - 1]).cpu().numpy().astype(np.uint8), ora.obs)
        launches = []
        np.testing.assert_array_equal(env.dump(), ora.dump())
        # a stepwise move between two runs of fused launches (ordinary stream order)
        a = agent.explore(env)
        oa, _ = ora.rules_act(orules, 77, t)
        np.testing.assert_array_equal(a.cpu().numpy(), oa)
        env.step(a); ora.step(oa)
        t += 1
        agent._step = t
    if name == "Hanabi-Full":      # (the rules hard-code 8 information tokens: illegal discards on Hanabi-Small)
        assert env.illegal_count() == 0


def test_overlapping_launches_at_benchmark_size_equal_serialised_launches():
    """The benchmark's pattern: launches of 64 steps over 65,536 games, back to back (they overlap on the
    device), every launch into its own buffers -- against the same launches with a device synchronisation
    after each (no overlap).  Every output bit must agree."""
    import torch
    from hanabi_agents_b200.env import HanabiParallelEnv
    n, T, L = 65536, 64, 5
    outs = []
    for serialise in (False, True):
        env = HanabiParallelEnv(n, "Hanabi-Full", 2, n_decks=2 * n, seed=11)
        agent, _ = _rules("piers")
        env.reset()
        bufs = [env.selfplay(agent, T, obs="alloc", legal="alloc", reward="alloc", step_type="alloc",
                             action="alloc") for _ in range(1)]
        torch.cuda.synchronize()
        for _ in range(L - 1):
            bufs.append(env.selfplay(agent, T, obs="alloc", legal="alloc", reward="alloc", step_type="alloc",
                                     action="alloc"))
            if serialise:
                torch.cuda.synchronize()
        torch.cuda.synchronize()
        outs.append((bufs, env.dump(), env.illegal_count(), agent.histogram))
    (a, da, ia, ha), (b, db, ib, hb_) = outs
    np.testing.assert_array_equal(da, db)
    assert ia == ib == 0 and ha == hb_
    for x, y in zip(a, b):
        for u, v in zip(x, y):
            assert torch.equal(u, v)


def test_sharded_envs_equal_one_env():
    """Two shards (game_id_base) reproduce the games of one big env: no data-path collective."""
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.game import make_decks
    n = 256
    decks = make_decks(5, 5, 600, 3)
    whole = HanabiParallelEnv(n, decks=decks)
    parts = [HanabiParallelEnv(n // 2, decks=decks, game_id_base=b, n_total=n) for b in (0, n // 2)]
    agent, _ = _rules("piers")
    agents = [_rules("piers")[0] for _ in parts]
    whole.reset()
    for p in parts:
        p.reset()
    for _ in range(120):
        whole.step(agent.explore(whole))
        for p, ag in zip(parts, agents):
            p.step(ag.explore(p))
    np.testing.assert_array_equal(whole.dump(), np.concatenate([p.dump() for p in parts]))
    import torch
    np.testing.assert_array_equal(whole.obs_packed.cpu().numpy(),
                                  torch.cat([p.obs_packed for p in parts]).cpu().numpy())


@pytest.mark.parametrize("players,n,T", [(2, 65536, 100), (2, 131072, 70), (5, 32768, 90)])
def test_full_size_properties_and_oracle(players, n, T):
    """BASELINE sizes per GPU: cfg 2 (65,536 Hanabi-Full 2p games), the shard of cfg 4 (1,048,576 games over
    8 GPUs) and of cfg 5 (262,144 Hanabi-Full 5p games over 8 GPUs), rule-based self-play in launches of
    at most 64 steps (the launches overlap on the device)."""
    env, ora = _pair("Hanabi-Full", players, n, 2 * n + 3, threads=32)
    agent, orules = _rules("piers")
    env.reset(); ora.reset()
    done = 0
    while done < T:                                # no outputs kept: state only
        k = min(64, T - done)
        env.selfplay(agent, k)
        done += k
    for t in range(T):
        oa, _ = ora.rules_act(orules, 77, t)
        ora.step(oa)
    d = env.dump()
    np.testing.assert_array_equal(d, ora.dump())
    # size-independent invariants (A.6): card conservation, token ranges
    C, P, H = 5, players, 5 if players <= 3 else 4
    hand_n = d[:, 8 + C:8 + C + P].sum(1)
    disc = d[:, 8 + C + P + 5 * P * H:8 + C + P + 5 * P * H + 25].sum(1)
    assert np.all(d[:, 3] + hand_n + disc + d[:, 8:8 + C].sum(1) == 50)
    assert np.all((d[:, 1] >= 0) & (d[:, 1] <= 8)) and np.all((d[:, 2] >= 0) & (d[:, 2] <= 3))
    assert env.illegal_count() == 0


@pytest.mark.parametrize("name,players", [("Hanabi-Full", 2), ("Hanabi-Full", 5)])
def test_osawa_discard_intended_flag(name, players):
    """Ruleset.osawa_discard_factory('intended') (SURVEY RB-3: the reading of ruleset.py:229-230 as a
    documented switch) against the oracle's iarg = 1, stepwise and fused; and it really differs from the
    quirk reading on the same games (a finished LAST colour makes the quirk discard un-hinted cards)."""
    from hanabi_agents_b200.rule_based import RulebasedAgent, Ruleset
    from oracle import hanabi as oh
    with pytest.raises(ValueError):
        Ruleset.osawa_discard_factory("other")
    assert Ruleset.osawa_discard_factory("quirk").iarg == Ruleset.osawa_discard.iarg == 0
    rules = [Ruleset.play_safe_card, Ruleset.osawa_discard_factory("intended"), Ruleset.tell_playable_card_outer,
             Ruleset.tell_unknown, Ruleset.discard_oldest_first]
    orules = [(oh.R_PLAY_SAFE_CARD, 0, 0), (oh.R_OSAWA_DISCARD, 0, 1), (oh.R_TELL_PLAYABLE_CARD_OUTER, 0, 0),
              (oh.R_TELL_UNKNOWN, 0, 0), (oh.R_DISCARD_OLDEST_FIRST, 0, 0)]
    n = 512
    env, ora = _pair(name, players, n, 4242)
    agent = RulebasedAgent(rules, seed=5)
    env.reset(); ora.reset()
    differs = 0
    for step in range(260):
        a = agent.explore(env.observations())
        oa, _ = ora.rules_act(orules, 5, step)
        np.testing.assert_array_equal(a.cpu().numpy(), oa)
        quirk, _ = ora.rules_act([(r, f, 0) for r, f, _ in orules], 5, step)
        differs += int((quirk != oa).sum())
        obs, legal, rew, st = env.step(a)
        ora.step(oa)
        _check(env, ora, obs, legal, rew, st)
    assert env.illegal_count() == 0
    assert differs > 0, "the two readings never disagreed: the flag is not exercised"
    env2, ora2 = _pair(name, players, n, 4242)
    agent2 = RulebasedAgent(rules, seed=5)
    env2.reset(); ora2.reset()
    T = 100
    _, _, _, _, act = env2.selfplay(agent2, T, obs=None, legal=None, reward=None, step_type=None, action="alloc")
    for t in range(T):
        oa, _ = ora2.rules_act(orules, 5, t)
        np.testing.assert_array_equal(act[t].cpu().numpy(), oa, err_msg="fused step %d" % t)
        ora2.step(oa)


@pytest.mark.parametrize("name,players", CONFIGS)
def test_minimal_and_seer_observation_types(name, players):
    """hb_env_observe_as (HLE's kMinimal / kSeer observation types) against the oracle's
    ho_encode_typed, every bit, every seat offset, along random legal play-outs."""
    import torch
    n = 256
    env, ora = _pair(name, players, n, 4 * n)
    env.reset(); ora.reset()
    rng = np.random.default_rng(7)
    for step in range(70):
        for t, tname in ((0, "minimal"), (1, "card_knowledge"), (2, "seer")):
            for offset in range(players):
                got, bits = env.observe_as(tname, offset, unpack=True)
                want = ora.observe_typed(t, offset)
                assert bits == want.shape[1]
                np.testing.assert_array_equal(got.cpu().numpy().astype(np.uint8), want,
                                              err_msg="%s offset %d step %d" % (tname, offset, step))
        packed, bits = env.observe_as("seer")
        assert packed.shape[1] % 4 == 0 and packed.shape[1] * 32 >= bits
        legal = ora.legal
        acts = np.array([rng.choice(np.nonzero(legal[g])[0]) for g in range(n)], np.int32)
        env.step(torch.as_tensor(acts, device="cuda"))
        ora.step(acts)
    with pytest.raises(ValueError):
        from hanabi_agents_b200 import _lib
        _lib.check(_lib.lib().hb_env_observe_as(env._h, 3, 0, packed.data_ptr(), packed.shape[1], None))
