"""GPU parity of the replay shard (csrc/replay.cu through the C ABI) against oracle/replay.py and
the fixtures produced by the reference's own ExperienceBuffer: bit-exact ring contents, cursor,
sampled indices / probabilities given identical uniforms, priorities and every sum-tree node."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "replay_ref.npz"))


def _all_rows(buf):
    tr = buf[np.arange(buf.capacity)]
    return dict(obs_tm1=tr.observation_tm1, act=tr.action_tm1, rew=tr.reward_t, obs_t=tr.observation_t,
                lms=tr.legal_moves_t, term=tr.terminal_t)


def test_reference_fixtures_through_the_numpy_api():
    from hanabi_agents_b200.rlax_dqn import ExperienceBuffer
    L, A, cap = int(GOLD["L"]), int(GOLD["A"]), int(GOLD["cap"])
    buf = ExperienceBuffer(L, A, 1, cap)
    for k, n in enumerate(GOLD["sizes"]):
        assert buf.get_update_indices(int(n)) == GOLD["upd%d" % k].tolist()
        buf.add_transitions(GOLD["in%d_o1" % k], GOLD["in%d_a" % k], GOLD["in%d_r" % k],
                            GOLD["in%d_o2" % k], GOLD["in%d_lm" % k], GOLD["in%d_t" % k])
        assert buf.size == int(GOLD["st%d_size" % k]) and buf.oldest_entry == int(GOLD["st%d_oldest" % k])
        for name, arr in _all_rows(buf).items():
            np.testing.assert_array_equal(arr, GOLD["st%d_%s" % (k, name)], err_msg="%d %s" % (k, name))
    tr = buf[GOLD["get_idx"]]
    for name, got in zip(tr._fields, tr):
        assert got.dtype == GOLD["get_" + name].dtype, name
        np.testing.assert_array_equal(got, GOLD["get_" + name])
    with pytest.raises(IndexError):
        buf[[cap]]


def _rand_tr(rng, n, L, A):
    return ((rng.random((n, L)) < .2).astype(np.int8), rng.integers(0, A, (n, 1)).astype(np.int8),
            rng.integers(-3, 3, (n, 1)).astype(np.float64), (rng.random((n, L)) < .2).astype(np.int8),
            (rng.random((n, A)) < .5).astype(np.int8), rng.random((n, 1)) < .1)


@pytest.mark.parametrize("L,A,cap", [(658, 20, 1000), (171, 11, 300), (1280, 48, 512)])
def test_priority_buffer_parity(L, A, cap):
    from hanabi_agents_b200.rlax_dqn import PriorityBuffer
    from oracle.replay import OraclePriorityBuffer
    rng = np.random.default_rng(cap)
    dev, ora = PriorityBuffer(L, A, 1, cap), OraclePriorityBuffer(L, A, 1, cap)
    assert dev.max_priority == np.float32(0.6) and dev.sum_tree.get_capacity() == ora.sum_tree.get_capacity()
    B = 64
    for it in range(40):
        n = int(rng.integers(1, cap // 3))
        tr = _rand_tr(rng, n, L, A)
        dev.add_transitions(*tr)
        ora.add_transitions(*tr)
        assert (dev.size, dev.oldest_entry) == (ora.size, ora.oldest_entry)
        np.testing.assert_array_equal(dev.sum_tree.heap()[1:], ora.sum_tree.heap()[1:])
        u = rng.uniform(0, 1.0 / B, B)
        idx, prios, batch = dev.sample_batch(B, u)
        oidx, oprios, obatch = ora.sample_batch(B, u)
        assert idx == oidx
        np.testing.assert_array_equal(prios, oprios)
        for g, o in zip(batch, obatch):
            np.testing.assert_array_equal(g, o)
        td = np.abs(rng.standard_normal(B)).astype(np.float32) * (it % 3)
        dev.update_priorities(idx, td)
        ora.update_priorities(oidx, td)
        np.testing.assert_array_equal(dev.sum_tree.heap()[1:], ora.sum_tree.heap()[1:])
        assert dev.max_priority == np.float32(ora.max_priority)
        assert dev.min_priority == np.float32(ora.min_priority)
    for name, arr in _all_rows(dev).items():
        np.testing.assert_array_equal(arr, dict(obs_tm1=ora._obs_tm1_buf, act=ora._act_tm1_buf, rew=ora._rew_t_buf,
                                                obs_t=ora._obs_t_buf, lms=ora._lms_t_buf,
                                                term=ora._terminal_t_buf)[name])


def test_masked_append_matches_add_experience():
    """DQNAgent.add_experience / add_experience_first bookkeeping (rlax_rainbow.py:292-308) on packed
    device rows vs the same statements on numpy arrays feeding the restated buffer."""
    import torch
    from hanabi_agents_b200.rlax_dqn import PriorityBuffer
    from oracle.replay import OraclePriorityBuffer
    from test_qnet_gpu import pack_bits
    L, A, cap, N = 658, 20, 5000, 3000
    rng = np.random.default_rng(4)
    dev, ora = PriorityBuffer(L, A, 1, cap), OraclePriorityBuffer(L, A, 1, cap)
    last_obs = np.zeros((N, L), np.int8)
    d_last = torch.zeros((N, dev.obs_stride), dtype=torch.int32, device="cuda")
    to_dev = lambda a: torch.as_tensor(a, device="cuda")  # noqa: E731
    for it in range(9):
        obs = (rng.random((N, L)) < .2).astype(np.int8)
        legal = (rng.random((N, A)) < .5).astype(np.int8)
        act = rng.integers(0, A, N).astype(np.int32)
        rew = rng.integers(-2, 2, N).astype(np.float32)
        st = rng.choice([0, 1, 1, 1, 2], N).astype(np.uint8) if it else np.zeros(N, np.uint8)
        d_obs = to_dev(pack_bits(obs, dev.obs_stride).view(np.int32))
        d_leg = to_dev(pack_bits(legal, dev.legal_words).view(np.int32))
        # reference statements
        first = st == 0
        nf = st != 0
        if nf.any():
            ora.add_transitions(last_obs[nf], act[nf].reshape(-1, 1), rew[nf].reshape(-1, 1), obs[nf],
                                legal[nf], (st[nf] == 2).reshape(-1, 1))
            last_obs[nf] = obs[nf]
        last_obs[first] = obs[first]
        # device: in one call, in the two halves (hb_rb_add_masked_begin / _finish), or without the host in
        # the loop (hb_rb_add_masked_async: the ring cursor lives on the device until somebody asks)
        if it % 3 == 0:
            n_added = dev.add_masked(d_last, d_obs, d_leg, to_dev(act), to_dev(rew), to_dev(st))
        elif it % 3 == 1:
            d_st = to_dev(st)
            dev.add_masked_begin(d_st, N)
            n_added = dev.add_masked_finish(d_last, d_obs, d_leg, to_dev(act), to_dev(rew), d_st)
        else:
            dev.add_masked_async(d_last, d_obs, d_leg, to_dev(act), to_dev(rew), to_dev(st))
            n_added = int(nf.sum())          # (nothing is returned; size / oldest_entry are checked below)
        dev.store_first(d_last, d_obs, to_dev(st))
        assert n_added == int(nf.sum())
        assert (dev.size, dev.oldest_entry) == (ora.size, ora.oldest_entry)
        np.testing.assert_array_equal(dev.unpack_obs(d_last).cpu().numpy(), last_obs)
        np.testing.assert_array_equal(dev.sum_tree.heap()[1:], ora.sum_tree.heap()[1:])
    # a run of asynchronous appends with nothing read back in between (the device-side cursor carries the
    # ring position from one to the next, around the end of the ring as well; with and without the copy of
    # the cursor having reached the host before the next call)
    for it in range(7):
        obs = (rng.random((N, L)) < .2).astype(np.int8)
        legal = (rng.random((N, A)) < .5).astype(np.int8)
        act = rng.integers(0, A, N).astype(np.int32)
        rew = rng.integers(-2, 2, N).astype(np.float32)
        st = rng.choice([0, 0, 0, 1, 2], N).astype(np.uint8)
        d_obs = to_dev(pack_bits(obs, dev.obs_stride).view(np.int32))
        d_leg = to_dev(pack_bits(legal, dev.legal_words).view(np.int32))
        first, nf = st == 0, st != 0
        ora.add_transitions(last_obs[nf], act[nf].reshape(-1, 1), rew[nf].reshape(-1, 1), obs[nf], legal[nf],
                            (st[nf] == 2).reshape(-1, 1))
        last_obs[nf] = obs[nf]
        last_obs[first] = obs[first]
        dev.add_masked_async(d_last, d_obs, d_leg, to_dev(act), to_dev(rew), to_dev(st))
        dev.store_first(d_last, d_obs, to_dev(st))
        if it % 2:
            torch.cuda.synchronize()
    assert (dev.size, dev.oldest_entry) == (ora.size, ora.oldest_entry)
    np.testing.assert_array_equal(dev.unpack_obs(d_last).cpu().numpy(), last_obs)
    np.testing.assert_array_equal(dev.sum_tree.heap()[1:], ora.sum_tree.heap()[1:])
    got = _all_rows(dev)
    np.testing.assert_array_equal(got["obs_tm1"], ora._obs_tm1_buf)
    np.testing.assert_array_equal(got["obs_t"], ora._obs_t_buf)
    np.testing.assert_array_equal(got["act"], ora._act_tm1_buf)
    np.testing.assert_array_equal(got["rew"], ora._rew_t_buf)
    np.testing.assert_array_equal(got["lms"], ora._lms_t_buf)
    np.testing.assert_array_equal(got["term"], ora._terminal_t_buf)


def test_uniform_sample_uses_numpy_global_rng():
    from hanabi_agents_b200.rlax_dqn import ExperienceBuffer
    buf = ExperienceBuffer(40, 5, 1, 64)
    rng = np.random.default_rng(0)
    buf.add_transitions(*_rand_tr(rng, 20, 40, 5))
    np.random.seed(3)
    tr = buf.sample(8)
    np.random.seed(3)
    idx = np.random.choice(np.arange(64)[:20], size=8)
    np.testing.assert_array_equal(tr.observation_tm1, buf[idx].observation_tm1)


def test_priority_buffer_against_the_reference_run():
    """The CUDA shard through the numpy API against fixtures recorded from the reference's OWN
    priority_buffer.py + sum tree (tests/golden/make_priority_golden.py): leaves, total, max / min
    priority, cursor, sampled indices and probabilities, ring contents after every operation,
    including wrap-around and duplicate indices in one update."""
    from hanabi_agents_b200.rlax_dqn import PriorityBuffer
    from test_replay_oracle import PGOLD as G, replay_priority_fixture
    cap, tree_cap = int(G["cap"]), int(G["tree_cap"])
    buf = PriorityBuffer(int(G["L"]), int(G["A"]), 1, cap, float(G["alpha"]))
    assert buf.sum_tree.get_capacity() == tree_cap
    replay_priority_fixture(
        buf, lambda b: np.array(b.sum_tree.get_values(list(range(tree_cap))), np.float32),
        lambda b, B, u: b.sample_batch(B, u), _all_rows)


@pytest.mark.parametrize("B,cap", [(256, 5000), (32, 300), (1000, 70000)])
def test_fused_sample_gather_matches_the_separate_calls(B, cap):
    """hb_rb_sample_gather (one launch: descent + probabilities + importance weights + row gather, the
    learner's front end) against hb_rb_sample + hb_rb_gather and the float32 statements of
    rlax_rainbow.py:188-189 on their output: indices / probabilities / rows bit-exact, weights to 1 ulp
    (powf on the device vs numpy)."""
    import torch
    from hanabi_agents_b200 import _lib
    from hanabi_agents_b200.rlax_dqn import PriorityBuffer
    L, A = 658, 20
    rng = np.random.default_rng(B)
    buf = PriorityBuffer(L, A, 1, cap)
    for _ in range(3):
        buf.add_transitions(*_rand_tr(rng, cap // 3, L, A))
        idx0, _, _ = buf.sample_batch(64, rng.uniform(0, 1 / 64, 64))
        buf.update_priorities(idx0, np.abs(rng.standard_normal(64)).astype(np.float32) * 2)
    u = rng.uniform(0, 1.0 / B, B)
    idx, p64, p32, (o1, a, r, o2, lm, t) = buf.sample_device(B, u)
    keys = torch.as_tensor(PriorityBuffer.stratified_keys(B, u).astype(np.float32), device="cuda")
    dyn_in = torch.tensor([0.4, 0.99, 1e-3, 0.9, 0.999, 3e-5, 1.0, 0.0], device="cuda")
    dyn_out = torch.zeros(8, device="cuda")
    e = lambda shape, dt: torch.zeros(shape, dtype=dt, device="cuda")  # noqa: E731
    f_idx, f_p32, f_p64, f_w = e(B, torch.int64), e(B, torch.float32), e(B, torch.float64), e(B, torch.float32)
    f_o1, f_o2 = e((B, buf.obs_stride), torch.int32), e((B, buf.obs_stride), torch.int32)
    f_a, f_r, f_lm, f_t = e(B, torch.int32), e(B, torch.float32), e((B, buf.legal_words), torch.int32), e(B, torch.uint8)
    _lib.check(_lib.lib().hb_rb_sample_gather(
        buf._h, keys.data_ptr(), B, dyn_in.data_ptr(), dyn_out.data_ptr(), 0.0, f_idx.data_ptr(), f_p32.data_ptr(),
        f_p64.data_ptr(), f_w.data_ptr(), f_o1.data_ptr(), f_a.data_ptr(), f_r.data_ptr(), f_o2.data_ptr(),
        f_lm.data_ptr(), f_t.data_ptr(), 0, _lib.current_stream_ptr()))
    _lib.check(_lib.lib().hb_rb_check(buf._h, _lib.current_stream_ptr()))
    for got, want in ((f_idx, idx), (f_p64, p64), (f_p32, p32), (f_o1, o1), (f_o2, o2), (f_a, a), (f_r, r),
                      (f_lm, lm), (f_t, t), (dyn_out, dyn_in)):
        assert torch.equal(got, want)
    w = np.power(np.float32(1.0) / p32.cpu().numpy(), np.float32(0.4)).astype(np.float32)
    w = w / w.max()
    np.testing.assert_allclose(f_w.cpu().numpy(), w, rtol=3e-7, atol=0)
