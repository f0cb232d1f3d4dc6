"""Pins for oracle/replay.py: fixtures produced by the reference's own ExperienceBuffer
(tests/golden/make_replay_golden.py) and the cases of the reference's
tests/rlax_dqn/test_experience_buffer.py / test_priority_buffer.py."""
import os

import numpy as np

from oracle.replay import OracleExperienceBuffer, OraclePriorityBuffer, priority_pow

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "replay_ref.npz"))


def _state(buf):
    return dict(obs_tm1=buf._obs_tm1_buf, obs_t=buf._obs_t_buf, act=buf._act_tm1_buf, rew=buf._rew_t_buf,
                lms=buf._lms_t_buf, term=buf._terminal_t_buf)


def test_against_reference_fixtures():
    L, A, cap = int(GOLD["L"]), int(GOLD["A"]), int(GOLD["cap"])
    buf = OracleExperienceBuffer(L, A, 1, cap)
    for k, n in enumerate(GOLD["sizes"]):
        assert buf.get_update_indices(int(n)) == GOLD["upd%d" % k].tolist()
        buf.add_transitions(GOLD["in%d_o1" % k], GOLD["in%d_a" % k], GOLD["in%d_r" % k],
                            GOLD["in%d_o2" % k], GOLD["in%d_lm" % k], GOLD["in%d_t" % k])
        assert buf.size == int(GOLD["st%d_size" % k]) and buf.oldest_entry == int(GOLD["st%d_oldest" % k])
        for name, arr in _state(buf).items():
            np.testing.assert_array_equal(arr, GOLD["st%d_%s" % (k, name)], err_msg="%d %s" % (k, name))
    tr = buf[GOLD["get_idx"]]
    for got, name in zip(tr, ("observation_tm1", "action_tm1", "reward_t", "observation_t",
                              "legal_moves_t", "terminal_t")):
        np.testing.assert_array_equal(got, GOLD["get_" + name])


def test_reference_unit_cases():
    """tests/rlax_dqn/test_experience_buffer.py:23-142 (`cur_idx` there is today's `oldest_entry`)."""
    buf = OracleExperienceBuffer(5, 3, 1, 7)
    assert buf._obs_tm1_buf.shape == (7, 5) and buf._lms_t_buf.shape == (7, 3) and buf.size == 0
    mk = lambda n, v: (np.full((n, 5), v, np.int8), np.full((n, 1), v, np.int8), np.full((n, 1), v, float),  # noqa: E731
                       np.full((n, 5), v + 1, np.int8), np.ones((n, 3), np.int8), np.zeros((n, 1), bool))
    buf.add_transitions(*mk(3, 1))
    assert (buf.size, buf.oldest_entry) == (3, 3)
    buf.add_transitions(*mk(4, 2))                      # exact fill: size == capacity, cursor 0
    assert (buf.size, buf.oldest_entry) == (7, 0)
    buf.add_transitions(*mk(1, 3))                      # overflow by one: element overwrites slot 0
    assert (buf.size, buf.oldest_entry) == (7, 1)
    assert buf._obs_tm1_buf[0, 0] == 3 and buf._obs_tm1_buf[1, 0] == 1
    buf.add_transitions(*mk(7, 4))                      # wrap split
    assert (buf.size, buf.oldest_entry) == (7, 1) and (buf._obs_tm1_buf == 4).all()


def test_priority_buffer_semantics():
    """priority_buffer.py:13-52 (the reference's own test asserts a stale max_priority == 1.0,
    tests/rlax_dqn/test_priority_buffer.py:27; the code sets alpha = 0.6)."""
    rng = np.random.default_rng(0)
    pb = OraclePriorityBuffer(6, 4, 1, 10)
    assert pb.max_priority == 0.6 and pb.min_priority == 0.6 and pb.sum_tree.get_capacity() == 16
    n = 7
    tr = lambda n: ((rng.random((n, 6)) < .5).astype(np.int8), rng.integers(0, 4, (n, 1)).astype(np.int8),  # noqa: E731
                    rng.random((n, 1)), (rng.random((n, 6)) < .5).astype(np.int8),
                    np.ones((n, 4), np.int8), np.zeros((n, 1), bool))
    pb.add_transitions(*tr(n))
    assert abs(pb.sum_tree.get_total_val() - n * np.float32(0.6)) < 1e-5
    idx, prios, batch = pb.sample_batch(4, rng.uniform(0, 0.25, 4))
    assert len(idx) == 4 and all(0 <= i < n for i in idx) and idx == sorted(idx)
    np.testing.assert_allclose(prios, (np.float32(0.6) + 1e-10) / pb.sum_tree.get_total_val())
    assert batch[0].shape == (4, 6)
    pb.update_priorities(idx, np.array([0.5, 2.0, 0.0, 1.0], np.float32))
    assert abs(pb.max_priority - 2.0 ** 0.6) < 1e-6 and pb.min_priority < 1e-5
    pb.add_transitions(*tr(5))                          # wraps: new rows get the new max priority
    assert pb.oldest_entry == 2 and pb.size == 10
    np.testing.assert_allclose(pb.sum_tree.get_values([7, 8, 9, 0, 1]), [np.float32(pb.max_priority)] * 5)


def test_priority_pow_is_within_one_ulp_of_numpy_float32():
    x = np.abs(np.random.default_rng(1).standard_normal(4096)).astype(np.float32) * 3
    ours = priority_pow(x, 0.6)
    theirs = (x + 1e-10) ** 0.6 if False else np.power(x + np.float32(1e-10), np.float32(0.6))
    assert theirs.dtype == np.float32
    ulp = np.spacing(ours)
    assert np.all(np.abs(ours.astype(np.float64) - theirs.astype(np.float64)) <= ulp)


# ---- PriorityBuffer pinned on the REFERENCE ITSELF --------------------------------------------
# tests/golden/priority_ref.npz was recorded by running /root/reference/hanabi_agents/rlax_dqn/
# priority_buffer.py:13-52 on the reference's own pybind sum tree (tests/golden/make_priority_golden.py).
PGOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "priority_ref.npz"))


def replay_priority_fixture(buf, leaves_of, sample, state_of):
    """Drives `buf` (oracle or CUDA shard) through the recorded operation sequence and checks every
    recorded quantity bit for bit.  leaves_of(buf) -> float32[tree_cap]; sample(buf, B, uniforms) ->
    (indices, prios f64, 6 arrays); state_of(buf) -> dict of the six ring arrays."""
    G = PGOLD
    tree_cap = int(G["tree_cap"])
    for k, op in enumerate(G["ops"]):
        if op == "add":
            buf.add_transitions(G["a%d_o1" % k], G["a%d_a" % k], G["a%d_r" % k], G["a%d_o2" % k],
                                G["a%d_lm" % k], G["a%d_t" % k])
        elif op == "sample":
            B = int(G["p%d_B" % k])
            np.random.seed(int(G["p%d_seed" % k]))              # the draw priority_buffer.py:40 made
            u = np.random.uniform(size=(B,), high=1. / B)
            idx, prios, tr = sample(buf, B, u)
            assert list(idx) == G["p%d_idx" % k].tolist(), "op %d sampled indices" % k
            np.testing.assert_array_equal(np.asarray(prios, np.float64), G["p%d_prios" % k], err_msg="op %d prios" % k)
            for name, got in zip(("observation_tm1", "action_tm1", "reward_t", "observation_t",
                                  "legal_moves_t", "terminal_t"), tr):
                np.testing.assert_array_equal(got, G["p%d_%s" % (k, name)], err_msg="op %d %s" % (k, name))
        else:
            buf.update_priorities(G["u%d_idx" % k], G["u%d_td" % k])
        leaves = leaves_of(buf)
        assert leaves.shape == (tree_cap,)
        np.testing.assert_array_equal(leaves, G["s%d_leaves" % k], err_msg="op %d (%s) leaves" % (k, op))
        assert np.float32(buf.sum_tree.get_total_val()) == G["s%d_total" % k], "op %d total" % k
        assert np.float32(buf.max_priority) == np.float32(G["s%d_max" % k]), "op %d max_priority" % k
        assert np.float32(buf.min_priority) == np.float32(G["s%d_min" % k]), "op %d min_priority" % k
        assert (buf.size, buf.oldest_entry) == (int(G["s%d_size" % k]), int(G["s%d_oldest" % k])), k
        for name, arr in state_of(buf).items():
            np.testing.assert_array_equal(arr, G["s%d_%s" % (k, name)], err_msg="op %d ring %s" % (k, name))


def test_priority_buffer_against_the_reference_run():
    G = PGOLD
    cap, tree_cap = int(G["cap"]), int(G["tree_cap"])
    buf = OraclePriorityBuffer(int(G["L"]), int(G["A"]), 1, cap, float(G["alpha"]))
    assert buf.sum_tree.get_capacity() == tree_cap
    assert list(G["ops"]).count("sample") >= 5 and list(G["ops"]).count("update") >= 5
    replay_priority_fixture(
        buf, lambda b: np.array(b.sum_tree.get_values(list(range(tree_cap))), np.float32),
        lambda b, B, u: b.sample_batch(B, u), _state)
