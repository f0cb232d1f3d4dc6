"""Pins for hanabi_agents_b200/rlax_dqn/jax_prng.py: the threefry stream behind the reference's
trace-time noise constants (hanabi_agents/rlax_dqn/noisy_mlp.py:127,140,79,87), restated without jax.
Known answers: the Random123 vectors jax tests itself with (random_test.py::testThreefry2x32) and the
values the jax documentation prints for PRNGKey(0) / PRNGKey(42)."""
import numpy as np

from hanabi_agents_b200.rlax_dqn import jax_prng as J


def _one(key, ctr):
    a, b = J.threefry2x32(np.uint32(key), np.uint32([ctr[0]]), np.uint32([ctr[1]]))
    return int(a[0]), int(b[0])


def test_threefry2x32_known_answers():
    assert _one([0, 0], [0, 0]) == (0x6b200159, 0x99ba4efe)
    assert _one([0xffffffff] * 2, [0xffffffff] * 2) == (0x1cb996fc, 0xbb002be7)
    assert _one([0x13198a2e, 0x03707344], [0x243f6a88, 0x85a308d3]) == (0xc4923a9c, 0x483df7a0)


def test_split_and_normal_match_the_jax_documentation():
    key = J.prng_key(0)
    assert key.tolist() == [0, 0] and J.prng_key(1234).tolist() == [0, 1234]
    ks = J.split(key)
    assert ks.tolist() == [[4146024105, 967050713], [2718843009, 1272950319]]
    assert J.normal(key, (1,))[0] == np.float32(-0.20584226)
    assert J.normal(ks[1], (1,))[0] == np.float32(-1.2515389)
    assert J.normal(J.prng_key(42), ())[()] == np.float32(-0.18471177)
    # odd sizes are padded inside threefry and cut again
    assert J.random_bits(key, (3,)).tolist() == [4146024105, 1351547692, 2718843009]


def test_noisy_mlp_noise_layout_and_statistics():
    sizes = [658, 512, 512, 20 * 51]
    noise = J.noisy_mlp_noise(sizes)
    assert [w.shape for w, _ in noise] == [(658, 512), (512, 512), (512, 1020)]
    assert [b.shape for _, b in noise] == [(512,), (512,), (1020,)]
    for w, b in noise:
        assert w.dtype == np.float32 and np.isfinite(w).all() and np.isfinite(b).all()
        assert abs(float(w.mean())) < 0.01 and abs(float(w.std()) - 1.0) < 0.01
    # one key per layer, weight and bias noise from consecutive draws of that layer's sequence
    rng = J.prng_key(1234)
    k0, rng = J.split(rng)
    seq, sub = J.split(k0)
    np.testing.assert_array_equal(noise[0][0], J.normal(sub, (658, 512)))
    seq, sub = J.split(seq)
    np.testing.assert_array_equal(noise[0][1], J.normal(sub, (512,)))
    # independent of params.seed: NoisyMLP's seed argument is never passed (rlax_rainbow.py:243)
    from hanabi_agents_b200.rlax_dqn.network import init_params
    a = init_params(171, [128], 11, 51, noisy=True, seed=1)
    b = init_params(171, [128], 11, 51, noisy=True, seed=2)
    np.testing.assert_array_equal(a[0]["eps_w"], b[0]["eps_w"])
    assert not np.array_equal(a[0]["w"], b[0]["w"])
