"""GPU parity of the Q-network path (csrc/qnet.cu through the C ABI) against oracle/learner.py on
the same parameters, noise constants, observations and uniforms.

Tolerances (north star: "Q-values and losses must match within 1e-3 relative"): q-values, logits
and losses rtol 1e-3 against the fp16-mirroring restatement (same fp16-rounded effective weights
and hidden activations, fp32 accumulation), with an absolute floor of 1e-3 x the row scale;
gradients 1e-2 in Frobenius norm (the backward GEMMs take fp16 dL/dz operands, which the
restatement does not round) and 1e-1 of the max-norm (ReLU units at ~0 may flip)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def pack_bits(x01, stride_words):
    """[n, L] 0/1 -> [n, stride_words] uint32, bit b at word b>>5, position b&31."""
    n, L = x01.shape
    pad = np.zeros((n, stride_words * 32), np.uint8)
    pad[:, :L] = x01
    return np.packbits(pad, axis=1, bitorder="little").view("<u4").reshape(n, stride_words).copy()


def _setup(obs_len, hidden, A, n_atoms, noisy, seed=1, vmax=25.0):
    import torch
    from hanabi_agents_b200.rlax_dqn.network import QNetwork, init_params
    from oracle.learner import OracleQNet
    stride = ((obs_len + 31) // 32 + 3) // 4 * 4
    layers = init_params(obs_len, hidden, A, n_atoms, noisy=noisy, seed=seed)
    # non-trivial biases so that the bias paths are exercised
    rng = np.random.default_rng(seed + 100)
    for p in layers:
        for k in p:
            if k.startswith("b"):
                p[k] = (0.1 * rng.standard_normal(p[k].shape)).astype(np.float32)
    net = QNetwork(obs_len, stride, hidden, A, n_atoms, vmax, noisy)
    net.set_params(layers)
    ora = OracleQNet(layers, A, n_atoms, vmax, noisy=noisy, mirror_fp16=True)
    return net, ora, stride


def _obs(rng, n, L, stride, p=0.15):
    import torch
    x = (rng.random((n, L)) < p).astype(np.uint8)
    return x, torch.as_tensor(pack_bits(x, stride).view(np.int32), device="cuda")


CONFIGS = [  # obs_len, hidden, A, n_atoms, noisy
    (658, [512, 512], 20, 51, True),     # cfg 4: Rainbow, Hanabi-Full 2p, 512x512
    (171, [512], 11, 51, True),          # cfg 1: Hanabi-Small 2p, default layers
    (1280, [512], 48, 51, True),         # cfg 5: Hanabi-Full 5p
    (658, [512], 20, 1, False),          # plain double-DQN head (rlax_dqn.py)
]


@pytest.mark.parametrize("obs_len,hidden,A,n_atoms,noisy", CONFIGS)
def test_forward_parity(obs_len, hidden, A, n_atoms, noisy):
    import torch
    net, ora, stride = _setup(obs_len, hidden, A, n_atoms, noisy)
    rng = np.random.default_rng(0)
    x, xp = _obs(rng, 300, obs_len, stride)
    q, logits = net.forward(xp, want_logits=True)
    torch.cuda.synchronize()
    ref_logits = ora.apply(ora.effective(), torch.tensor(x))
    ref_q = ora.q_values(ref_logits)
    if n_atoms > 1:
        scale = float(ref_logits.abs().max())
        torch.testing.assert_close(logits.cpu(), ref_logits, rtol=1e-3, atol=1e-3 * scale)
    scale = float(ref_q.abs().max())
    torch.testing.assert_close(q.cpu(), ref_q, rtol=1e-3, atol=1e-3 * scale)
    # target network = snapshot of the online one
    qt = net.forward(xp, use_target=True)
    assert torch.equal(qt, q)


def test_effective_weights_roundtrip_and_layout():
    net, ora, _ = _setup(658, [512, 512], 20, 51, True)
    got = net.get_params(with_noise=True)
    for g, p, e in zip(got, ora.params, ora.eps):
        for k in p:
            np.testing.assert_array_equal(g[k], p[k].numpy())
        for k in e:
            np.testing.assert_array_equal(g[k], e[k].numpy())


@pytest.mark.parametrize("obs_len,hidden,A,n_atoms,noisy", CONFIGS[:2] + CONFIGS[3:])
def test_policy_parity(obs_len, hidden, A, n_atoms, noisy):
    import torch
    net, ora, stride = _setup(obs_len, hidden, A, n_atoms, noisy)
    rng = np.random.default_rng(1)
    n = 2000
    x, xp = _obs(rng, n, obs_len, stride)
    legal = (rng.random((n, A)) < 0.6).astype(np.uint8)
    legal[np.arange(n), rng.integers(0, A, n)] = 1
    lp = torch.as_tensor(pack_bits(legal, (A + 31) // 32).view(np.int32), device="cuda")
    u = torch.as_tensor((rng.random(n) * 0.999).astype(np.float32), device="cuda")
    for eps in (0.0, 0.1, 1.0):
        act, q = net.act(xp, lp, greedy=False, epsilon=eps, uniform=u, want_q=True)
        torch.cuda.synchronize()
        qc = q.cpu()
        assert torch.isinf(qc[torch.tensor(legal) == 0]).all()
        # the policy arithmetic is exact given the same q: apply the restated policy to the device q
        want = ora.explore_from_q(qc, torch.tensor(legal), eps, u.cpu())
        assert torch.equal(act.cpu().long(), want)
        assert (torch.tensor(legal)[torch.arange(n), act.cpu().long()] == 1).all()
    act = net.act(xp, lp, greedy=True, uniform=u)
    want = ora.greedy_from_q(qc, u.cpu())
    assert torch.equal(act.cpu().long(), want)
    # and the device q itself agrees with the restatement
    ref_q = ora.masked_q(torch.tensor(x), torch.tensor(legal))
    fin = torch.isfinite(ref_q)
    torch.testing.assert_close(qc[fin], ref_q[fin], rtol=1e-3, atol=1e-3 * float(ref_q[fin].abs().max()))
    # built-in uniform stream: legal actions, epsilon respected
    act2 = net.act(xp, lp, greedy=False, epsilon=1.0, seed=7, counter=3)
    assert (torch.tensor(legal)[torch.arange(n), act2.cpu().long()] == 1).all()
    assert len(set(act2.cpu().tolist())) > 3


def _batch(rng, B, obs_len, stride, A):
    import torch
    x1, x1p = _obs(rng, B, obs_len, stride)
    x2, x2p = _obs(rng, B, obs_len, stride)
    a = rng.integers(0, A, B).astype(np.int32)
    r = rng.integers(-3, 2, B).astype(np.float32)
    term = (rng.random(B) < 0.2).astype(np.uint8)
    prios = (rng.random(B) * 2.0 / B + 1e-4).astype(np.float32)
    dev = lambda v: torch.as_tensor(v, device="cuda")  # noqa: E731
    return (x1, x2, a, r, term, prios), (x1p, dev(a), dev(r), x2p, dev(term), dev(prios))


@pytest.mark.parametrize("obs_len,hidden,A,n_atoms,noisy,B", [c + (b,) for c, b in zip(CONFIGS, (256, 32, 256, 256))])
def test_learner_step_parity(obs_len, hidden, A, n_atoms, noisy, B):
    import torch
    net, ora, stride = _setup(obs_len, hidden, A, n_atoms, noisy, seed=3)
    rng = np.random.default_rng(2)
    # make the target differ from the online network: one Adam step on both, target left behind
    lr, eps = 1e-3, 3.125e-5
    for it in range(3):
        (x1, x2, a, r, term, prios), dev = _batch(rng, B, obs_len, stride, A)
        t = lambda v: torch.tensor(v)  # noqa: E731
        kw = dict(prios=t(prios), beta=0.4, discount=0.99) if n_atoms > 1 else dict(weights=None, discount=0.99)
        mean_ref, per_ref, g_ref = ora.grads(t(x1), t(a), t(r), t(x2), t(term), **kw)
        if n_atoms > 1:
            loss, mean = net.grads(dev[0], dev[1], dev[2], dev[3], dev[4], prios=dev[5], beta=0.4, discount=0.99)
        else:
            loss, mean = net.grads(dev[0], dev[1], dev[2], dev[3], dev[4], discount=0.99)
        torch.cuda.synchronize()
        torch.testing.assert_close(loss.cpu(), per_ref, rtol=1e-3, atol=1e-3 * float(per_ref.abs().max()))
        assert abs(float(mean) - mean_ref) <= 1e-3 * abs(mean_ref) + 1e-6
        got = net.get_params(grads=True)
        for l, (gg, gr) in enumerate(zip(got, g_ref)):
            for k_dev, k_ref in (("dw", "w"), ("db", "b")):
                ref = gr[k_ref].numpy()
                # 1 % in Frobenius norm.  The max-norm bound is looser: a hidden unit whose
                # pre-activation is ~0 can land on different sides of the ReLU in the two fp
                # evaluations, which moves one row of dW by that unit's whole contribution.
                err = gg[k_dev] - ref
                assert np.linalg.norm(err) <= 1e-2 * np.linalg.norm(ref) + 1e-12, (it, l, k_dev)
                assert np.abs(err).max() <= 1e-1 * max(np.abs(ref).max(), 1e-12), (it, l, k_dev)
        net.adam(lr=lr, eps=eps)
        if it == 0:
            # identical starting state: the Adam update itself can be compared (units of lr)
            ora.adam(g_ref, lr=lr, eps=eps)
            for l, (gp, rp) in enumerate(zip(net.get_params(), ora.params)):
                for k in rp:
                    assert np.abs(gp[k] - rp[k].numpy()).max() <= 0.35 * lr, (l, k)
        # later iterations start from the device's parameters (online != target from here on)
        for gp, rp in zip(net.get_params(), ora.params):
            for k in rp:
                rp[k] = torch.tensor(gp[k])
    # after the updates the online net differs from the target snapshot
    x, xp = _obs(rng, 64, obs_len, stride)
    assert not torch.equal(net.forward(xp), net.forward(xp, use_target=True))
    net.sync_target()
    assert torch.equal(net.forward(xp), net.forward(xp, use_target=True))


def test_adam_multi_step_against_oracle():
    """Known gradients written into the flat bucket: fused Adam + fan-out (dw = dw_mu = dW,
    dw_sigma = dW * eps) + bias correction over several steps."""
    import torch
    net, ora, _ = _setup(171, [512], 11, 51, True, seed=5)
    g = torch.Generator(device="cuda").manual_seed(0)
    for step in range(4):
        net.grad_bucket.copy_(torch.randn(net.n_grad, device="cuda", generator=g) * 1e-3)
        torch.cuda.synchronize()
        eff = net.get_params(grads=True)
        grads = []
        for ge, e in zip(eff, ora.eps):
            dw, db = torch.tensor(ge["dw"]), torch.tensor(ge["db"])
            grads.append({"w": dw, "w_mu": dw, "w_sigma": dw * e["eps_w"],
                          "b": db, "b_mu": db, "b_sigma": db * e["eps_b"]})
        net.adam(lr=1e-3, eps=3.125e-5, grad_scale=0.5)
        ora.adam([{k: 0.5 * v for k, v in gl.items()} for gl in grads], lr=1e-3, eps=3.125e-5)
        for gp, rp in zip(net.get_params(), ora.params):
            for k in rp:
                np.testing.assert_allclose(gp[k], rp[k].numpy(), rtol=1e-5, atol=2e-7)


def test_learner_step_with_changing_batch_sizes():
    """One network, batches of 256, 40 and 256 rows again: the forward chain's row-block counters
    (csrc/tc_chain.cuh) are per batch geometry; losses must match the oracle each time."""
    import torch
    obs_len, hidden, A, n_atoms = CONFIGS[0][:4]
    net, ora, stride = _setup(obs_len, hidden, A, n_atoms, True, seed=7)
    rng = np.random.default_rng(9)
    t = lambda v: torch.tensor(v)  # noqa: E731
    for B in (256, 40, 256, 40):
        (x1, x2, a, r, term, prios), dev = _batch(rng, B, obs_len, stride, A)
        mean_ref, per_ref, _ = ora.grads(t(x1), t(a), t(r), t(x2), t(term), prios=t(prios), beta=0.4, discount=0.99)
        loss, mean = net.grads(dev[0], dev[1], dev[2], dev[3], dev[4], prios=dev[5], beta=0.4, discount=0.99)
        torch.cuda.synchronize()
        torch.testing.assert_close(loss.cpu(), per_ref, rtol=1e-3, atol=1e-3 * float(per_ref.abs().max()))
        assert abs(float(mean) - mean_ref) <= 1e-3 * abs(mean_ref) + 1e-6


def test_learner_step_parity_on_the_per_layer_paths():
    """The learner step has switches that restore the round-1 launch structure (one GEMM launch per layer
    instead of mlp_chain_kernel, the dense head backward GEMM instead of the product inside the C51 loss,
    the last dZ product untransposed): the same parity tests must hold there.  The switches are read once
    per process, hence the child process."""
    import os
    import subprocess
    import sys
    if os.environ.get("HB_FWD_CHAIN") == "0":
        pytest.skip("already inside the child")
    env = dict(os.environ, HB_FWD_CHAIN="0", HB_DENSE_HEAD_BW="1", HB_DZ0_PLAIN="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-q", "-x", "-m", "gpu",
                        "-k", "test_learner_step_parity and not per_layer_paths", "-p", "no:cacheprovider"],
                       env=env, capture_output=True, text=True, timeout=600,
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout
