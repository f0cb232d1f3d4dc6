"""Pins for oracle/learner.py (torch-CPU restatement of the reference's learner math)."""
import numpy as np
import torch

from oracle.learner import OracleQNet, categorical_l2_project, init_params


def test_c51_projection_known_answer():
    """The worked example of rainbow/rainbow_agent.py:262-266,400-401."""
    supports = torch.tensor([[0., 2, 4, 6, 8], [1, 3, 4, 5, 6]])
    weights = torch.tensor([[0.1, 0.6, 0.1, 0.1, 0.1], [0.1, 0.2, 0.5, 0.1, 0.1]])
    target = torch.tensor([4., 5, 6, 7, 8])
    want = torch.tensor([[0.8, 0.0, 0.1, 0.0, 0.1], [0.8, 0.1, 0.1, 0.0, 0.0]])
    got = torch.stack([categorical_l2_project(supports[i], weights[i], target) for i in range(2)])
    torch.testing.assert_close(got, want, rtol=0, atol=1e-6)


def test_projection_conserves_mass_and_identity():
    z = torch.linspace(-25, 25, 51)
    p = torch.softmax(torch.randn(51, generator=torch.Generator().manual_seed(0)), -1)
    torch.testing.assert_close(categorical_l2_project(z, p, z), p, rtol=0, atol=1e-6)
    m = categorical_l2_project(1.0 + 0.99 * z, p, z)
    assert abs(float(m.sum()) - 1.0) < 1e-5


def test_policy_semantics():
    q = torch.tensor([[0.1, 0.5, 0.5, -float("inf")], [0.3, -float("inf"), 0.2, 0.1]])
    legal = torch.tensor([[1, 1, 1, 0], [1, 0, 1, 1]])
    # epsilon 0: only maxima; ties split evenly -> u < 0.5 picks the first maximum
    a = OracleQNet.explore_from_q(q, legal, 0.0, torch.tensor([0.25, 0.9]))
    assert a.tolist() == [1, 0]
    a = OracleQNet.explore_from_q(q, legal, 0.0, torch.tensor([0.75, 0.0]))
    assert a.tolist() == [2, 0]
    # epsilon 1: uniform over legal moves, never an illegal one
    u = torch.linspace(0, 0.998, 50)
    a = OracleQNet.explore_from_q(q[1:].repeat(50, 1), legal[1:].repeat(50, 1), 1.0, u)
    assert set(a.tolist()) == {0, 2, 3}
    g = OracleQNet.greedy_from_q(q, torch.tensor([0.6, 0.6]))
    assert g.tolist() == [2, 0]


def test_adam_matches_torch_optimizer():
    layers = init_params(20, [16], 4, 5, noisy=True, seed=3)
    net = OracleQNet(layers, 4, 5, 25.0, mirror_fp16=False)
    ref = [{k: v.clone().requires_grad_(True) for k, v in l.items()} for l in net.params]
    opt = torch.optim.Adam([v for l in ref for v in l.values()], lr=1e-3, betas=(0.9, 0.999), eps=3.125e-5)
    g = torch.Generator().manual_seed(1)
    for _ in range(3):
        grads = [{k: torch.randn(v.shape, generator=g) for k, v in l.items()} for l in net.params]
        for l, gl in zip(ref, grads):
            for k in l:
                l[k].grad = gl[k].clone()
        opt.step()
        net.adam(grads)
    for l, rl in zip(net.params, ref):
        for k in l:
            torch.testing.assert_close(l[k], rl[k].detach(), rtol=1e-5, atol=1e-6)


def test_noisy_gradient_fanout():
    """dw == dw_mu and dw_sigma == dw * eps."""
    layers = init_params(12, [8], 3, 5, noisy=True, seed=5)
    net = OracleQNet(layers, 3, 5, 25.0, mirror_fp16=False)
    g = torch.Generator().manual_seed(2)
    B = 6
    obs = (torch.rand(B, 12, generator=g) < 0.3).float()
    obs2 = (torch.rand(B, 12, generator=g) < 0.3).float()
    a = torch.randint(0, 3, (B,), generator=g)
    r = torch.randint(-1, 2, (B,), generator=g).float()
    _, per, grads = net.grads(obs, a, r, obs2, torch.zeros(B), prios=torch.full((B,), 1.0 / B))
    assert per.shape == (B,) and (per > 0).all()
    for gl, e in zip(grads, net.eps):
        torch.testing.assert_close(gl["w"], gl["w_mu"])
        torch.testing.assert_close(gl["w_sigma"], gl["w"] * e["eps_w"])
        torch.testing.assert_close(gl["b_sigma"], gl["b"] * e["eps_b"])


def test_double_q_td():
    layers = init_params(10, [8], 4, 1, noisy=False, seed=7)
    net = OracleQNet(layers, 4, 1, noisy=False, mirror_fp16=False)
    g = torch.Generator().manual_seed(4)
    B = 5
    obs = (torch.rand(B, 10, generator=g) < 0.5).float()
    obs2 = (torch.rand(B, 10, generator=g) < 0.5).float()
    a = torch.randint(0, 4, (B,), generator=g)
    r = torch.ones(B)
    term = torch.tensor([0, 1, 0, 0, 1])
    loss, td, _ = net.grads(obs, a, r, obs2, term, weights=torch.ones(B), discount=0.9)
    eff = net.effective()
    q1, q2 = net.apply(eff, obs), net.apply(eff, obs2)
    for b in range(B):
        boot = 0.0 if term[b] else 0.9 * float(q2[b, int(torch.argmax(q2[b]))])   # target == online here
        assert abs(float(td[b]) - (1.0 + boot - float(q1[b, a[b]]))) < 1e-5
    assert abs(loss - float((0.5 * td * td).mean())) < 1e-6
