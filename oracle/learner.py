"""CPU restatement (torch, fp32) of the reference's Q-network, policy and learner step.

TEST INFRASTRUCTURE -- not part of the product.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module.

PARITY UNPINNED: the reference implements this path with jax / dm-haiku / rlax /
jax.experimental.optix (no versions pinned anywhere; none of them is installed here) and holds no
test for the network, the loss or the optimiser.  This file restates the published algorithms at
the reference's own call sites (path:line under /root/reference/hanabi_agents):
  NoisyLinear / NoisyMLP         rlax_dqn/noisy_mlp.py:49-91, 154-185
  DQNPolicy.policy / eval_policy rlax_dqn/rlax_rainbow.py:34-59, 93-150
  categorical double-Q loss      rlax_dqn/rlax_rainbow.py:172-200  (rlax.categorical_double_q_learning,
                                 rlax.categorical_l2_project)
  double-Q + L2 loss             rlax_dqn/rlax_dqn.py:170-205       (rlax.double_q_learning, rlax.l2_loss)
  Adam                           rlax_dqn/rlax_rainbow.py:215-216,257 (optix.adam)
The one known answer the reference holds for this path -- the worked example of the C51
projection in rainbow/rainbow_agent.py:262-266,400-401 -- is checked in tests/test_learner_oracle.py.

Noise: NoisyLinear draws eps inside a jitted trace (noisy_mlp.py:79,87), so eps are constants and
a layer is  y = x (w + w_mu + w_sigma*eps_w) + (b + b_mu + b_sigma*eps_b).

`mirror_fp16`: the CUDA path feeds fp16 operands to the tensor cores (effective weights and hidden
activations rounded to fp16, fp32 accumulation).  With mirror_fp16=True this restatement rounds at
the same points, which is what the 1e-3 parity tests compare against; with False it is plain fp32.
"""
import numpy as np
import torch

KINDS = ("w", "b", "w_mu", "b_mu", "w_sigma", "b_sigma")


def truncated_normal(rng, shape, stddev):
    """hk.initializers.TruncatedNormal(stddev): N(0,1) truncated to [-2, 2], scaled."""
    out = rng.standard_normal(shape)
    bad = np.abs(out) > 2
    while bad.any():
        out[bad] = rng.standard_normal(int(bad.sum()))
        bad = np.abs(out) > 2
    return (out * stddev).astype(np.float32)


def init_params(obs_len, hidden, num_actions, n_atoms, noisy=True, seed=1234):
    """Parameters + noise constants, in the reference's layout (w: [in, out])."""
    rng = np.random.default_rng(seed)
    sizes = [obs_len] + list(hidden) + [num_actions * n_atoms]
    layers = []
    for i in range(len(sizes) - 1):
        fi, fo = sizes[i], sizes[i + 1]
        sd = 1.0 / np.sqrt(fi)                       # noisy_mlp.py:57-78
        p = {"w": truncated_normal(rng, (fi, fo), sd), "b": np.zeros(fo, np.float32)}
        if noisy:
            p["w_mu"] = truncated_normal(rng, (fi, fo), sd)
            p["w_sigma"] = truncated_normal(rng, (fi, fo), sd)
            p["b_mu"] = np.zeros(fo, np.float32)
            p["b_sigma"] = np.zeros(fo, np.float32)
            p["eps_w"] = rng.standard_normal((fi, fo)).astype(np.float32)
            p["eps_b"] = rng.standard_normal(fo).astype(np.float32)
        layers.append(p)
    return layers


def _round16(x):
    return x.half().float()


def _ste16(x):
    """fp16 rounding with a straight-through gradient."""
    return x + (_round16(x) - x).detach()


def categorical_l2_project(z_p, probs, z_q):
    """rlax.categorical_l2_project (1-D): projects (z_p, probs) onto the support z_q."""
    kq = z_q.shape[0]
    d_pos = torch.roll(z_q, -1) - z_q
    d_neg = z_q - torch.roll(z_q, 1)
    z_p = torch.clamp(z_p, z_q[0], z_q[-1])[None, :]
    z_qc = z_q[:, None]
    d_pos = d_pos[:, None]
    d_neg = d_neg[:, None]
    d_neg = torch.where(d_neg > 0, 1.0 / d_neg, torch.zeros_like(d_neg))
    d_pos = torch.where(d_pos > 0, 1.0 / d_pos, torch.zeros_like(d_pos))
    delta_qp = z_p - z_qc
    d_sign = (delta_qp >= 0).to(probs.dtype)
    delta_hat = d_sign * delta_qp * d_pos - (1.0 - d_sign) * delta_qp * d_neg
    assert delta_hat.shape[0] == kq
    return torch.sum(torch.clamp(1.0 - delta_hat, 0.0, 1.0) * probs[None, :], dim=-1)


class OracleQNet:
    def __init__(self, layers, num_actions, n_atoms=51, vmax=25.0, noisy=True, mirror_fp16=True):
        self.noisy, self.A, self.n_atoms, self.vmax = noisy, num_actions, n_atoms, float(vmax)
        self.mirror = mirror_fp16
        self.params = [{k: torch.tensor(np.asarray(v, np.float32)) for k, v in l.items()
                        if not k.startswith("eps")} for l in layers]
        self.eps = [{k: torch.tensor(np.asarray(v, np.float32)) for k, v in l.items()
                     if k.startswith("eps")} for l in layers]
        self.target = None
        self.sync_target()
        self.m = [{k: torch.zeros_like(v) for k, v in l.items()} for l in self.params]
        self.v = [{k: torch.zeros_like(v) for k, v in l.items()} for l in self.params]
        self.t = 0
        self.atoms = torch.linspace(-self.vmax, self.vmax, n_atoms) if n_atoms > 1 else None

    # -- network --------------------------------------------------------------------------
    def effective(self, params=None):
        params = self.params if params is None else params
        out = []
        for p, e in zip(params, self.eps):
            if self.noisy:
                W = p["w"] + p["w_mu"] + p["w_sigma"] * e["eps_w"]
                b = p["b"] + p["b_mu"] + p["b_sigma"] * e["eps_b"]
            else:
                W, b = p["w"], p["b"]
            out.append((W, b))
        return out

    def sync_target(self):
        self.target = [(W.detach().clone(), b.detach().clone()) for W, b in self.effective()]

    def apply(self, eff, x):
        """network.apply: x [M, L] 0/1 -> logits [M, A, n_atoms] (or q [M, A] when n_atoms == 1)."""
        h = x.float()
        for i, (W, b) in enumerate(eff):
            Wc = _ste16(W) if self.mirror else W
            h = h @ Wc + b
            if i < len(eff) - 1:
                h = torch.relu(h)
                if self.mirror:
                    h = _ste16(h)
        return h.view(h.shape[0], self.A, self.n_atoms) if self.n_atoms > 1 else h

    def q_values(self, logits):
        if self.n_atoms == 1:
            return logits
        probs = torch.softmax(logits, -1)
        return torch.mean(probs * self.atoms, -1)              # rlax_rainbow.py:117-118 (mean!)

    # -- policy (rlax_rainbow.py:34-59, 93-150) ---------------------------------------------
    def masked_q(self, obs, legal, target=False):
        with torch.no_grad():
            q = self.q_values(self.apply(self.target if target else self.effective(), obs))
        return torch.where(legal.bool(), q, torch.full_like(q, -float("inf")))

    @staticmethod
    def explore_from_q(q, legal, epsilon, u):
        """legal_epsilon_greedy(eps).sample with the uniforms given (u ~ U[0, 0.999))."""
        legal = legal.float()
        greedy = (q == q.max(-1, keepdim=True).values).float()
        greedy = greedy / greedy.sum(-1, keepdim=True)
        probs = (1 - epsilon) * greedy + epsilon * legal / legal.sum(-1, keepdim=True)
        cpi = torch.zeros_like(probs)
        c = torch.zeros(probs.shape[0])
        for a in range(probs.shape[1]):                       # sequential fp32 cumsum
            c = c + probs[:, a]
            cpi[:, a] = c
        cpi[:, -1] = 1.0
        gt = (u[:, None] > cpi)
        return torch.argmin(gt.int(), -1)

    @staticmethod
    def greedy_from_q(q, u):
        """rlax.greedy().sample: uniform among the maxima; pick = floor(u * n_max)."""
        is_max = q == q.max(-1, keepdim=True).values
        n_max = is_max.sum(-1)
        pick = torch.minimum((u * n_max.float()).floor().long(), n_max - 1)
        rank = torch.cumsum(is_max.int(), -1) - 1
        return torch.argmax((is_max & (rank == pick[:, None])).int(), -1)

    # -- learner --------------------------------------------------------------------------
    def loss(self, obs_tm1, a_tm1, r_t, obs_t, term_t, prios=None, weights=None, beta=0.4,
             discount=0.99, mask_terminal=False, tf_variant=False, legal_t=None):
        """Returns (mean_loss, per_sample) with the autograd graph attached to self.params.
        tf_variant: the target of rainbow/rainbow_agent.py:183-210 -- the next action is the argmax of the
        TARGET network's q = sum(p z) over the legal next moves (`legal_t` 0/1), gamma masked by term_t."""
        eff = self.effective()
        B = obs_tm1.shape[0]
        if prios is not None:
            w = (1.0 / prios.float()) ** beta                 # rlax_rainbow.py:188-189
            w = w / w.max()
        elif weights is not None:
            w = weights.float()
        else:
            w = torch.ones(B)
        if self.n_atoms > 1:
            logits_tm1 = self.apply(eff, obs_tm1)
            with torch.no_grad():
                logits_t = self.apply(self.target, obs_t)
                if tf_variant:
                    q_sel = torch.sum(torch.softmax(logits_t, -1) * self.atoms, -1)      # :199-201
                    q_sel = torch.where(legal_t.bool(), q_sel, torch.full_like(q_sel, -float("inf")))
                    mask_terminal = True
                else:
                    q_sel = self.q_values(self.apply(eff, obs_t))
            losses = []
            for b in range(B):
                a_star = int(torch.argmax(q_sel[b]))
                g = 0.0 if (mask_terminal and bool(term_t[b])) else discount
                target_z = r_t[b] + g * self.atoms
                p = torch.softmax(logits_t[b, a_star], -1)
                m = categorical_l2_project(target_z, p, self.atoms).detach()
                logp = torch.log_softmax(logits_tm1[b, int(a_tm1[b])], -1)
                losses.append(-(m * logp).sum())
            per = torch.stack(losses)
            return torch.mean(per * w), per
        q_tm1 = self.apply(eff, obs_tm1)
        with torch.no_grad():
            q_t = self.apply(self.target, obs_t)
            q_t = torch.where(term_t.bool()[:, None], torch.zeros_like(q_t), q_t)   # rlax_dqn.py:178
            q_sel = self.apply(eff, obs_t)
            a_star = torch.argmax(q_sel, -1)
            target = r_t + discount * q_t[torch.arange(B), a_star]
        td = target - q_tm1[torch.arange(B), a_tm1.long()]
        return torch.mean(w * 0.5 * td * td), td

    def grads(self, *args, **kw):
        for l in self.params:
            for v in l.values():
                v.requires_grad_(True)
                v.grad = None
        mean_loss, per = self.loss(*args, **kw)
        mean_loss.backward()
        g = [{k: v.grad.detach().clone() for k, v in l.items()} for l in self.params]
        for l in self.params:
            for v in l.values():
                v.requires_grad_(False)
        return float(mean_loss.detach()), per.detach(), g

    def adam_tf(self, grads, lr=2.5e-5, b1=0.9, b2=0.999, eps=3.125e-5):
        """tf.train.AdamOptimizer: lr_t = lr sqrt(1 - b2^t) / (1 - b1^t); theta -= lr_t m / (sqrt(v) + eps)."""
        self.t += 1
        lr_t = lr * (1 - b2 ** self.t) ** 0.5 / (1 - b1 ** self.t)
        with torch.no_grad():
            for p, g, m, v in zip(self.params, grads, self.m, self.v):
                for k in p:
                    m[k] = b1 * m[k] + (1 - b1) * g[k]
                    v[k] = b2 * v[k] + (1 - b2) * g[k] * g[k]
                    p[k] -= lr_t * m[k] / (torch.sqrt(v[k]) + eps)

    def adam(self, grads, lr=1e-3, b1=0.9, b2=0.999, eps=3.125e-5):
        """optix.adam: scale_by_adam (bias-corrected) then scale(-lr)."""
        self.t += 1
        bc1, bc2 = 1 - b1 ** self.t, 1 - b2 ** self.t
        with torch.no_grad():
            for p, g, m, v in zip(self.params, grads, self.m, self.v):
                for k in p:
                    m[k] = b1 * m[k] + (1 - b1) * g[k]
                    v[k] = b2 * v[k] + (1 - b2) * g[k] * g[k]
                    p[k] -= lr * (m[k] / bc1) / (torch.sqrt(v[k] / bc2) + eps)
