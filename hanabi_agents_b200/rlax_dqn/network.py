"""Q-network handle: NoisyMLP / MLP forward, policy, gradients and Adam on the B200.

Thin ctypes wrapper of the hb_net_* entry points (include/hanabi_b200.h).  Mirrors what
DQNAgent keeps as `network`, `online_params`, `trg_params`, `optimizer`, `opt_state`
(hanabi_agents/rlax_dqn/rlax_rainbow.py:237-259).  Parameters are exchanged in the reference's
haiku layout (`w`: [in, out]), keyed like the haiku tree (`noisy_mlp/~/noisy_linear_%d`,
noisy_mlp.py:151) so that reference pickles can be loaded.
"""
import ctypes as C

import numpy as np

from .. import _lib

KIND = {"w": 0, "b": 1, "w_mu": 2, "b_mu": 3, "w_sigma": 4, "b_sigma": 5, "eps_w": 6, "eps_b": 7,
        "dw": 8, "db": 9}


def truncated_normal(rng, shape, stddev):
    """hk.initializers.TruncatedNormal(stddev) (noisy_mlp.py:57-78): N(0, 1) cut at +-2, scaled."""
    out = rng.standard_normal(shape)
    bad = np.abs(out) > 2
    while bad.any():
        out[bad] = rng.standard_normal(int(bad.sum()))
        bad = np.abs(out) > 2
    return (out * stddev).astype(np.float32)


def init_params(obs_len, hidden, num_actions, n_atoms, noisy=True, seed=1234):
    """Initial parameters and the trace-time noise constants, reference layout.

    The NOISE is the reference's own: NoisyMLP always seeds jax.random.PRNGKey(1234), whatever
    params.seed is, and draws the constants while it is traced (noisy_mlp.py:127,140,79,87); jax_prng
    restates that threefry stream, so a pickle written by the reference's save_weights evaluates to
    the reference's network after restore_weights.  The initial WEIGHTS come from haiku's initialisers
    on another jax stream (rlax_rainbow.py:250-251): their distributions are the reference's
    (truncated normal with stddev 1/sqrt(fan_in) for w, w_mu, w_sigma; zero biases), their stream is
    numpy's default_rng(seed)."""
    from .jax_prng import noisy_mlp_noise
    rng = np.random.default_rng(seed)
    sizes = [obs_len] + list(hidden) + [num_actions * n_atoms]
    noise = noisy_mlp_noise(sizes) if noisy else None
    layers = []
    for i in range(len(sizes) - 1):
        fi, fo = sizes[i], sizes[i + 1]
        sd = 1.0 / np.sqrt(fi)
        p = {"w": truncated_normal(rng, (fi, fo), sd), "b": np.zeros(fo, np.float32)}
        if noisy:
            p["w_mu"] = truncated_normal(rng, (fi, fo), sd)
            p["w_sigma"] = truncated_normal(rng, (fi, fo), sd)
            p["b_mu"] = np.zeros(fo, np.float32)
            p["b_sigma"] = np.zeros(fo, np.float32)
            p["eps_w"], p["eps_b"] = noise[i]
        layers.append(p)
    return layers


class _DevArray:
    """__cuda_array_interface__ view of library-owned device memory (zero copy into torch)."""

    def __init__(self, ptr, n, owner):
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class QNetwork:
    def __init__(self, obs_len, obs_stride, hidden, num_actions, n_atoms=51, atom_vmax=25.0, noisy=True,
                 device=None):
        import torch
        self.torch = torch
        if device is None:
            device = torch.cuda.current_device() if torch.cuda.is_available() else 0
        self.device_index = int(device)
        self.device = torch.device("cuda", self.device_index)
        self.obs_len, self.obs_stride = int(obs_len), int(obs_stride)
        self.hidden = [int(h) for h in hidden]
        self.A, self.n_atoms, self.noisy = int(num_actions), int(n_atoms), bool(noisy)
        self.legal_words = (self.A + 31) // 32
        hid = (C.c_int32 * max(1, len(self.hidden)))(*self.hidden)
        self._h = C.c_void_p()
        _lib.check(_lib.lib().hb_net_create(self.obs_len, self.obs_stride, hid, len(self.hidden), self.A,
                                            self.n_atoms, float(atom_vmax), int(self.noisy),
                                            self.device_index, C.byref(self._h)))
        L = _lib.lib()
        self.n_layers = int(L.hb_net_info(self._h, 0))
        self.n_grad = int(L.hb_net_info(self._h, 1))
        self.n_last = int(L.hb_net_info(self._h, 40 + self.n_layers - 1))
        self._grad = None

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                _lib.lib().hb_net_destroy(h)
            except Exception:
                pass
            self._h = None

    def _s(self):
        return _lib.current_stream_ptr()

    # -- parameters -----------------------------------------------------------------------
    def layer_shape(self, layer):
        L = _lib.lib()
        return int(L.hb_net_info(self._h, 10 + layer)), int(L.hb_net_info(self._h, 20 + layer))

    def set_params(self, layers, sync_target=True):
        """layers: list (one per layer) of dicts {w, b[, w_mu, b_mu, w_sigma, b_sigma, eps_w, eps_b]}."""
        if len(layers) != self.n_layers:
            raise ValueError("expected %d layers" % self.n_layers)
        for i, p in enumerate(layers):
            for k, v in p.items():
                a = np.ascontiguousarray(np.asarray(v, dtype=np.float32))
                _lib.check(_lib.lib().hb_net_set_param(self._h, i, KIND[k], a.ctypes.data, a.size))
        _lib.check(_lib.lib().hb_net_refresh(self._h, self._s()))
        if sync_target:
            self.sync_target()

    def get_params(self, with_noise=False, grads=False, target=False):
        out = []
        fn = _lib.lib().hb_net_get_target_param if target else _lib.lib().hb_net_get_param
        kinds = ["w", "b"] + (["w_mu", "b_mu", "w_sigma", "b_sigma"] if self.noisy else [])
        if with_noise and self.noisy:
            kinds += ["eps_w", "eps_b"]
        if grads:
            kinds = ["dw", "db"]
        for i in range(self.n_layers):
            fi, fo = self.layer_shape(i)
            p = {}
            for k in kinds:
                a = np.empty((fi, fo) if KIND[k] % 2 == 0 else (fo,), np.float32)
                _lib.check(fn(self._h, i, KIND[k], a.ctypes.data, a.size))
                p[k] = a
            out.append(p)
        return out

    def haiku_tree(self, target=False):
        """Parameters keyed like the reference's haiku tree (noisy_mlp.py:151)."""
        fmt = "noisy_mlp/~/noisy_linear_%d" if self.noisy else "mlp/~/linear_%d"
        return {fmt % i: p for i, p in enumerate(self.get_params(target=target))}

    def load_haiku_tree(self, tree, noise=None):
        layers = []
        for i in range(self.n_layers):
            key = [k for k in tree if k.endswith("_%d" % i)]
            if len(key) != 1:
                raise KeyError("layer %d not found in parameter tree" % i)
            p = {k: np.asarray(v, np.float32) for k, v in tree[key[0]].items()}
            if noise is not None:
                p.update(noise[i])
            layers.append(p)
        self.set_params(layers, sync_target=False)

    def sync_target(self):
        _lib.check(_lib.lib().hb_net_sync_target(self._h, self._s()))

    @property
    def grad_bucket(self):
        """Flat fp32 gradient bucket as a torch tensor (zero copy): what DDP-style training all-reduces."""
        if self._grad is None:
            ptr = _lib.lib().hb_net_grad(self._h)
            self._grad = self.torch.as_tensor(_DevArray(ptr, self.n_grad, self), device=self.device)
        return self._grad

    # -- compute --------------------------------------------------------------------------
    def act(self, obs_packed, legal_packed, greedy=False, epsilon=0.0, uniform=None, seed=0, counter=0,
            row0=0, use_target=False, want_q=False):
        torch = self.torch
        n = obs_packed.shape[0]
        action = torch.empty(n, dtype=torch.int32, device=self.device)
        q = torch.empty((n, self.A), dtype=torch.float32, device=self.device) if want_q else None
        _lib.check(_lib.lib().hb_net_act(
            self._h, obs_packed.data_ptr(), legal_packed.data_ptr(), self.legal_words, n, int(greedy),
            float(epsilon), None if uniform is None else uniform.data_ptr(), int(seed), int(counter),
            int(row0), int(use_target), action.data_ptr(), None if q is None else q.data_ptr(), self._s()))
        return (action, q) if want_q else action

    def forward(self, obs_packed, use_target=False, want_logits=False):
        torch = self.torch
        n = obs_packed.shape[0]
        q = torch.empty((n, self.A), dtype=torch.float32, device=self.device)
        logits = None
        if want_logits and self.n_atoms > 1:
            logits = torch.empty((n, self.n_last), dtype=torch.float32, device=self.device)
        _lib.check(_lib.lib().hb_net_forward(self._h, obs_packed.data_ptr(), n, int(use_target),
                                             q.data_ptr(), None if logits is None else logits.data_ptr(),
                                             self._s()))
        if logits is not None:
            logits = logits.view(n, self.A, 64)[:, :, :self.n_atoms]
        return (q, logits) if want_logits else q

    def grads(self, obs_tm1, a_tm1, r_t, obs_t, term_t=None, prios=None, weights=None, beta=0.4,
              discount=0.99, mask_terminal=False):
        """Forward + loss + backward; returns (per-sample loss / td [B], mean loss [1]) tensors."""
        torch = self.torch
        B = obs_tm1.shape[0]
        loss = torch.empty(B, dtype=torch.float32, device=self.device)
        mean = torch.empty(1, dtype=torch.float32, device=self.device)
        ptr = lambda t: None if t is None else t.data_ptr()  # noqa: E731
        _lib.check(_lib.lib().hb_net_grads(
            self._h, obs_tm1.data_ptr(), a_tm1.data_ptr(), r_t.data_ptr(), obs_t.data_ptr(), ptr(term_t),
            ptr(prios), ptr(weights), B, float(beta), float(discount), int(mask_terminal),
            loss.data_ptr(), mean.data_ptr(), self._s()))
        return loss, mean

    def adam(self, lr=1e-3, b1=0.9, b2=0.999, eps=3.125e-5, grad_scale=1.0):
        _lib.check(_lib.lib().hb_net_adam(self._h, float(lr), float(b1), float(b2), float(eps),
                                          float(grad_scale), self._s()))
