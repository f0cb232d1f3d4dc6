"""DQNAgent with the reference's verbs (hanabi_agents/rlax_dqn/rlax_rainbow.py:220-365), running
on the B200: tcgen05 Q-network, device-resident replay shard, fused learner step.

Same constructor (`observation_spec`, `action_spec`, `params`), same methods
(`explore`, `exploit`, `add_experience_first`, `add_experience`, `update`, `save_weights`,
`restore_weights`, `requires_vectorized_observation`) and the same state (`train_step`,
`experience`, `params`).  `observations[1] == (obs, legal)` may be
  * bit-packed int32 CUDA tensors as HanabiParallelEnv emits them (zero copy), or
  * 0/1 arrays [N, L] / [N, A] (numpy or CUDA) as the reference's callers pass (packed on the way in).
Actions come back as a numpy array for numpy input (as the reference returns) and as a CUDA int32
tensor for CUDA input.

Data parallel: when torch.distributed is initialised, every rank owns its games and its replay
shard; `update()` all-reduces the flat gradient bucket once (the only collective) and scales by
1/world inside the fused Adam, so parameters stay identical on every rank.  Ranks must call
`update()` in lockstep (one collective per call).  Divergence from a single-process run at the same
global batch: every rank stratifies over ITS shard's tree and divides its
importance weights by ITS batch maximum (rlax_rainbow.py:188-189 on the local batch); the global
batch is the union of the per-rank batches, not one stratified draw over one tree.
"""
import os
import pickle

import numpy as np

from .buffers import ExperienceBuffer, PriorityBuffer, _is_cuda
from .network import QNetwork, init_params
from .params import RlaxRainbowParams


class DQNAgent:
    def __init__(self, observation_spec, action_spec, params: RlaxRainbowParams = RlaxRainbowParams(),
                 device=None, noisy=True, adam_eps=3.125e-5, process_group=None):
        import torch
        self.torch = torch
        if not callable(params.epsilon):          # rlax_rainbow.py:227-232
            eps = params.epsilon
            params = params._replace(epsilon=lambda ts: eps)
        if not callable(params.beta_is):
            beta = params.beta_is
            params = params._replace(beta_is=lambda ts: beta)
        self.params = params
        if device is None:
            device = torch.cuda.current_device() if torch.cuda.is_available() else 0
        self.device = torch.device("cuda", int(device))
        self.n_envs, self.obs_len = int(observation_spec.shape[0]), int(observation_spec.shape[1])
        self.num_actions = int(action_spec.num_values)
        self.obs_stride = ((self.obs_len + 31) // 32 + 3) // 4 * 4
        self.legal_words = (self.num_actions + 31) // 32
        self.network = QNetwork(self.obs_len, self.obs_stride, list(params.layers), self.num_actions,
                                params.n_atoms, float(params.atom_vmax), noisy=noisy, device=int(device))
        # trg_params = network.init(...); online_params = trg_params   (rlax_rainbow.py:250-252)
        self.network.set_params(init_params(self.obs_len, list(params.layers), self.num_actions,
                                            params.n_atoms, noisy=noisy, seed=params.seed))
        self.adam_eps = float(adam_eps)              # optix.adam(lr, eps=3.125e-5)  :257
        self.train_step = 0
        self._rng_counter = 0                        # stands in for hk.PRNGSequence(seed)  :234
        cls = PriorityBuffer if params.use_priority else ExperienceBuffer
        self.experience = cls(self.obs_len, self.num_actions, 1, params.experience_buffer_size,
                              device=int(device))
        self.last_obs = torch.zeros((self.n_envs, self.obs_stride), dtype=torch.int32, device=self.device)
        self.requires_vectorized_observation = lambda: True
        self.process_group = process_group
        self.last_loss = None
        self._learner = None

    # -- helpers ----------------------------------------------------------------------------
    def _pack_pair(self, observations):
        obs, legal = observations[1]
        exp = self.experience
        return (exp._packed(obs, self.obs_len, self.obs_stride),
                exp._packed(legal, self.num_actions, self.legal_words), _is_cuda(obs))

    def _dev(self, x, dtype):
        return self.experience._vec(x, dtype)

    def _world(self):
        dist = self.torch.distributed
        if dist.is_available() and dist.is_initialized():
            return dist.get_world_size(self.process_group), dist.get_rank(self.process_group)
        return 1, 0

    def _act(self, observations, greedy):
        obs, legal, on_device = self._pack_pair(observations)
        world, rank = self._world()
        self._rng_counter += 1
        act = self.network.act(obs, legal, greedy=greedy,
                               epsilon=0.0 if greedy else float(self.params.epsilon(self.train_step)),
                               seed=self.params.seed, counter=self._rng_counter,
                               row0=rank * self.n_envs)
        return act if on_device else act.cpu().numpy()

    # -- the reference's verbs --------------------------------------------------------------
    def exploit(self, observations):               # rlax_rainbow.py:277-282
        return self._act(observations, True)

    def explore(self, observations):               # rlax_rainbow.py:284-290
        return self._act(observations, False)

    def add_experience_first(self, observations, step_types):      # :292-295
        obs, _, _ = self._pack_pair(observations)
        self.experience.store_first(self.last_obs, obs, self._dev(step_types, self.torch.uint8))

    def add_experience(self, observations, actions, rewards, step_types):   # :297-308
        torch = self.torch
        obs, legal, _ = self._pack_pair(observations)
        return self.experience.add_masked(self.last_obs, obs, legal, self._dev(actions, torch.int32),
                                          self._dev(rewards, torch.float32),
                                          self._dev(step_types, torch.uint8))

    def add_experience_async(self, observations, actions, rewards, step_types):
        """add_experience without the count coming back to the host (no synchronisation); returns None."""
        torch = self.torch
        obs, legal, _ = self._pack_pair(observations)
        self.experience.add_masked_async(self.last_obs, obs, legal, self._dev(actions, torch.int32),
                                         self._dev(rewards, torch.float32), self._dev(step_types, torch.uint8))

    def add_experience_begin(self, observations, step_types):
        """add_experience in two halves (the session starts every seat's row count before it waits for the
        first): begin on the current stream, add_experience_finish with the same arguments on the same one."""
        obs, _, _ = self._pack_pair(observations)
        self.experience.add_masked_begin(self._dev(step_types, self.torch.uint8), obs.shape[0])

    def add_experience_finish(self, observations, actions, rewards, step_types):
        torch = self.torch
        obs, legal, _ = self._pack_pair(observations)
        return self.experience.add_masked_finish(self.last_obs, obs, legal, self._dev(actions, torch.int32),
                                                 self._dev(rewards, torch.float32),
                                                 self._dev(step_types, torch.uint8))

    def _make_learner(self):
        import ctypes as C
        from .. import _lib
        from .network import _DevArray
        L = _lib.lib()
        h = C.c_void_p()
        B = self.params.train_batch_size
        _lib.check(L.hb_learner_create(self.network._h, self.experience._h, B, self.device.index, C.byref(h)))
        self._learner = h
        torch = self.torch
        self._loss_view = torch.as_tensor(_DevArray(L.hb_learner_loss(h), B, self), device=self.device)
        self._mean_view = torch.as_tensor(_DevArray(L.hb_learner_mean_loss(h), 1, self), device=self.device)

    def __del__(self):
        h = getattr(self, "_learner", None)
        if h is not None and h.value:
            try:
                from .. import _lib
                _lib.lib().hb_learner_destroy(h)
            except Exception:
                pass
            self._learner = None

    _fused_dp = False

    def _setup_allreduce(self):
        """The one collective of the job: sum of the flat gradient bucket over the ranks.  Kernels over
        NVLink peer memory (dist.NvlinkAllReduce) when the peers can be mapped -- then fused into the
        update, one all-reduce per layer beside the backward pass (HB_ALLREDUCE=nvlink, the default;
        HB_ALLREDUCE=nvlink-unfused keeps one whole-bucket all-reduce between two graphs) -- NCCL
        otherwise or when HB_ALLREDUCE=nccl."""
        if getattr(self, "_comm", None) is None:
            self._comm = False
            self.allreduce_impl = "nccl"
            if os.environ.get("HB_ALLREDUCE", "nvlink") != "nccl":
                try:
                    from ..dist import NvlinkAllReduce
                    self._comm = NvlinkAllReduce(self.network.grad_bucket, self.process_group)
                    self.allreduce_impl = "nvlink"
                    if os.environ.get("HB_ALLREDUCE", "nvlink") == "nvlink":
                        from .. import _lib
                        world, _ = self._world()
                        _lib.check(_lib.lib().hb_learner_set_comm(self._learner, self._comm._h, world))
                        self._fused_dp = True
                        self.allreduce_impl = "nvlink, per layer inside the update"
                except RuntimeError as e:   # peers cannot be mapped: every rank takes this branch (dist.py)
                    import warnings
                    self._comm = False
                    self.allreduce_fallback_reason = repr(e)
                    warnings.warn("hanabi_agents_b200: NVLink all-reduce unavailable, using "
                                  "torch.distributed.all_reduce instead: %r" % (e,), RuntimeWarning)

    def _allreduce_grads(self):
        if getattr(self, "_comm", None) is None:
            self._setup_allreduce()
        if self._comm:
            self._comm()
        else:
            self.torch.distributed.all_reduce(self.network.grad_bucket, group=self.process_group)

    def update(self, uniforms=None):
        """One training step (rlax_rainbow.py:310-339): sample_batch -> update_q -> update_priorities
        -> target sync.  Two C calls (each a replayed CUDA graph) around the gradient all-reduce.
        `uniforms`: optional explicit draws for the stratified keys (default: numpy's global RNG,
        exactly as the reference draws them).  Returns the mean loss as a 1-element CUDA tensor."""
        from .. import _lib
        L = _lib.lib()
        if getattr(self, "_learner", None) is None:
            self._make_learner()
        B = self.params.train_batch_size
        exp = self.experience
        stream = _lib.current_stream_ptr()
        if self.params.use_priority:
            keys = np.ascontiguousarray(PriorityBuffer.stratified_keys(B, uniforms), dtype=np.float32)
            kp, ip = keys.ctypes.data, None
        else:                                        # uniform sample, prios = ones  (:315-318)
            # np.random.choice(np.arange(capacity)[:size], B) draws randint(0, size, B) from the global RNG:
            # the same stream without building a capacity-sized array every step
            idx = np.ascontiguousarray(np.random.randint(0, exp.size, size=B), dtype=np.int64)
            kp, ip = None, idx.ctypes.data
        world, _ = self._world()
        beta = float(self.params.beta_is(self.train_step))
        lr = self.params.learning_rate
        lr = float(lr(self.train_step)) if callable(lr) else float(lr)
        if world > 1 and getattr(self, "_comm", None) is None:
            self._setup_allreduce()
        if world == 1 or self._fused_dp:
            # one call, one graph: Adam runs layer by layer beside the backward pass; data parallel: each
            # layer's gradient is all-reduced over NVLink on its own channel in between (csrc/comm.cu)
            _lib.check(L.hb_learner_step(self._learner, kp, ip, beta, float(self.params.discount), 0,
                                         lr, 0.9, 0.999, self.adam_eps, stream))
        else:
            _lib.check(L.hb_learner_grads(self._learner, kp, ip, beta, float(self.params.discount), 0, stream))
            self._allreduce_grads()
            _lib.check(L.hb_learner_apply(self._learner, lr, 0.9, 0.999, self.adam_eps, 1.0 / world, stream))
        if self.train_step % self.params.target_update_period == 0:   # :336-337
            self.network.sync_target()
        self.train_step += 1
        self.last_loss = self._mean_view
        return self._mean_view

    def __repr__(self):
        return f"<rlax_dqn.DQNAgent(params={self.params})>"

    # -- weights (rlax_rainbow.py:344-365) ----------------------------------------------------
    def save_weights(self, path, fname_part):
        """Pickles of the haiku-style parameter trees, file names as the reference writes them."""
        for kind in ("online", "target"):
            tree = self.network.haiku_tree(target=kind == "target")
            with open(os.path.join(path, "rlax_rainbow_" + fname_part + "_" + kind + ".pkl"), "wb") as of:
                pickle.dump(tree, of)

    def restore_weights(self, online_weights_file, trg_weights_file):
        with open(trg_weights_file, "rb") as iwf:
            trg = pickle.load(iwf)
        with open(online_weights_file, "rb") as iwf:
            online = pickle.load(iwf)
        self.network.load_haiku_tree(trg)
        self.network.sync_target()
        self.network.load_haiku_tree(online)
