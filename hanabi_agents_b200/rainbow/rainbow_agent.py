"""RainbowAgent with the reference's Dopamine-style verbs (hanabi_agents/rainbow/dqn_agent.py:284-332,
rainbow_agent.py:75-246) on the B200 kernels -- SURVEY section 8f row 3.

Same constructor arguments and verbs as the reference class:

    action = agent.begin_episode(current_player, legal_actions, observation)
    action = agent.step(reward, current_player, legal_actions, observation)
    agent.end_episode(final_rewards)

with `legal_actions` a float vector of 0 (legal) / -inf (illegal) (run_experiment.py:229-248) and
`observation` the 0/1 canonical vector.  What differs from the rlax agent and is reproduced here:
the next action of the target comes from the TARGET network over the LEGAL next moves
(rainbow_agent.py:199-204), gamma^n is masked by the terminal flag (:191-195), priorities are
sqrt(loss + 1e-10) (:233), new rows enter at DEFAULT_PRIORITY = 100
(prioritized_replay_memory.py:38), a batch is `batch_size` independent draws from the sum tree (:105-131),
importance weights are 1/sqrt(priority + 1e-10) / max (:236-239), the optimiser is
tf.train.AdamOptimizer, the network a plain ReLU MLP (rainbow_template :45-72).

Documented divergences: `update_horizon` must be 1 (the configured value, configs/hanabi_rainbow.gin);
each replay row is a complete transition, so nothing is ever "invalid" to sample (replay_memory.py
rejects the rows next to its write cursor); the staged batch of the TF graph (a batch sampled one
training step earlier, dqn_agent.py:433-440) is sampled fresh; the sum tree is the fp32 device tree,
the reference's third_party/dopamine/sum_tree.py float64 numpy.
"""
import math
import random

import numpy as np

from .. import _lib
from ..rlax_dqn.buffers import PriorityBuffer
from ..rlax_dqn.network import QNetwork, truncated_normal  # noqa: F401

DEFAULT_PRIORITY = 100.0        # prioritized_replay_memory.py:38


def linearly_decaying_epsilon(decay_period, step, warmup_steps, epsilon):   # dqn_agent.py:44-60
    steps_left = decay_period + warmup_steps - step
    bonus = (1.0 - epsilon) * steps_left / decay_period
    bonus = np.clip(bonus, 0.0, 1.0 - epsilon)
    return epsilon + bonus


def init_params_tf(obs_len, hidden, num_actions, n_atoms, seed=0):
    """slim.fully_connected defaults (Glorot-uniform weights, zero biases) for the hidden layers,
    variance_scaling(factor=1/sqrt(3), FAN_IN, uniform) for the head (rainbow_agent.py:59-70).  The
    stream is numpy's default_rng(seed); TF's own stream cannot be reproduced without TF."""
    rng = np.random.default_rng(seed)
    sizes = [obs_len] + list(hidden) + [num_actions * n_atoms]
    layers = []
    for i in range(len(sizes) - 1):
        fi, fo = sizes[i], sizes[i + 1]
        if i < len(sizes) - 2:
            lim = math.sqrt(6.0 / (fi + fo))
        else:
            lim = math.sqrt(3.0 * (1.0 / math.sqrt(3.0)) / fi)
        layers.append({"w": rng.uniform(-lim, lim, (fi, fo)).astype(np.float32),
                       "b": np.zeros(fo, np.float32)})
    return layers


class RainbowAgent:
    def __init__(self, num_actions=None, observation_size=None, num_players=None, num_atoms=51, vmax=25.,
                 gamma=0.99, update_horizon=1, min_replay_history=500, update_period=4,
                 target_update_period=500, epsilon_train=0.0, epsilon_eval=0.0, epsilon_decay_period=1000,
                 learning_rate=0.000025, optimizer_epsilon=0.00003125, epsilon_fn=linearly_decaying_epsilon,
                 layer_size=512, num_layers=1, replay_capacity=1000000, batch_size=32, seed=0, device=None):
        import torch
        self.torch = torch
        if update_horizon != 1:
            raise NotImplementedError("update_horizon = %d: only the configured 1-step update is built"
                                      % update_horizon)
        self.num_actions, self.observation_size = int(num_actions), int(observation_size)
        self.num_players = int(num_players)
        self.num_atoms, self.vmax = int(num_atoms), float(vmax)
        self.gamma, self.update_horizon = float(gamma), int(update_horizon)
        self.cumulative_gamma = math.pow(gamma, update_horizon)         # dqn_agent.py:162
        self.min_replay_history = min_replay_history
        self.update_period, self.target_update_period = update_period, target_update_period
        self.epsilon_fn, self.epsilon_train, self.epsilon_eval = epsilon_fn, epsilon_train, epsilon_eval
        self.epsilon_decay_period = epsilon_decay_period
        self.learning_rate, self.optimizer_epsilon = float(learning_rate), float(optimizer_epsilon)
        self.batch_size = int(batch_size)
        self.eval_mode = False
        self.training_steps = 0
        self.add_count = 0
        if device is None:
            device = torch.cuda.current_device() if torch.cuda.is_available() else 0
        self.device = torch.device("cuda", int(device))
        self.obs_stride = ((self.observation_size + 31) // 32 + 3) // 4 * 4
        self.legal_words = (self.num_actions + 31) // 32
        hidden = [int(layer_size)] * int(num_layers)
        self.network = QNetwork(self.observation_size, self.obs_stride, hidden, self.num_actions,
                                self.num_atoms, self.vmax, noisy=False, device=int(device))
        self.network.set_params(init_params_tf(self.observation_size, hidden, self.num_actions,
                                               self.num_atoms, seed))
        self.network.sync_target()
        # the replay row is a whole transition; priorities: sqrt(loss + 1e-10) = (|x| + 1e-10)^0.5
        self._replay = PriorityBuffer(self.observation_size, self.num_actions, 1, int(replay_capacity),
                                      alpha=0.5, device=int(device))
        _lib.check(_lib.lib().hb_rb_set_insert_priority(self._replay._h, DEFAULT_PRIORITY))
        self.transitions = [[] for _ in range(self.num_players)]      # dqn_agent.py:170
        self.action = None
        self.last_loss = None
        B = self.batch_size
        self._loss = torch.zeros(B, dtype=torch.float32, device=self.device)
        self._mean = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._zero_u = torch.zeros(1, dtype=torch.float32, device=self.device)

    # -- helpers ----------------------------------------------------------------------------
    def _pack(self, observation, legal_actions):
        torch = self.torch
        obs = np.ascontiguousarray(np.asarray(observation).reshape(1, -1) != 0, dtype=np.int8)
        legal = np.ascontiguousarray((np.asarray(legal_actions, dtype=np.float32) == 0.0).reshape(1, -1),
                                     dtype=np.int8)                    # 0 = legal, -inf = illegal
        return (self._replay._packed(obs, self.observation_size, self.obs_stride),
                self._replay._packed(legal, self.num_actions, self.legal_words))

    def _select_action(self, observation, legal_actions, packed=None):   # dqn_agent.py:385-418
        if self.eval_mode:
            epsilon = self.epsilon_eval
        else:
            epsilon = self.epsilon_fn(self.epsilon_decay_period, self.training_steps,
                                      self.min_replay_history, self.epsilon_train)
        if random.random() <= epsilon:
            legal_action_indices = np.where(np.asarray(legal_actions) == 0.0)
            return int(np.random.choice(legal_action_indices[0]))
        obs, legal = packed if packed is not None else self._pack(observation, legal_actions)
        # tf.argmax(q + legal_actions): the first maximum = uniform draw 0 of the greedy sampler
        a = self.network.act(obs, legal, greedy=True, uniform=self._zero_u)
        return int(a.item())

    def _record_transition(self, current_player, reward, packed, action, begin=False):   # :333-356
        self.transitions[current_player].append((reward, packed[0], packed[1], action, begin))

    # -- the reference's verbs --------------------------------------------------------------
    def begin_episode(self, current_player, legal_actions, observation):   # :284-299
        self._train_step()
        packed = self._pack(observation, legal_actions)
        self.action = self._select_action(observation, legal_actions, packed)
        self._record_transition(current_player, 0, packed, self.action, begin=True)
        return self.action

    def step(self, reward, current_player, legal_actions, observation):    # :301-320
        self._train_step()
        packed = self._pack(observation, legal_actions)
        self.action = self._select_action(observation, legal_actions, packed)
        self._record_transition(current_player, reward, packed, self.action)
        return self.action

    def end_episode(self, final_rewards):                                  # :322-331
        self._post_transitions(terminal_rewards=final_rewards)

    def _post_transitions(self, terminal_rewards):                         # :358-383
        torch = self.torch
        if self.eval_mode:                       # _store_transition is a no-op in eval mode (:461)
            self.transitions = [[] for _ in range(self.num_players)]
            return
        for player in range(self.num_players):
            tr = self.transitions[player]
            n = len(tr)
            if n == 0:
                continue
            obs = torch.cat([t[1] for t in tr], 0)
            legal = torch.cat([t[2] for t in tr], 0)
            # row j: (o_j, a_j, r_{j+1} | terminal reward, o_{j+1}, l_{j+1}, terminal)
            rewards = [tr[j + 1][0] for j in range(n - 1)] + [terminal_rewards[player]]
            nxt_obs = torch.cat([obs[1:], torch.zeros_like(obs[:1])], 0)
            nxt_legal = torch.cat([legal[1:], torch.zeros_like(legal[:1])], 0)
            term = np.zeros(n, np.uint8)
            term[-1] = 1
            self._replay.add_transitions(obs, np.array([t[3] for t in tr], np.int32),
                                         np.array(rewards, np.float32), nxt_obs, nxt_legal, term)
            self.add_count += n
            self.transitions[player] = []

    def _train_step(self):                                                 # :420-446
        if self.eval_mode:
            return
        if self.add_count > self.min_replay_history and self.training_steps % self.update_period == 0:
            self._train()
        if self.training_steps % self.target_update_period == 0:
            self.network.sync_target()
        self.training_steps += 1

    def _train(self, uniforms=None):
        """One run of _train_op (rainbow_agent.py:213-246).  `uniforms`: the batch_size draws of
        sum_tree.sample() (default: python's random.random, as third_party/dopamine/sum_tree.py)."""
        torch = self.torch
        L = _lib.lib()
        B = self.batch_size
        rep = self._replay
        if uniforms is None:
            uniforms = [random.random() for _ in range(B)]
        keys = torch.as_tensor(np.asarray(uniforms, dtype=np.float32), device=self.device)
        idx = torch.empty(B, dtype=torch.int64, device=self.device)
        p64 = torch.empty(B, dtype=torch.float64, device=self.device)
        p32 = torch.empty(B, dtype=torch.float32, device=self.device)
        s = _lib.current_stream_ptr()
        _lib.check(L.hb_rb_sample(rep._h, keys.data_ptr(), B, idx.data_ptr(), p64.data_ptr(), p32.data_ptr(), s))
        o1, a, r, o2, lm, t = rep.gather_packed(idx)
        prio = rep.sum_tree.get_values(idx)                                # tf_get_priority  :235
        w = 1.0 / torch.sqrt(prio + 1e-10)
        w = (w / w.max()).contiguous()                                     # :236-239
        _lib.check(L.hb_net_grads_rainbow(self.network._h, o1.data_ptr(), a.data_ptr(), r.data_ptr(),
                                          o2.data_ptr(), lm.data_ptr(), t.data_ptr(), w.data_ptr(), B,
                                          float(self.cumulative_gamma), self._loss.data_ptr(),
                                          self._mean.data_ptr(), s))
        rep.update_priorities(idx, self._loss)                             # sqrt(loss + 1e-10)  :232-233
        _lib.check(L.hb_net_adam_tf(self.network._h, self.learning_rate, 0.9, 0.999, self.optimizer_epsilon,
                                    1.0, s))
        self.last_loss = self._mean
        self._last_batch = (idx, w)
        return self._mean
