"""RulebasedAgent with the reference's verbs (hanabi_agents/rule_based/rule_based.py:5-48),
acting on every game of a HanabiParallelEnv with one kernel launch."""
import ctypes as C

import numpy as np

from .. import _lib
from .ruleset import Rule


class RulebasedAgent:
    def __init__(self, rules, seed=1234):
        for r in rules:
            if not isinstance(r, Rule):
                raise TypeError("rules must be Ruleset rules, got %r" % (r,))
        if len(rules) > 16:
            raise ValueError("at most 16 rules")
        self.rules = list(rules)
        self.seed = int(seed)
        self._step = 0
        self._ids = np.array([r.rule_id for r in rules], np.int32)
        self._farg = np.array([r.farg for r in rules], np.float64)
        self._iarg = np.array([r.iarg for r in rules], np.int32)
        self._hist = None
        self._actions = None

    @staticmethod
    def _env_of(observations):
        # observations is what HanabiParallelEnv.observations() returns, or the env itself
        env = observations[0] if isinstance(observations, tuple) else observations
        if not hasattr(env, "_h"):
            raise TypeError("RulebasedAgent acts on a HanabiParallelEnv (observations()[0])")
        return env

    def _buffers(self, env):
        torch = env.torch
        if self._hist is None or self._hist.device != env.device:
            self._hist = torch.zeros(len(self.rules) + 1, dtype=torch.int64, device=env.device)
        if self._actions is None or self._actions.numel() != env.n or self._actions.device != env.device:
            self._actions = torch.zeros(env.n, dtype=torch.int32, device=env.device)

    def explore(self, observations):
        """Move uid for every game (int32 CUDA tensor).  rule_based.py:27-31"""
        env = self._env_of(observations)
        self._buffers(env)
        _lib.check(_lib.lib().hb_rules_act(
            env._h, self._ids.ctypes.data, self._farg.ctypes.data, self._iarg.ctypes.data,
            len(self.rules), self.seed, self._step, self._actions.data_ptr(), None,
            self._hist.data_ptr(), _lib.current_stream_ptr()))
        self._step += 1
        return self._actions

    def explore_host_async(self, env, out):
        """explore() for a host-side loop: the moves of every game land in `out` (pinned int32[N] torch
        CPU tensor) chunk by chunk behind `env.step_host_async`; valid after `env.host_wait()`."""
        self._buffers(env)
        _lib.check(_lib.lib().hb_rules_act_host_async(
            env._h, self._ids.ctypes.data, self._farg.ctypes.data, self._iarg.ctypes.data,
            len(self.rules), self.seed, self._step, out.data_ptr(), self._hist.data_ptr()))
        self._step += 1
        return out

    def exploit(self, observations):  # rule_based.py:33-34
        return self.explore(observations)

    def rule_args(self):
        """(ids, fargs, iargs) host arrays, for the fused self-play entry point."""
        return self._ids, self._farg, self._iarg

    @property
    def histogram(self):  # rule_based.py:9-10,19-23: hits per rule, last = legal_random fallback
        if self._hist is None:
            return [0] * (len(self.rules) + 1)
        return [int(x) for x in self._hist.cpu().tolist()]

    @property
    def totalCalls(self):
        return sum(self.histogram)

    def requires_vectorized_observation(self):  # rule_based.py:37-38
        return False

    def add_experience_first(self, o, st):  # rule_based.py:41-48: no learning
        pass

    def add_experience(self, o, a, r, st):
        pass

    def update(self):
        pass
