"""CPU restatement (numpy) of the reference's replay buffers -- TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.

Follows /root/reference/hanabi_agents/rlax_dqn/experience_buffer.py:5-97 and
priority_buffer.py:13-52 (the live code; the rest of that file is commented out / unused).
PINNED: tests/test_replay_oracle.py replays the cases of the reference's own
tests/rlax_dqn/test_experience_buffer.py (fill, exact fill, overflow-by-one wrap, getter) and
compares with fixtures produced by the reference's experience_buffer.py itself, loaded standalone
(tests/golden/make_replay_golden.py; `np.int` is shimmed because the file predates numpy 1.24).
priority_buffer.py cannot be imported (it imports jax at line 3 without using it), so its 40 live
lines are restated here on top of the pinned sum-tree oracle.

One deliberate precision statement: update_priorities computes `(td + 1e-10) ** alpha` on a float32
array; numpy's float32 power is platform code (SVML / libm) and not correctly rounded everywhere.
The restatement evaluates the power in float64 and rounds once, i.e. the correctly rounded float32
power; tests/test_replay_oracle.py checks numpy's own float32 result stays within 1 ulp of it.
"""
import numpy as np

from . import OracleSumTree


class OracleExperienceBuffer:
    def __init__(self, observation_len, action_len, reward_len, capacity):
        self._obs_tm1_buf = np.zeros((capacity, observation_len), dtype=np.byte)
        self._act_tm1_buf = np.zeros((capacity, 1), dtype=np.byte)
        self._obs_t_buf = np.zeros((capacity, observation_len), dtype=np.byte)
        self._lms_t_buf = np.zeros((capacity, action_len), dtype=np.byte)
        self._rew_t_buf = np.zeros((capacity, reward_len), dtype=np.float64)
        self._terminal_t_buf = np.zeros((capacity, 1), dtype=bool)
        self.oldest_entry = 0
        self.capacity = capacity
        self.size = 0

    def get_update_indices(self, batch_size):                      # :21-24
        if self.oldest_entry + batch_size <= self.capacity:
            return list(range(self.oldest_entry, self.oldest_entry + batch_size))
        return list(range(self.oldest_entry, self.capacity)) + \
            list(range(0, batch_size - self.capacity + self.oldest_entry))

    def add_transitions(self, observation_tm1, action_tm1, reward_t, observation_t, legal_moves_t,
                        terminal_t):                               # :26-81
        batch_size = len(observation_tm1)
        bufs = (self._obs_tm1_buf, self._act_tm1_buf, self._obs_t_buf, self._lms_t_buf,
                self._rew_t_buf, self._terminal_t_buf)
        vals = (observation_tm1, action_tm1, observation_t, legal_moves_t, reward_t, terminal_t)
        o = self.oldest_entry
        if o + batch_size <= self.capacity:
            for b, v in zip(bufs, vals):
                b[o:o + batch_size, :] = v
            if o + batch_size == self.capacity:
                self.size = self.capacity
            self.oldest_entry = (o + batch_size) % self.capacity
            self.size = max(self.size, self.oldest_entry)
        else:
            tail = o + batch_size - self.capacity
            for b, v in zip(bufs, vals):
                b[o:, :] = v[:batch_size - tail]
                b[:tail, :] = v[-tail:]
            self.oldest_entry = tail
            self.size = self.capacity

    def __getitem__(self, indices):                                # :83-87
        return (self._obs_tm1_buf[indices], self._act_tm1_buf[indices], self._rew_t_buf[indices],
                self._obs_t_buf[indices], self._lms_t_buf[indices], self._terminal_t_buf[indices])


def priority_pow(td_abs, alpha):
    """(td + 1e-10) ** alpha on float32, correctly rounded (see module docstring)."""
    x = np.asarray(td_abs, dtype=np.float32) + np.float32(1e-10)
    return np.power(x.astype(np.float64), np.float64(np.float32(alpha))).astype(np.float32)


class OraclePriorityBuffer(OracleExperienceBuffer):
    def __init__(self, observation_len, action_len, reward_len, capacity, alpha=0.6):   # :14-20
        super().__init__(observation_len, action_len, reward_len, capacity)
        self.sum_tree = OracleSumTree(capacity)
        self.max_priority = alpha
        self.min_priority = alpha
        self.alpha = alpha

    def add_transitions(self, observation_tm1, action_tm1, reward_t, observation_t, legal_moves_t,
                        terminal_t):                               # :22-34
        batch_size = len(observation_tm1)
        self.sum_tree.update_values(self.get_update_indices(batch_size),
                                    [self.max_priority for _ in range(batch_size)])
        super().add_transitions(observation_tm1, action_tm1, reward_t, observation_t, legal_moves_t,
                                terminal_t)

    def sample_batch(self, batch_size, uniforms):                  # :36-46, uniforms explicit
        keys = np.linspace(1. / batch_size, 1, batch_size)
        keys -= np.asarray(uniforms, dtype=np.float64)
        indices = self.sum_tree.get_indices(keys.astype(np.float32))   # pybind narrows to float
        prios = (np.array(self.sum_tree.get_values(indices), dtype=np.float64) + 1e-10) / \
            self.sum_tree.get_total_val()
        return indices, prios, self[indices]

    def update_priorities(self, indices, priorities):              # :48-52
        priorities = priority_pow(priorities, self.alpha)
        self.max_priority = max(self.max_priority, np.max(priorities))
        self.min_priority = min(self.min_priority, np.min(priorities))
        self.sum_tree.update_values(indices, priorities)
