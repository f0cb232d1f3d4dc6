"""ExperienceBuffer / PriorityBuffer with the reference's surface, resident on the B200.

Mirrors hanabi_agents/rlax_dqn/experience_buffer.py:5-97 and priority_buffer.py:13-52: same
constructor arguments, attributes (`capacity`, `size`, `oldest_entry`, `sum_tree`, `max_priority`,
`min_priority`, `alpha`) and methods.  Rows live in HBM bit-packed; numpy int8 arguments are copied
in and packed (compat path, what the reference's callers pass), bit-packed CUDA tensors are taken
as they are (fast path, what HanabiParallelEnv emits).
"""
import ctypes as C

import numpy as np

from .. import _lib
from ..sum_tree import SumTreef
from .transition import Transition


def _is_cuda(x):
    return hasattr(x, "is_cuda") and x.is_cuda


class ExperienceBuffer:
    _PRIORITIZED = False

    def __init__(self, observation_len, action_len, reward_len, capacity, device=None, alpha=0.6):
        import torch
        self.torch = torch
        if device is None:
            device = torch.cuda.current_device() if torch.cuda.is_available() else 0
        self.device_index = int(device)
        self.device = torch.device("cuda", self.device_index)
        self.observation_len, self.action_len = int(observation_len), int(action_len)
        self.reward_len = int(reward_len)
        if self.reward_len != 1:
            raise ValueError("reward_len must be 1 (as every caller of the reference passes)")
        self.obs_stride = ((self.observation_len + 31) // 32 + 3) // 4 * 4
        self.legal_words = (self.action_len + 31) // 32
        self._h = C.c_void_p()
        _lib.check(_lib.lib().hb_rb_create(self.observation_len, self.obs_stride, self.action_len,
                                           int(capacity), int(self._PRIORITIZED), float(alpha),
                                           self.device_index, C.byref(self._h)))
        self.capacity = int(capacity)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                _lib.lib().hb_rb_destroy(h)
            except Exception:
                pass
            self._h = None

    def _s(self):
        return _lib.current_stream_ptr()

    # -- reference attributes -----------------------------------------------------------------
    @property
    def size(self):
        return int(_lib.lib().hb_rb_info(self._h, 1))

    @property
    def oldest_entry(self):
        return int(_lib.lib().hb_rb_info(self._h, 2))

    def get_update_indices(self, batch_size):   # experience_buffer.py:21-24
        o = self.oldest_entry
        if o + batch_size <= self.capacity:
            return list(range(o, o + batch_size))
        return list(range(o, self.capacity)) + list(range(0, batch_size - self.capacity + o))

    # -- helpers --------------------------------------------------------------------------------
    def _packed(self, x, bits, stride):
        """int8/bool [n, bits] (numpy or CUDA) -> packed int32 [n, stride] CUDA; packed passes through."""
        torch = self.torch
        if _is_cuda(x) and x.dtype == torch.int32 and x.dim() == 2 and x.shape[1] == stride:
            return x.contiguous()
        if not _is_cuda(x):
            x = torch.as_tensor(np.ascontiguousarray(np.asarray(x).reshape(len(x), -1), dtype=np.int8),
                                device=self.device)
        x = x.to(torch.int8).contiguous()
        if x.dim() != 2 or x.shape[1] != bits:
            raise ValueError("expected [n, %d] rows, got %s" % (bits, tuple(x.shape)))
        out = torch.empty((x.shape[0], stride), dtype=torch.int32, device=self.device)
        _lib.check(_lib.lib().hb_pack_bits(x.data_ptr(), bits, out.data_ptr(), stride, x.shape[0], self._s()))
        return out

    def _vec(self, x, dtype):
        torch = self.torch
        if not _is_cuda(x):
            x = torch.as_tensor(np.ascontiguousarray(np.asarray(x).reshape(-1)), device=self.device)
        return x.reshape(-1).to(dtype).contiguous()

    def unpack_obs(self, packed):
        torch = self.torch
        out = torch.empty((packed.shape[0], self.observation_len), dtype=torch.int8, device=self.device)
        _lib.check(_lib.lib().hb_unpack_bits(packed.data_ptr(), self.obs_stride, self.observation_len,
                                             out.data_ptr(), packed.shape[0], self._s()))
        return out

    def unpack_legal(self, packed):
        torch = self.torch
        out = torch.empty((packed.shape[0], self.action_len), dtype=torch.int8, device=self.device)
        _lib.check(_lib.lib().hb_unpack_bits(packed.data_ptr(), self.legal_words, self.action_len,
                                             out.data_ptr(), packed.shape[0], self._s()))
        return out

    # -- reference methods --------------------------------------------------------------------
    def add_transitions(self, observation_tm1, action_tm1, reward_t, observation_t, legal_moves_t,
                        terminal_t):
        """experience_buffer.py:26-81 (PriorityBuffer: priority_buffer.py:22-34)."""
        torch = self.torch
        o1 = self._packed(observation_tm1, self.observation_len, self.obs_stride)
        o2 = self._packed(observation_t, self.observation_len, self.obs_stride)
        lm = self._packed(legal_moves_t, self.action_len, self.legal_words)
        a = self._vec(action_tm1, torch.int32)
        r = self._vec(reward_t, torch.float32)
        t = self._vec(terminal_t, torch.uint8)
        n = o1.shape[0]
        if not (o2.shape[0] == lm.shape[0] == a.numel() == r.numel() == t.numel() == n):
            raise ValueError("transition arrays differ in length")
        _lib.check(_lib.lib().hb_rb_add(self._h, o1.data_ptr(), a.data_ptr(), r.data_ptr(), o2.data_ptr(),
                                        lm.data_ptr(), t.data_ptr(), n, self._s()))

    def add_masked(self, last_obs, obs, legal, actions, rewards, step_types):
        """The body of DQNAgent.add_experience (rlax_rainbow.py:297-308) on packed device rows.
        Returns the number of transitions appended."""
        n_added = C.c_int64()
        _lib.check(_lib.lib().hb_rb_add_masked(
            self._h, last_obs.data_ptr(), obs.data_ptr(), legal.data_ptr(), actions.data_ptr(),
            rewards.data_ptr(), step_types.data_ptr(), obs.shape[0], C.byref(n_added), self._s()))
        return int(n_added.value)

    def add_masked_begin(self, step_types, n):
        """First half of add_masked (hb_rb_add_masked_begin): the row count is on its way to the host."""
        _lib.check(_lib.lib().hb_rb_add_masked_begin(self._h, step_types.data_ptr(), int(n), self._s()))

    def add_masked_finish(self, last_obs, obs, legal, actions, rewards, step_types):
        """Second half, on the stream of the first: waits for the count and appends."""
        n_added = C.c_int64()
        _lib.check(_lib.lib().hb_rb_add_masked_finish(
            self._h, last_obs.data_ptr(), obs.data_ptr(), legal.data_ptr(), actions.data_ptr(),
            rewards.data_ptr(), step_types.data_ptr(), obs.shape[0], C.byref(n_added), self._s()))
        return int(n_added.value)

    def add_masked_async(self, last_obs, obs, legal, actions, rewards, step_types):
        """add_masked without a host round trip (hb_rb_add_masked_async): nothing is returned; `size` and
        `oldest_entry` catch up when they are next read."""
        _lib.check(_lib.lib().hb_rb_add_masked_async(
            self._h, last_obs.data_ptr(), obs.data_ptr(), legal.data_ptr(), actions.data_ptr(),
            rewards.data_ptr(), step_types.data_ptr(), obs.shape[0], self._s()))

    def store_first(self, last_obs, obs, step_types):
        _lib.check(_lib.lib().hb_rb_store_first(self._h, last_obs.data_ptr(), obs.data_ptr(),
                                                step_types.data_ptr(), obs.shape[0], self._s()))

    def gather_packed(self, indices):
        """Device gather: (obs_tm1 packed, act i32, rew f32, obs_t packed, lms packed, term u8)."""
        torch = self.torch
        if not _is_cuda(indices):
            indices = torch.as_tensor(np.ascontiguousarray(np.asarray(indices, dtype=np.int64).reshape(-1)),
                                      device=self.device)
        idx = indices.to(torch.int64).contiguous()
        n = idx.numel()
        e = lambda shape, dt: torch.empty(shape, dtype=dt, device=self.device)  # noqa: E731
        o1, o2 = e((n, self.obs_stride), torch.int32), e((n, self.obs_stride), torch.int32)
        lm, a = e((n, self.legal_words), torch.int32), e((n,), torch.int32)
        r, t = e((n,), torch.float32), e((n,), torch.uint8)
        _lib.check(_lib.lib().hb_rb_gather(self._h, idx.data_ptr(), n, o1.data_ptr(), a.data_ptr(),
                                           r.data_ptr(), o2.data_ptr(), lm.data_ptr(), t.data_ptr(), self._s()))
        return o1, a, r, o2, lm, t

    def __getitem__(self, indices):
        """experience_buffer.py:83-87: Transition of numpy arrays in the reference's dtypes."""
        o1, a, r, o2, lm, t = self.gather_packed(indices)
        _lib.check(_lib.lib().hb_rb_check(self._h, self._s()))
        return Transition(
            self.unpack_obs(o1).cpu().numpy(), a.cpu().numpy().astype(np.int8).reshape(-1, 1),
            r.cpu().numpy().astype(np.float64).reshape(-1, 1), self.unpack_obs(o2).cpu().numpy(),
            self.unpack_legal(lm).cpu().numpy(), t.cpu().numpy().astype(bool).reshape(-1, 1))

    def sample(self, batch_size):
        """Uniform sample with numpy's global RNG (experience_buffer.py:89-97)."""
        indices = np.random.choice(np.arange(0, self.capacity, dtype=np.int64)[:self.size], size=batch_size)
        return self[indices]


class PriorityBuffer(ExperienceBuffer):
    _PRIORITIZED = True

    def __init__(self, observation_len, action_len, reward_len, capacity, alpha=0.6, device=None):
        super().__init__(observation_len, action_len, reward_len, capacity, device=device, alpha=alpha)
        self.alpha = alpha
        self.sum_tree = SumTreef._from_handle(_lib.lib().hb_rb_tree(self._h), self.device_index, self)

    def _range(self):
        mx, mn = C.c_float(), C.c_float()
        _lib.check(_lib.lib().hb_rb_priority_range(self._h, C.byref(mx), C.byref(mn), self._s()))
        return float(mx.value), float(mn.value)

    @property
    def max_priority(self):
        return self._range()[0]

    @property
    def min_priority(self):
        return self._range()[1]

    @staticmethod
    def stratified_keys(batch_size, uniforms=None):
        """priority_buffer.py:37-40: keys = linspace(1/B, 1, B) - U[0, 1/B) (float64; numpy's global
        RNG when `uniforms` is None, exactly as the reference draws them)."""
        keys = np.linspace(1. / batch_size, 1, batch_size)
        if uniforms is None:
            uniforms = np.random.uniform(size=(batch_size,), high=1. / batch_size)
        return keys - np.asarray(uniforms, dtype=np.float64)

    def sample_device(self, batch_size, uniforms=None):
        """Device-side sample_batch: (indices i64, prios f64, prios f32, packed transition tuple)."""
        torch = self.torch
        keys = self.stratified_keys(batch_size, uniforms).astype(np.float32)   # pybind narrows to fp32
        k = torch.as_tensor(keys, device=self.device)
        idx = torch.empty(batch_size, dtype=torch.int64, device=self.device)
        p64 = torch.empty(batch_size, dtype=torch.float64, device=self.device)
        p32 = torch.empty(batch_size, dtype=torch.float32, device=self.device)
        _lib.check(_lib.lib().hb_rb_sample(self._h, k.data_ptr(), batch_size, idx.data_ptr(),
                                           p64.data_ptr(), p32.data_ptr(), self._s()))
        return idx, p64, p32, self.gather_packed(idx)

    def sample_batch(self, batch_size, uniforms=None):
        """priority_buffer.py:36-46: (indices list, prios float64 ndarray, Transition)."""
        idx, p64, _, _ = self.sample_device(batch_size, uniforms)
        indices = idx.cpu().numpy()
        return indices.tolist(), p64.cpu().numpy(), self[indices]

    def update_priorities(self, indices, priorities):
        """priority_buffer.py:48-52."""
        torch = self.torch
        if not _is_cuda(indices):
            indices = torch.as_tensor(np.ascontiguousarray(np.asarray(indices, dtype=np.int64).reshape(-1)),
                                      device=self.device)
        if not _is_cuda(priorities):
            priorities = torch.as_tensor(np.ascontiguousarray(np.asarray(priorities, dtype=np.float32).reshape(-1)),
                                         device=self.device)
        idx = indices.to(torch.int64).contiguous()
        td = priorities.to(torch.float32).contiguous()
        if idx.numel() != td.numel():
            raise ValueError("indices and priorities differ in length")
        _lib.check(_lib.lib().hb_rb_update_priorities(self._h, idx.data_ptr(), td.data_ptr(), idx.numel(),
                                                      self._s()))
