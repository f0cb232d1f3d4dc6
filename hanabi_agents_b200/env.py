"""Batched Hanabi environment on the B200 (host-side mirror of the env API the reference's
agents assume: per-turn batches of `observations[1] == (obs[N,L], legal[N,A])`, `rewards[N]`,
`step_types[N]` in dm_env coding -- rlax_dqn/rlax_rainbow.py:277-308).

All tensors returned are torch CUDA tensors owned by the environment and overwritten by the
next call.  Observations and legal moves are bit-packed (`obs_packed[N, obs_stride]` uint32,
`legal_packed[N, legal_words]` uint32); `unpack_obs` / `unpack_legal` give the reference's
int8 0/1 layout when an API needs it.
"""
import ctypes as C

import numpy as np

from . import _lib
from .game import game_params, make_decks

FIRST, MID, LAST = 0, 1, 2


class ObservationSpec:
    """dm_env-style spec as DQNAgent uses it (rlax_rainbow.py:251,264,274)."""

    def __init__(self, shape, dtype=np.int8):
        self.shape, self.dtype = tuple(shape), dtype

    def generate_value(self):
        return np.zeros(self.shape, self.dtype)


class ActionSpec:
    """dm_env DiscreteArray-style spec (rlax_rainbow.py:249: action_spec.num_values)."""

    def __init__(self, num_values):
        self.num_values = int(num_values)
        self.shape, self.dtype = (), np.int32


class HanabiParallelEnv:
    def __init__(self, n_games, game="Hanabi-Full", players=2, device=None, decks=None,
                 n_decks=None, seed=1234, game_id_base=0, n_total=None, **overrides):
        import torch
        self.torch = torch
        params = game_params(game, players)
        params.update(overrides)
        self.params = params
        self.n = int(n_games)
        if device is None:
            device = torch.cuda.current_device() if torch.cuda.is_available() else 0
        self.device_index = int(device)
        self.device = torch.device("cuda", self.device_index)
        self._h = C.c_void_p()
        L = _lib.lib()
        _lib.check(L.hb_env_create(params["players"], params["colors"], params["ranks"],
                                   params["hand_size"], params["max_info"], params["max_life"],
                                   self.n, self.device_index, C.byref(self._h)))
        info = lambda k: int(L.hb_env_info(self._h, k))  # noqa: E731
        self.obs_len, self.obs_stride, self.num_moves = info(0), info(1), info(2)
        self.legal_words, self.deck_size, self.state_words = info(3), info(4), info(5)
        self.dump_len, self.players = info(7), info(8)
        if n_total is not None or game_id_base:
            _lib.check(L.hb_env_set_shard(self._h, int(game_id_base), int(n_total or self.n)))
        self.game_id_base, self.n_total = int(game_id_base), int(n_total or self.n)
        # the step outputs are views of ONE device allocation (`outputs_blob`), so that a host-side
        # consumer fetches a whole step with a single copy; `split_outputs` carves the same views
        # out of a host copy of the blob
        def up(x):
            return (x + 255) // 256 * 256
        sizes = [self.n * self.obs_stride * 4, self.n * self.legal_words * 4, self.n * 4, self.n]
        self._blob_off = [0]
        for b in sizes:
            self._blob_off.append(self._blob_off[-1] + up(b))
        self.outputs_blob = torch.zeros(self._blob_off[-1], dtype=torch.uint8, device=self.device)
        self.obs_packed, self.legal_packed, self.reward, self.step_type = self.split_outputs(self.outputs_blob)
        if decks is None:
            decks = make_decks(params["colors"], params["ranks"], n_decks or 4 * self.n_total, seed)
        self.set_decks(decks)

    def split_outputs(self, blob):
        """(obs_packed, legal_packed, reward, step_type) views of a uint8 blob laid out like
        `outputs_blob` (device or pinned host)."""
        torch = self.torch
        o = self._blob_off
        n = self.n
        obs = blob[o[0]:o[0] + n * self.obs_stride * 4].view(torch.int32).view(n, self.obs_stride)
        legal = blob[o[1]:o[1] + n * self.legal_words * 4].view(torch.int32).view(n, self.legal_words)
        reward = blob[o[2]:o[2] + n * 4].view(torch.float32)
        step_type = blob[o[3]:o[3] + n]
        return obs, legal, reward, step_type

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                _lib.lib().hb_env_destroy(h)
            except Exception:
                pass
            self._h = None

    # -- specs the agents read ----------------------------------------------------------
    def observation_spec(self):
        return ObservationSpec((self.n, self.obs_len))

    def action_spec(self):
        return ActionSpec(self.num_moves)

    # -- control ------------------------------------------------------------------------
    def _s(self):
        return _lib.current_stream_ptr()

    def set_decks(self, decks):
        """decks: [n_decks, deck_size] card indices (numpy -> validated on the host; torch CUDA
        tensor -> taken as is)."""
        L = _lib.lib()
        if hasattr(decks, "is_cuda") and decks.is_cuda:
            d = decks.to(self.torch.uint8).contiguous()
            assert d.dim() == 2 and d.shape[1] == self.deck_size
            _lib.check(L.hb_env_set_decks(self._h, d.data_ptr(), d.shape[0], self._s()))
            self.torch.cuda.current_stream().synchronize()
        else:
            d = np.ascontiguousarray(decks, dtype=np.uint8)
            if d.ndim != 2 or d.shape[1] != self.deck_size:
                raise ValueError("decks must be [n_decks, %d]" % self.deck_size)
            _lib.check(L.hb_env_set_decks_host(self._h, d.ctypes.data, d.shape[0]))
        self.n_decks = int(d.shape[0])

    def reset(self):
        _lib.check(_lib.lib().hb_env_reset(self._h, self.obs_packed.data_ptr(),
                                           self.legal_packed.data_ptr(),
                                           self.step_type.data_ptr(), self._s()))
        self.reward.zero_()
        return self.obs_packed, self.legal_packed, self.step_type

    def step(self, actions):
        """actions: int32[N] move uids (torch CUDA tensor, or anything numpy can read)."""
        torch = self.torch
        if not (hasattr(actions, "is_cuda") and actions.is_cuda):
            actions = torch.as_tensor(np.ascontiguousarray(actions, dtype=np.int32), device=self.device)
        a = actions.to(torch.int32).contiguous()
        if a.numel() != self.n:
            raise ValueError("expected %d actions, got %d" % (self.n, a.numel()))
        _lib.check(_lib.lib().hb_env_step(self._h, a.data_ptr(), self.obs_packed.data_ptr(),
                                          self.legal_packed.data_ptr(), self.reward.data_ptr(),
                                          self.step_type.data_ptr(), self._s()))
        return self.obs_packed, self.legal_packed, self.reward, self.step_type

    def step_host(self, actions, obs=None, legal=None, reward=None, step_type=None):
        """numpy in / numpy out through the host-buffer C entry point (copies inside)."""
        a = np.ascontiguousarray(actions, dtype=np.int32)
        obs = np.empty((self.n, self.obs_stride), np.uint32) if obs is None else obs
        legal = np.empty((self.n, self.legal_words), np.uint32) if legal is None else legal
        reward = np.empty(self.n, np.float32) if reward is None else reward
        step_type = np.empty(self.n, np.uint8) if step_type is None else step_type
        _lib.check(_lib.lib().hb_env_step_host(self._h, a.ctypes.data, obs.ctypes.data,
                                               legal.ctypes.data, reward.ctypes.data,
                                               step_type.ctypes.data))
        return obs, legal, reward, step_type

    # ---- chunked host pipeline (pinned host buffers; see include/hanabi_b200.h) ----------------
    @property
    def obs_words(self):
        """Words per observation row without the alignment padding: ceil(obs_len / 32)."""
        return (self.obs_len + 31) // 32

    def host_buffers(self):
        """Pinned host arrays for `step_host_async`: dict of torch CPU tensors
        actions int32[N], obs int32[N, obs_words], legal int32[N, legal_words], reward f32[N],
        step_type u8[N], next_actions int32[N]."""
        torch = self.torch
        n = self.n
        # the four outputs are views of ONE pinned allocation, back to back: with chunks=1 the step's
        # outputs then come down in a single copy
        sizes = [n * self.obs_words * 4, n * self.legal_words * 4, n * 4, n]
        blob = torch.zeros(sum(sizes), dtype=torch.uint8).pin_memory()
        o = [0, sizes[0], sizes[0] + sizes[1], sizes[0] + sizes[1] + sizes[2]]
        mk = lambda shape, dt: torch.zeros(shape, dtype=dt).pin_memory()
        return {"actions": mk((n,), torch.int32),
                "obs": blob[o[0]:o[1]].view(torch.int32).view(n, self.obs_words),
                "legal": blob[o[1]:o[2]].view(torch.int32).view(n, self.legal_words),
                "reward": blob[o[2]:o[3]].view(torch.float32),
                "step_type": blob[o[3]:],
                "next_actions": mk((n,), torch.int32), "outputs_blob": blob}

    def step_host_async(self, actions, obs=None, legal=None, reward=None, step_type=None, chunks=1):
        """`environment.step(actions)` with HOST arrays (torch CPU tensors, pinned for overlap), enqueued
        as a chunked pipeline; the arrays are valid after `host_wait()`."""
        p = lambda t: None if t is None else t.data_ptr()
        _lib.check(_lib.lib().hb_env_step_host_async(self._h, actions.data_ptr(), p(obs), p(legal),
                                                     p(reward), p(step_type), int(chunks)))

    def host_wait(self):
        _lib.check(_lib.lib().hb_env_host_wait(self._h))

    # ---- the same loop double-buffered (include/hanabi_b200.h: hb_env_step_host_begin) ----------
    def step_host_begin(self, actions):
        """First half of `environment.step(actions)` with host arrays: actions up + the step kernel."""
        _lib.check(_lib.lib().hb_env_step_host_begin(self._h, actions.data_ptr()))

    def step_host_outputs(self, obs=None, legal=None, reward=None, step_type=None):
        """Second half: the step's outputs start their way to the host arrays (alternate two sets);
        complete after `host_wait_outputs(slot)` / `host_wait()`.  Returns the staging slot."""
        p = lambda t: None if t is None else t.data_ptr()
        _lib.check(_lib.lib().hb_env_step_host_outputs(self._h, p(obs), p(legal), p(reward), p(step_type)))
        return int(_lib.lib().hb_env_host_slot(self._h))

    def host_wait_actions(self):
        """Returns when the agent's next actions (explore_host_async) are on the host."""
        _lib.check(_lib.lib().hb_env_host_wait_actions(self._h))

    def host_wait_outputs(self, slot):
        _lib.check(_lib.lib().hb_env_host_wait_outputs(self._h, int(slot)))

    def observe(self, offset=0, out=None):
        """HanabiObservation of the player `offset` seats after the mover, for every game
        (bit-packed rows + legal masks; `out` = (obs, legal) reuses caller-owned buffers)."""
        torch = self.torch
        if out is not None:
            obs, legal = out
        else:
            obs = torch.empty_like(self.obs_packed)
            legal = torch.empty_like(self.legal_packed)
        _lib.check(_lib.lib().hb_env_observe(self._h, int(offset), obs.data_ptr(), legal.data_ptr(),
                                             self._s()))
        return obs, legal

    OBSERVATION_TYPES = {"minimal": 0, "card_knowledge": 1, "seer": 2}

    def observe_as(self, observation_type, offset=0, unpack=False):
        """The observation of the player `offset` seats after the mover in another of HLE's observation
        types ("minimal": no card-knowledge section; "seer": own cards visible; "card_knowledge": what
        `observe` / `step` give).  Returns (rows, length in bits): bit-packed int32 [N, stride] CUDA rows, or
        int8 0/1 [N, length] with unpack=True."""
        torch = self.torch
        t = self.OBSERVATION_TYPES[observation_type] if isinstance(observation_type, str) else int(observation_type)
        L = _lib.lib()
        bits = int(L.hb_env_obs_len_as(self._h, t))
        stride = ((bits + 31) // 32 + 3) // 4 * 4
        rows = torch.empty((self.n, stride), dtype=torch.int32, device=self.device)
        _lib.check(L.hb_env_observe_as(self._h, t, int(offset), rows.data_ptr(), stride, self._s()))
        if not unpack:
            return rows, bits
        out = torch.empty((self.n, bits), dtype=torch.int8, device=self.device)
        _lib.check(L.hb_unpack_bits(rows.data_ptr(), stride, bits, out.data_ptr(), self.n, self._s()))
        return out, bits

    def unpack_obs(self, packed=None):
        torch = self.torch
        packed = self.obs_packed if packed is None else packed
        n = packed.shape[0]
        out = torch.empty((n, self.obs_len), dtype=torch.int8, device=self.device)
        _lib.check(_lib.lib().hb_env_unpack_obs(self._h, packed.data_ptr(), out.data_ptr(), n, self._s()))
        return out

    def unpack_legal(self, packed=None):
        torch = self.torch
        packed = self.legal_packed if packed is None else packed
        n = packed.shape[0]
        out = torch.empty((n, self.num_moves), dtype=torch.int8, device=self.device)
        _lib.check(_lib.lib().hb_env_unpack_legal(self._h, packed.data_ptr(), out.data_ptr(), n, self._s()))
        return out

    def observations_packed(self):
        """Same tuple with the bit-packed rows (zero copy): what DQNAgent takes on the fast path."""
        return (self, (self.obs_packed, self.legal_packed))

    def observations(self):
        """The `observations` tuple the agents index: observations[1] == (obs, legal) int8 0/1
        (rlax_rainbow.py:278,285); observations[0] is this environment (what the rule-based
        agents read the game state from, in place of HLE observation objects)."""
        return (self, (self.unpack_obs(), self.unpack_legal()))

    def illegal_count(self):
        out = C.c_int64()
        _lib.check(_lib.lib().hb_env_illegal_count(self._h, C.byref(out), self._s()))
        return int(out.value)

    def dump(self):
        torch = self.torch
        out = torch.empty((self.n, self.dump_len), dtype=torch.int32, device=self.device)
        _lib.check(_lib.lib().hb_env_dump(self._h, out.data_ptr(), self._s()))
        return out.cpu().numpy()

    def scores(self):
        return self.dump()[:, 6]

    def selfplay(self, agent, n_steps, step0=None, obs=None, legal=None, reward=None,
                 step_type=None, action=None):
        """Fused self-play: n_steps x {agent acts, env steps, observation encoded} in ONE launch
        with the game state kept in registers.  Buffers are [n_steps, N, ...] CUDA tensors
        (allocated when None is passed as the string "alloc"; None = that output is skipped)."""
        torch = self.torch
        T = int(n_steps)

        def buf(x, shape, dt):
            if isinstance(x, str) and x == "alloc":
                return torch.empty(shape, dtype=dt, device=self.device)
            return x
        obs = buf(obs, (T, self.n, self.obs_stride), torch.int32)
        legal = buf(legal, (T, self.n, self.legal_words), torch.int32)
        reward = buf(reward, (T, self.n), torch.float32)
        step_type = buf(step_type, (T, self.n), torch.uint8)
        action = buf(action, (T, self.n), torch.int32)
        agent._buffers(self)
        ids, fa, ia = agent.rule_args()
        if step0 is None:
            step0 = agent._step
        ptr = lambda t: None if t is None else t.data_ptr()  # noqa: E731
        _lib.check(_lib.lib().hb_env_selfplay(
            self._h, ids.ctypes.data, fa.ctypes.data, ia.ctypes.data, len(ids), agent.seed,
            int(step0), T, ptr(obs), ptr(legal), ptr(reward), ptr(step_type), ptr(action),
            agent._hist.data_ptr(), self._s()))
        agent._step = int(step0) + T
        return obs, legal, reward, step_type, action
