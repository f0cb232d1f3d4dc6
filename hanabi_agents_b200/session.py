"""Self-play driver: N lockstep Hanabi games on one B200, >= 2 agents taking turns.

The reference's agents are driven by an external session loop; its in-repo analogue is
hanabi_agents/rainbow/run_experiment.py:280-436 (run_one_episode / run_one_phase /
run_one_iteration).  This is that loop batched over every game of a HanabiParallelEnv, with the
per-player accounting (`has_played`, `reward_since_last_action`, episode return and length,
:296-331) kept on the device by two small kernels per tick (csrc/session.cu), so a tick is

    begin_turn -> seat.add_experience_first/add_experience -> seat.explore -> env.step -> end_turn

with no host round trip of per-game data.

Seats, not players: seat k = tick mod P moves in every game.  A game that ended spends one tick
on the env's reset step (its mover's action is ignored), which keeps all games on the same
rotation; in that tick EVERY seat that played the episode is handed the LAST observation from its
own point of view, its last action and the reward since -- what `agent.end_episode(
reward_since_last_action)` (:335) delivers in the reference.
"""
import ctypes as C
import os

import numpy as np

from . import _lib

FIRST, MID, LAST = 0, 1, 2


class HanabiParallelSession:
    def __init__(self, env, agents, lenient_score=False):
        import torch
        self.torch = torch
        self.env = env
        self.agents = list(agents)
        if len(self.agents) != env.players:
            raise ValueError("need one agent per seat: %d agents for %d players"
                             % (len(self.agents), env.players))
        self.P, self.n = env.players, env.n
        self.lenient_score = bool(lenient_score)   # LENIENT_SCORE  run_experiment.py:45,306
        self._h = C.c_void_p()
        L = _lib.lib()
        _lib.check(L.hb_session_create(self.P, self.n, env.device_index, C.byref(self._h)))

        def view(which, seat, typestr, n=self.n):
            return _lib.dev_tensor(L.hb_session_buffer(self._h, which, seat), n, typestr, self, env.device)
        self.reward_since_last_action = [view(0, k, "<f4") for k in range(self.P)]
        self.last_action = [view(1, k, "<i4") for k in range(self.P)]
        self.step_types = [view(2, k, "|u1") for k in range(self.P)]
        self.has_played = [view(3, k, "|u1") for k in range(self.P)]
        self.episode_return = view(4, 0, "<f4")
        self.episode_length = view(5, 0, "<i4")
        self.episodes_done = view(6, 0, "<i4")
        self.first_return = view(7, 0, "<f4")
        self.first_length = view(8, 0, "<i4")
        # an agent object sitting at several seats needs one `last_obs` per seat
        self._seat_last_obs = {}
        for k, a in enumerate(self.agents):
            if hasattr(a, "last_obs") and sum(1 for b in self.agents if b is a) > 1:
                self._seat_last_obs[k] = torch.zeros_like(a.last_obs)
        self._learners = []
        for a in self.agents:
            if a.requires_vectorized_observation() and all(a is not b for b in self._learners):
                self._learners.append(a)
        self.tick = 0
        self.is_reset = False
        self._side_obs = None
        # Training ticks: the replay inserts of a tick (both seats' masked appends, their sum-tree range
        # updates, two host round trips for the row counts) do not feed the move that is chosen in it, so
        # they run on a second stream beside the acting pass (HB_SESSION_OVERLAP=0: one after the other, the
        # reference's order of calls).  The acting GEMMs are persistent CTAs that take whole SMs: they leave
        # HB_SESSION_RESERVE_SMS (default 16 of the 148) of them to the inserts, for that call only.
        self._overlap = os.environ.get("HB_SESSION_OVERLAP", "1") != "0"
        self._ins_stream = None
        self._ins_event = None
        self._seat_streams = []
        self._reserve = int(os.environ.get("HB_SESSION_RESERVE_SMS", "16"))
        # HB_SESSION_ASYNC_INSERT=0: the inserts report their row counts to the host (two waits per tick)
        self._async_inserts = os.environ.get("HB_SESSION_ASYNC_INSERT", "1") != "0"

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                _lib.lib().hb_session_destroy(h)
            except Exception:
                pass
            self._h = None

    # -- helpers ----------------------------------------------------------------------------
    def _s(self):
        return _lib.current_stream_ptr()

    def _bind(self, seat):
        lo = self._seat_last_obs.get(seat)
        if lo is not None:
            self.agents[seat].last_obs = lo

    def _obs_for(self, agent, offset=0):
        env = self.env
        if not agent.requires_vectorized_observation():
            return (env, None)
        if offset == 0:
            return env.observations_packed()
        if self._side_obs is None:
            self._side_obs = {}
        if offset not in self._side_obs:     # one pair per offset: the seats' inserts may run side by side
            self._side_obs[offset] = (self.torch.empty_like(env.obs_packed),
                                      self.torch.empty_like(env.legal_packed))
        return (env, env.observe(offset, out=self._side_obs[offset]))

    # -- control ----------------------------------------------------------------------------
    def reset(self, clear_stats=True):
        """environment.reset() for every game; nobody has played, no reward pending (:292-299)."""
        self.env.reset()
        _lib.check(_lib.lib().hb_session_reset(self._h, int(clear_stats), self._s()))
        self.tick = 0
        self.is_reset = True

    def step(self, train=True):
        """One tick: seat `tick mod P` moves in every game.  Returns its actions (CUDA int32)."""
        if not self.is_reset:
            self.reset()
        L = _lib.lib()
        env, k = self.env, self.tick % self.P
        agent = self.agents[k]
        _lib.check(L.hb_session_begin_turn(self._h, k, env.step_type.data_ptr(), self._s()))
        obs = self._obs_for(agent)

        def inserts():
            if agent.requires_vectorized_observation():
                self._bind(k)
                agent.add_experience_first(obs, self.step_types[k])
                agent.add_experience(obs, self.last_action[k], self.reward_since_last_action[k],
                                     self.step_types[k])
            for j in range(self.P):          # end_episode for the seats that are not up (:335)
                other = self.agents[j]
                if j == k or not other.requires_vectorized_observation():
                    continue
                self._bind(j)
                other.add_experience(self._obs_for(other, (j - k) % self.P), self.last_action[j],
                                     self.reward_since_last_action[j], self.step_types[j])
            self._bind(k)

        if train and self._overlap and self._learners:
            torch = self.torch
            main = torch.cuda.current_stream(env.device)
            if self._ins_stream is None:
                self._ins_stream = torch.cuda.Stream(device=env.device)
                self._ins_event = torch.cuda.Event()
            self._ins_event.record(main)             # observations and per-seat accounting are in place
            self._bind(k)
            net = getattr(getattr(agent, "network", None), "_h", None)
            if net is not None:
                _lib.check(L.hb_net_set_act_reserve(net, self._reserve))
            try:
                actions = agent.explore(obs)         # enqueued first: the inserts' host waits run under it
            finally:
                if net is not None:
                    _lib.check(L.hb_net_set_act_reserve(net, 0))
            seats = [j for j in [k] + [j for j in range(self.P) if j != k]
                     if self.agents[j].requires_vectorized_observation()]
            split = (len(set(id(self.agents[j]) for j in seats)) == len(seats) and
                     all(hasattr(self.agents[j], "add_experience_begin") for j in seats))
            if not split:                            # a shared agent / another agent type: one after the other
                with torch.cuda.stream(self._ins_stream):
                    self._ins_stream.wait_event(self._ins_event)
                    inserts()
                main.wait_stream(self._ins_stream)   # the env step overwrites what the inserts read
            else:
                # one stream per seat; every seat's row count is started before the host waits for the first
                while len(self._seat_streams) < self.P:
                    self._seat_streams.append(torch.cuda.Stream(device=env.device))
                # the environment's packed device rows need no torch work on the way in: the library's launches
                # are pointed at the seat's stream directly (torch.cuda.stream(...) costs 15 us a time)
                packed = obs[1] is not None and getattr(obs[1][0], "is_cuda", False) and obs[1][0].dtype == torch.int32

                def on(j):
                    st = self._seat_streams[j]
                    return _lib.use_stream(st.cuda_stream) if packed else torch.cuda.stream(st)
                seat_obs = {}
                nosync = self._async_inserts and all(hasattr(self.agents[j], "add_experience_async") for j in seats)
                for j in seats:
                    with on(j):
                        self._seat_streams[j].wait_event(self._ins_event)
                        if j == k:
                            agent.add_experience_first(obs, self.step_types[k])
                            seat_obs[j] = obs
                        else:
                            seat_obs[j] = self._obs_for(self.agents[j], (j - k) % self.P)
                        if nosync:   # ring cursor on the device: the tick has no host wait at all
                            self.agents[j].add_experience_async(seat_obs[j], self.last_action[j],
                                                                self.reward_since_last_action[j],
                                                                self.step_types[j])
                        else:
                            self.agents[j].add_experience_begin(seat_obs[j], self.step_types[j])
                if not nosync:
                    for j in seats:
                        with on(j):
                            self.agents[j].add_experience_finish(seat_obs[j], self.last_action[j],
                                                                 self.reward_since_last_action[j],
                                                                 self.step_types[j])
                for j in seats:
                    main.wait_stream(self._seat_streams[j])
        elif train:
            inserts()
            actions = agent.explore(obs)
        else:
            actions = agent.exploit(obs)
        if not (hasattr(actions, "is_cuda") and actions.is_cuda):
            actions = self.torch.as_tensor(np.ascontiguousarray(actions, dtype=np.int32), device=env.device)
        actions = actions.to(self.torch.int32).contiguous()
        env.step(actions)
        _lib.check(L.hb_session_end_turn(self._h, k, actions.data_ptr(), env.reward.data_ptr(),
                                         int(self.lenient_score), self._s()))
        self.tick += 1
        return actions

    def run(self, n_ticks, train=True):
        for _ in range(int(n_ticks)):
            self.step(train)

    def stats(self):
        """Episodes closed since the last clear: count, mean / std of the undiscounted return,
        mean length (the 'episode_returns' / 'episode_lengths' statistics of run_one_phase :359-368)."""
        out = (C.c_double * 4)()
        _lib.check(_lib.lib().hb_session_stats(self._h, out, self._s()))
        s, n, ln, sq = (float(x) for x in out)
        mean = s / n if n else float("nan")
        var = max(sq / n - mean * mean, 0.0) if n else float("nan")
        return {"episodes": int(n), "mean_return": mean, "std_return": var ** 0.5,
                "mean_length": ln / n if n else float("nan"), "sum_returns": s}

    def clear_stats(self):
        self.episodes_done.zero_()
        _lib.dev_tensor(_lib.lib().hb_session_buffer(self._h, 9, 0), 4, "<f8", self, self.env.device).zero_()

    def update_learners(self):
        """One learner update of every learning agent (run_experiment.py:323-327 loops over the agents).
        The agents share nothing (own network, own replay shard), and one update is a chain of small
        kernels that leaves most of the device idle: in a single-process job they run side by side, each
        from its own stream.  Data-parallel jobs keep them in line (one collective order for all ranks)."""
        ls = self._learners
        side = self._overlap and len(ls) > 1 and all(
            getattr(a, "_world", None) is not None and a._world()[0] == 1 for a in ls)
        if not side:
            for a in ls:
                a.update()
            return
        torch = self.torch
        main = torch.cuda.current_stream(self.env.device)
        if self._ins_event is None:
            self._ins_event = torch.cuda.Event()
        while len(self._seat_streams) < len(ls):
            self._seat_streams.append(torch.cuda.Stream(device=self.env.device))
        self._ins_event.record(main)
        for i, a in enumerate(ls):
            with _lib.use_stream(self._seat_streams[i].cuda_stream):   # update() launches through the library only
                self._seat_streams[i].wait_event(self._ins_event)
                a.update()
        for i in range(len(ls)):
            main.wait_stream(self._seat_streams[i])

    def train(self, n_iter, n_sim_steps, n_train_steps, n_warmup=0):
        """n_iter x {n_sim_steps ticks of acting + recording, n_train_steps learner updates of
        every learning agent}; the first n_warmup ticks only fill the replay shards."""
        if not self.is_reset:
            self.reset()
        self.run(n_warmup, train=True)
        for _ in range(int(n_iter)):
            self.run(n_sim_steps, train=True)
            for _ in range(int(n_train_steps)):
                self.update_learners()
        return self.stats()

    def run_eval(self, check_every=16, max_ticks=4096):
        """Greedy evaluation (agent.eval_mode = True, run_experiment.py:417-431): fresh games, every
        seat exploits, until each of the N games has finished once.  Returns mean return / length of
        those first episodes.  Training state of the games is discarded (as a new episode is there)."""
        torch = self.torch
        self.reset(clear_stats=True)
        ticks = 0
        while ticks < max_ticks:
            self.run(check_every, train=False)
            ticks += check_every
            if bool((self.episodes_done > 0).all().item()):
                break
        done = self.episodes_done > 0
        n = int(done.sum().item())
        ret = self.first_return[done].to(torch.float64)
        ln = self.first_length[done].to(torch.float64)
        self.is_reset = False
        return {"episodes": n, "eval_episode_returns": float(ret.mean().item()) if n else float("nan"),
                "eval_episode_lengths": float(ln.mean().item()) if n else float("nan"), "ticks": ticks}
