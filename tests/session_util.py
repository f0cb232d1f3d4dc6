"""Recording agents with a deterministic policy, for the session parity tests."""
import numpy as np

FIRST, MID, LAST = 0, 1, 2


def pick(legal_row, tick, game):
    """Deterministic legal move: the ((7 tick + 13 game + 3) mod n_legal)-th legal uid."""
    ids = np.flatnonzero(np.asarray(legal_row) != 0)
    if len(ids) == 0:
        return 0
    return int(ids[(7 * tick + 13 * game + 3) % len(ids)])


class OracleRecorder:
    """Agent protocol of oracle.session.OracleSession; logs every recorded transition."""

    def __init__(self, name, session_tick):
        self.name = name
        self.events = []          # (tick, game, step_type, action, reward, obs bytes, legal bytes)
        self._tick = session_tick     # callable -> current tick

    def requires_vectorized_observation(self):
        return True

    def add_experience_first(self, obs, legal, st):
        for g in np.flatnonzero(st == FIRST):
            self.events.append((self._tick(), int(g), "seed", 0, 0.0, obs[g].tobytes(), b""))

    def add_experience(self, obs, legal, actions, rewards, st):
        for g in np.flatnonzero(st != FIRST):
            self.events.append((self._tick(), int(g), int(st[g]), int(actions[g]), float(rewards[g]),
                                obs[g].tobytes(), legal[g].tobytes()))

    def explore(self, obs, legal, tick):
        return np.array([pick(legal[g], tick, g) for g in range(len(legal))], np.int32)

    exploit = explore


class DeviceRecorder:
    """Same log, with the verbs of the reference's agents (rlax_rainbow.py:277-308), fed by
    hanabi_agents_b200.session.HanabiParallelSession with device tensors."""

    def __init__(self, name, session_tick):
        self.name = name
        self.events = []
        self._tick = session_tick     # callable -> current tick

    def requires_vectorized_observation(self):
        return True

    @staticmethod
    def _unpack(observations):
        env, (obs, legal) = observations
        o = env.unpack_obs(obs).cpu().numpy().astype(np.uint8)
        l = env.unpack_legal(legal).cpu().numpy().astype(np.uint8)
        return o, l

    def add_experience_first(self, observations, step_types):
        o, _ = self._unpack(observations)
        st = step_types.cpu().numpy()
        for g in np.flatnonzero(st == FIRST):
            self.events.append((self._tick(), int(g), "seed", 0, 0.0, o[g].tobytes(), b""))

    def add_experience(self, observations, actions, rewards, step_types):
        o, l = self._unpack(observations)
        st = step_types.cpu().numpy()
        a, r = actions.cpu().numpy(), rewards.cpu().numpy()
        for g in np.flatnonzero(st != FIRST):
            self.events.append((self._tick(), int(g), int(st[g]), int(a[g]), float(r[g]),
                                o[g].tobytes(), l[g].tobytes()))

    def explore(self, observations):
        _, l = self._unpack(observations)
        t = self._tick()
        return np.array([pick(l[g], t, g) for g in range(len(l))], np.int32)

    exploit = explore


class DopamineRecorder:
    """begin_episode / step / end_episode recorder for oracle.session.run_one_episode
    (dqn_agent.py:284-332): what each PLAYER is told, in order."""

    def __init__(self, game):
        self.game = game
        self.move = 0
        self.per_player = {}
        self.last_action = {}

    def _act(self, player, legal):
        a = pick(legal, self.move, self.game)
        self.move += 1
        self.last_action[player] = a
        return a

    def begin_episode(self, player, legal, obs):
        self.per_player.setdefault(player, []).append(("seed", 0, 0.0, obs.tobytes()))
        return self._act(player, legal)

    def step(self, reward, player, legal, obs):
        self.per_player[player].append((MID, self.last_action[player], float(reward), obs.tobytes()))
        return self._act(player, legal)

    def end_episode(self, final_rewards):
        for player in sorted(self.per_player):
            self.per_player[player].append((LAST, self.last_action[player], float(final_rewards[player]), None))
