"""CPU restatement of the self-play session -- TEST INFRASTRUCTURE (never imported by the product).

Two things live here:

* `run_one_episode(env, game, agent)`: ONE game driven exactly as
  hanabi_agents/rainbow/run_experiment.py:280-336 drives it (begin_episode / step / end_episode,
  `has_played`, `reward_since_last_action`), over the oracle environment.  It pins the batched
  restatement below: parity unpinned beyond that, because the session the rlax agents are run
  by (`hanabi_multiagent_framework`) is not part of the reference checkout.

* `OracleSession`: the same accounting for N lockstep games with the env's auto-reset tick, in
  per-game Python loops -- what hanabi_agents_b200.session.HanabiParallelSession must reproduce.
"""
import numpy as np

FIRST, MID, LAST = 0, 1, 2


def run_one_episode(env, agent, lenient=False):
    """run_experiment.py:280-336 for a one-game oracle env (N == 1).  `agent` has the dopamine
    verbs begin_episode(player, legal, obs) / step(reward, player, legal, obs) /
    end_episode(final_rewards).  Returns (step_number, total_reward)."""
    assert env.n == 1
    players = env.game.players
    obs, legal, _ = env.reset()
    current_player = 0                                     # HLE: player 0 starts
    action = agent.begin_episode(current_player, legal[0].copy(), obs[0].copy())
    is_done = False
    total_reward = 0.0
    step_number = 0
    has_played = {current_player}
    reward_since_last_action = np.zeros(players)
    while not is_done:
        obs, legal, reward, step_type = env.step(np.array([action], np.int32))
        is_done = step_type[0] == LAST
        r = float(reward[0])
        modified_reward = max(r, 0) if lenient else r
        total_reward += modified_reward
        reward_since_last_action += modified_reward
        step_number += 1
        if is_done:
            break
        current_player = (current_player + 1) % players
        if current_player in has_played:
            action = agent.step(reward_since_last_action[current_player], current_player,
                                legal[0].copy(), obs[0].copy())
        else:
            action = agent.begin_episode(current_player, legal[0].copy(), obs[0].copy())
            has_played.add(current_player)
        reward_since_last_action[current_player] = 0
    agent.end_episode(reward_since_last_action.copy())
    return step_number, total_reward


class OracleSession:
    """agents[k]: object with requires_vectorized_observation(), add_experience_first(obs, legal,
    st), add_experience(obs, legal, actions, rewards, st), explore(obs, legal, tick) and
    exploit(...) -- numpy arrays, one row per game."""

    def __init__(self, env, agents, lenient_score=False):
        self.env, self.agents = env, list(agents)
        self.P, self.n = env.game.players, env.n
        assert len(self.agents) == self.P
        self.lenient = bool(lenient_score)
        self.reset()

    def reset(self):
        self.env.reset()
        self.has_played = [set() for _ in range(self.n)]
        self.rsl = np.zeros((self.P, self.n), np.float32)       # reward_since_last_action
        self.last_action = np.zeros((self.P, self.n), np.int32)
        self.total = np.zeros(self.n, np.float32)
        self.steps = np.zeros(self.n, np.int32)
        self.episodes = []                                       # (game, length, return) in closing order
        self.tick = 0

    def step_types(self, k):
        st = np.zeros((self.P, self.n), np.uint8)
        for g in range(self.n):
            if self.env.step_type[g] == LAST:
                for j in range(self.P):
                    st[j, g] = LAST if j in self.has_played[g] else FIRST
            else:
                st[k, g] = MID if k in self.has_played[g] else FIRST
        return st

    def step(self, train=True):
        env, k = self.env, self.tick % self.P
        agent = self.agents[k]
        st = self.step_types(k)
        obs, legal = env.obs.copy(), env.legal.copy()
        if train:
            if agent.requires_vectorized_observation():
                agent.add_experience_first(obs, legal, st[k].copy())
                agent.add_experience(obs, legal, self.last_action[k].copy(), self.rsl[k].copy(), st[k].copy())
            for j in range(self.P):
                other = self.agents[j]
                if j == k or not other.requires_vectorized_observation():
                    continue
                o, l = env.observe((j - k) % self.P)
                other.add_experience(o, l, self.last_action[j].copy(), self.rsl[j].copy(), st[j].copy())
            actions = agent.explore(obs, legal, self.tick)
        else:
            actions = agent.exploit(obs, legal, self.tick)
        actions = np.asarray(actions, np.int32)
        prev = env.step_type.copy()
        _, _, reward, _ = env.step(actions)
        for g in range(self.n):
            if prev[g] == LAST:
                self.episodes.append((g, int(self.steps[g]), float(self.total[g])))
                self.total[g] = 0
                self.steps[g] = 0
                self.has_played[g] = set()
                self.rsl[:, g] = 0
            else:
                r = float(reward[g])
                if self.lenient:
                    r = max(r, 0.0)
                self.last_action[k, g] = actions[g]
                self.has_played[g].add(k)
                self.rsl[k, g] = 0
                self.rsl[:, g] += r
                self.total[g] += r
                self.steps[g] += 1
        self.tick += 1
        return actions
