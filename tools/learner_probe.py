#!/usr/bin/env python
"""Is DQNAgent.update bound by the host (numpy keys, ctypes, graph launches) or by the device?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from hanabi_agents_b200.env import HanabiParallelEnv
from hanabi_agents_b200.rule_based import RulebasedAgent, predefined_rules
from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams

n, T = 65536, 8
env = HanabiParallelEnv(n, "Hanabi-Full", 2, n_decks=4 * n, seed=1234)
ra = RulebasedAgent(predefined_rules.piers_rules, seed=1234)
env.reset()
obs, legal, reward, stype, action = env.selfplay(ra, T, obs="alloc", legal="alloc", reward="alloc",
                                                 step_type="alloc", action="alloc")
params = RlaxRainbowParams(train_batch_size=256, layers=[512, 512], experience_buffer_size=2 ** 19,
                           n_atoms=51, atom_vmax=25, seed=1234)
agent = DQNAgent(env.observation_spec(), env.action_spec(), params)
for t in range(T - 1):
    agent.experience.add_transitions(obs[t], action[t + 1], reward[t + 1], obs[t + 1], legal[t + 1],
                                     (stype[t + 1] == 2).to(torch.uint8))
np.random.seed(0)
for _ in range(10):
    agent.update()
torch.cuda.synchronize()
k = 500
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for _ in range(k):
    agent.update()
e1.record()
t_host = time.perf_counter() - t0
torch.cuda.synchronize()
t_all = time.perf_counter() - t0
print("host enqueue %.1f us/update, wall %.1f us/update, device events %.1f us/update"
      % (t_host / k * 1e6, t_all / k * 1e6, e0.elapsed_time(e1) / k * 1e3))
# host-only cost of the key generation
from hanabi_agents_b200.rlax_dqn.buffers import PriorityBuffer
t0 = time.perf_counter()
for _ in range(k):
    keys = np.ascontiguousarray(PriorityBuffer.stratified_keys(256, None), dtype=np.float32)
print("stratified keys %.1f us" % ((time.perf_counter() - t0) / k * 1e6))
# pure host cost: three updates right after a device synchronisation (the staging ring has four slots)
hs = []
for _ in range(50):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        agent.update()
    hs.append((time.perf_counter() - t0) / 3 * 1e6)
print("host cost of one update call (device idle): median %.1f us" % sorted(hs)[len(hs) // 2])
# the two phases separately (device events around 300 calls of each)
from hanabi_agents_b200 import _lib
L = _lib.lib()
stream = _lib.current_stream_ptr()
keys = np.ascontiguousarray(PriorityBuffer.stratified_keys(256, None), dtype=np.float32)
def phase_a():
    _lib.check(L.hb_learner_grads(agent._learner, keys.ctypes.data, None, 0.5, 0.99, 0, stream))
def phase_b():
    _lib.check(L.hb_learner_apply(agent._learner, 1e-4, 0.9, 0.999, agent.adam_eps, 1.0, stream))
def both():
    phase_a()
    phase_b()
def fused():
    _lib.check(L.hb_learner_step(agent._learner, keys.ctypes.data, None, 0.5, 0.99, 0, 1e-4, 0.9, 0.999,
                                 agent.adam_eps, stream))
for name, fn in (("phase A (sample..grads, priorities)", phase_a), ("phase B (adam)", phase_b),
                 ("phase A + phase B (data-parallel path, no all-reduce)", both),
                 ("hb_learner_step (one graph, per-layer Adam)", fused)):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(300):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print("%s: %.1f us" % (name, e0.elapsed_time(e1) / 300 * 1e3))
