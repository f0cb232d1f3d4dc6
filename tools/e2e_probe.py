#!/usr/bin/env python
"""Where the host-in-the-loop step time goes: the double-buffered loop with / without the output copy,
the output copy alone, and the single-buffered loop."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hanabi_agents_b200.env import HanabiParallelEnv
from hanabi_agents_b200.rule_based import RulebasedAgent, predefined_rules

n = 65536
env = HanabiParallelEnv(n, "Hanabi-Full", 2, n_decks=4 * n, seed=1234)
agent = RulebasedAgent(predefined_rules.piers_rules, seed=1234)
env.reset()
sets = [env.host_buffers(), env.host_buffers()]
acts = [sets[0]["actions"], sets[0]["next_actions"], sets[1]["actions"]]
acts[0].copy_(agent.explore(env).cpu())
torch.cuda.synchronize()
T = [0]

def step(outputs=True, wait_all=False):
    t = T[0]
    cur, nxt, hs = acts[t % 3], acts[(t + 1) % 3], sets[t & 1]
    env.step_host_begin(cur)
    agent.explore_host_async(env, nxt)
    if outputs:
        env.step_host_outputs(hs["obs"], hs["legal"], hs["reward"], hs["step_type"])
    if wait_all:
        env.host_wait()
    else:
        env.host_wait_actions()
    T[0] = t + 1

def timed(fn, k=300):
    for _ in range(5):
        fn()
    env.host_wait()
    t0 = time.perf_counter()
    for _ in range(k):
        fn()
    env.host_wait()
    return (time.perf_counter() - t0) / k * 1e6

print("double-buffered, outputs, wait actions only : %.1f us/step" % timed(lambda: step(True, False)))
print("double-buffered, outputs, wait everything   : %.1f us/step" % timed(lambda: step(True, True)))
print("no output copy (small dependent work only)  : %.1f us/step" % timed(lambda: step(False, False)))
dev = torch.empty(sets[0]["outputs_blob"].numel(), dtype=torch.uint8, device="cuda")
s = torch.cuda.Stream()
def copy_only():
    with torch.cuda.stream(s):
        sets[0]["outputs_blob"].copy_(dev, non_blocking=True)
for _ in range(5):
    copy_only()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(300):
    copy_only()
torch.cuda.synchronize()
print("output copy alone (%.2f MB)                  : %.1f us" % (dev.numel() / 1e6, (time.perf_counter() - t0) / 300 * 1e6))
