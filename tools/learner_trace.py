#!/usr/bin/env python
"""Timeline of one learner update as the device ran it (CUPTI through torch.profiler: start, duration and
stream of every kernel of the captured graph).  python tools/learner_trace.py [n_updates]"""
import json, os, sys, tempfile
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
for _ in range(50):
    agent.update()
torch.cuda.synchronize()
k = int(sys.argv[1]) if len(sys.argv) > 1 else 6
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(k):
        agent.update()
    torch.cuda.synchronize()
path = os.path.join(tempfile.mkdtemp(), "trace.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy")]
ev.sort(key=lambda e: e["ts"])
starts = [i for i, e in enumerate(ev) if "sample_gather" in e["name"]]
print("kernels traced: %d, updates: %d" % (len(ev), len(starts)))
if len(starts) >= 3:
    a, b = starts[-2], starts[-1]
    t0 = ev[a]["ts"]
    print("update period %.1f us" % (ev[b]["ts"] - t0))
    for e in ev[a:b]:
        nm = e["name"].replace("(anonymous namespace)::", "").replace("void ", "")
        print("  +%6.1f  %5.1f us  s%-3s %s" % (e["ts"] - t0, e["dur"], e.get("args", {}).get("stream", "?"), nm[:70]))
