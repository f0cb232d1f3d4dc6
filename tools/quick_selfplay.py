#!/usr/bin/env python
"""Quick timing + output checksum of the fused self-play kernel (bench configuration: piers,
Hanabi-Full 2p, 65,536 games, 64 steps per launch).  The checksum covers every output of the first
four launches after reset: a kernel variant that changes a bit changes the checksum.

  python tools/quick_selfplay.py [--launches 200] [--games 65536] [--preset piers] [--players 2]
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hanabi_agents_b200.env import HanabiParallelEnv
from hanabi_agents_b200.rule_based import RulebasedAgent, predefined_rules

ap = argparse.ArgumentParser()
ap.add_argument("--launches", type=int, default=200)
ap.add_argument("--games", type=int, default=65536)
ap.add_argument("--preset", default="piers")
ap.add_argument("--players", type=int, default=2)
ap.add_argument("--fuse", type=int, default=64)
a = ap.parse_args()
n, T = a.games, a.fuse
env = HanabiParallelEnv(n, "Hanabi-Full", a.players, n_decks=4 * n, seed=1234)
agent = RulebasedAgent(getattr(predefined_rules, a.preset + "_rules"), seed=1234)
env.reset()
dev = env.device
obs = torch.empty((T, n, env.obs_stride), dtype=torch.int32, device=dev)
legal = torch.empty((T, n, env.legal_words), dtype=torch.int32, device=dev)
reward = torch.empty((T, n), dtype=torch.float32, device=dev)
stype = torch.empty((T, n), dtype=torch.uint8, device=dev)
action = torch.empty((T, n), dtype=torch.int32, device=dev)
run = lambda: env.selfplay(agent, T, obs=obs, legal=legal, reward=reward, step_type=stype, action=action)
cs = 0
for _ in range(4):
    run()
    w = torch.arange(1, obs.numel() + 1, device=dev, dtype=torch.int64).view_as(obs)
    cs = (cs * 1000003 + int((obs.long() * (w % 65521)).sum().item()) + 3 * int(legal.long().sum().item())
          + 5 * int(action.long().sum().item()) + 7 * int(reward.double().sum().item())
          + 11 * int(stype.long().sum().item())) % (1 << 61)
    del w
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.launches):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.launches
print(json.dumps({"preset": a.preset, "players": a.players, "games": n, "ms_per_launch": ms,
                  "env_steps_per_s": n * T / (ms * 1e-3), "checksum": cs}))
