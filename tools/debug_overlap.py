#!/usr/bin/env python
"""Diagnose overlapping self-play launches: which oracle step count does each game's state match?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from hanabi_agents_b200.env import HanabiParallelEnv
from hanabi_agents_b200.game import make_decks
from hanabi_agents_b200.rule_based import RulebasedAgent, predefined_rules
from oracle.hanabi import HanabiOracleEnv, PRESETS, game_params

n, pool, plan = 65536, 4 * 65536, [64, 36]
gp = game_params("Hanabi-Full", 2)
decks = make_decks(5, 5, pool, 3)
env = HanabiParallelEnv(n, "Hanabi-Full", 2, decks=decks)
ora = HanabiOracleEnv(n, threads=32, **gp)
ora.set_decks(decks)
agent = RulebasedAgent(predefined_rules.piers_rules, seed=77)
env.reset(); ora.reset()
for k in plan:
    env.selfplay(agent, k)
d = env.dump()
snaps = {}
for t in range(140):
    oa, _ = ora.rules_act(PRESETS["piers"], 77, t)
    ora.step(oa)
    if t + 1 in (36, 64, 100, 128, 136):
        snaps[t + 1] = ora.dump().copy()
bad = np.nonzero((d != snaps[100]).any(axis=1))[0]
print("bad games", len(bad), "groups", sorted(set(bad // 64))[:20], "n groups", len(set(bad // 64)))
for T, s in snaps.items():
    print("  matches oracle@%d: %d" % (T, int((d[bad] == s[bad]).all(axis=1).sum())))
g = bad[0]
print("game", g, "gpu", d[g][:20], "ora100", snaps[100][g][:20], "ora64", snaps[64][g][:20])
