#!/usr/bin/env python
"""selfplay_kernel throughput at several games-per-GPU counts (how it responds to resident warps)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hanabi_agents_b200.env import HanabiParallelEnv
from hanabi_agents_b200.rule_based import RulebasedAgent, predefined_rules
T = 64
for n in (65536, 131072, 262144):
    env = HanabiParallelEnv(n, "Hanabi-Full", 2, n_decks=2 * n, seed=1234)
    agent = RulebasedAgent(predefined_rules.piers_rules, seed=1234)
    env.reset()
    dev = env.device
    bufs = dict(obs=torch.empty((T, n, env.obs_stride), dtype=torch.int32, device=dev),
                legal=torch.empty((T, n, env.legal_words), dtype=torch.int32, device=dev),
                reward=torch.empty((T, n), dtype=torch.float32, device=dev),
                step_type=torch.empty((T, n), dtype=torch.uint8, device=dev),
                action=torch.empty((T, n), dtype=torch.int32, device=dev))
    for _ in range(4):
        env.selfplay(agent, T, **bufs)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 40
    e0.record()
    for _ in range(reps):
        env.selfplay(agent, T, **bufs)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print("games %7d: %.4f ms per 64-step launch, %.4g env-steps/s, illegal %d" % (n, ms, n * T / (ms * 1e-3), env.illegal_count()))
    del env, bufs
    torch.cuda.empty_cache()
