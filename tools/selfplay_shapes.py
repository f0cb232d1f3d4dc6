"""Fused self-play throughput per game shape (the parity-test configurations of BASELINE.json that are
not the bench line): env-steps/s and algorithmic GB/s, CUDA-event timed.

  python tools/selfplay_shapes.py [games] [launches]
"""
import sys
import torch
from hanabi_agents_b200.env import HanabiParallelEnv
from hanabi_agents_b200.rule_based import RulebasedAgent, predefined_rules

CASES = [("Hanabi-Full", 2), ("Hanabi-Full", 3), ("Hanabi-Full", 4), ("Hanabi-Full", 5), ("Hanabi-Small", 2)]
T = 64


def main():
    games = [int(sys.argv[1])] if len(sys.argv) > 1 else [32768, 65536, 262144]
    launches = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    for name, players in CASES:
        for n in games:
            env = HanabiParallelEnv(n, name, players, n_decks=2 * n, seed=1234)
            agent = RulebasedAgent(predefined_rules.piers_rules, seed=1234)
            env.reset()
            dev = env.device
            obs = torch.empty((T, n, env.obs_stride), dtype=torch.int32, device=dev)
            legal = torch.empty((T, n, env.legal_words), dtype=torch.int32, device=dev)
            reward = torch.empty((T, n), dtype=torch.float32, device=dev)
            stype = torch.empty((T, n), dtype=torch.uint8, device=dev)
            action = torch.empty((T, n), dtype=torch.int32, device=dev)
            run = lambda: env.selfplay(agent, T, obs=obs, legal=legal, reward=reward, step_type=stype, action=action)
            for _ in range(3):
                run()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(launches):
                run()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / launches
            bytes_step = env.obs_words * 4 + env.legal_words * 4 + 4 + 1 + 4
            alg = n * (T * bytes_step + 2 * env.state_words * 4)
            print("%-12s %dp  games %7d  %.3f ms/launch  %.2fe9 env-steps/s  %6.0f GB/s algorithmic (%d B/step)  illegal %d"
                  % (name, players, n, ms, n * T / ms / 1e6, alg / ms / 1e6, bytes_step, env.illegal_count()), flush=True)
            del env, obs, legal, reward, stype, action


main()
