"""Small driver for ncu: acting forward on 65,536 packed rows (Rainbow 658-512-512-20x51) and a
few learner updates.  python tools/prof_qnet.py [act|learn] [reps]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams
    what = sys.argv[1] if len(sys.argv) > 1 else "act"
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    n = 65536
    env = HanabiParallelEnv(n, "Hanabi-Full", 2, seed=1234)
    params = RlaxRainbowParams(train_batch_size=256, layers=[512, 512], experience_buffer_size=2 ** 17)
    agent = DQNAgent(env.observation_spec(), env.action_spec(), params)
    env.reset()
    obs = env.observations_packed()
    if what == "act":
        for _ in range(reps):
            a = agent.explore(obs)
        torch.cuda.synchronize()
        print("acted", int(a.sum()))
        return
    agent.add_experience_first(obs, env.step_type)
    for _ in range(3):
        a = agent.explore(obs)
        env.step(a)
        obs = env.observations_packed()
        agent.add_experience(obs, a, env.reward, env.step_type)
        agent.add_experience_first(obs, env.step_type)
    np.random.seed(0)
    for _ in range(reps):
        m = agent.update()
    torch.cuda.synchronize()
    print("loss", float(m))


if __name__ == "__main__":
    main()
