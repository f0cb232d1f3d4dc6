#!/usr/bin/env python
"""Where the host spends a training tick (cProfile over 300 ticks)."""
import cProfile, os, pstats, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hanabi_agents_b200.env import HanabiParallelEnv
from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams
from hanabi_agents_b200.session import HanabiParallelSession
n = 65536
env = HanabiParallelEnv(n, "Hanabi-Full", 2, n_decks=4 * n, seed=1234)
params = RlaxRainbowParams(train_batch_size=256, layers=[512, 512], experience_buffer_size=2 ** 20,
                           n_atoms=51, atom_vmax=25, seed=1234)
agents = [DQNAgent(env.observation_spec(), env.action_spec(), params),
          DQNAgent(env.observation_spec(), env.action_spec(), params._replace(seed=1235))]
ses = HanabiParallelSession(env, agents)
ses.reset()
ses.run(40, train=True)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
ses.run(300, train=True)
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(28)
