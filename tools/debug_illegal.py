import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from hanabi_agents_b200.env import HanabiParallelEnv
from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams
from hanabi_agents_b200.session import HanabiParallelSession
n = int(sys.argv[1]); shared = "--shared" in sys.argv; learn = "--nolearn" not in sys.argv
env = HanabiParallelEnv(n, "Hanabi-Full", 2, n_decks=4 * n, seed=1234)
params = RlaxRainbowParams(train_batch_size=256, layers=[512, 512], experience_buffer_size=2 ** 21, seed=1)
a0 = DQNAgent(env.observation_spec(), env.action_spec(), params)
a1 = a0 if shared else DQNAgent(env.observation_spec(), env.action_spec(), params._replace(seed=2))
ses = HanabiParallelSession(env, [a0, a1])
ses.reset()
prev = 0
for t in range(340):
    lg = env.legal_packed.clone(); st = env.step_type.clone(); obs = env.obs_packed.clone()
    act = ses.step(train=True)
    if learn and t >= 20:
        a0.update()
        if a1 is not a0: a1.update()
    ill = env.illegal_count()
    if ill != prev:
        a = act.cpu().numpy(); l = lg.cpu().numpy().view(np.uint32)[:, 0]; s = st.cpu().numpy()
        bad = np.flatnonzero((((l >> a) & 1) == 0) & (s != 2))
        print("tick", t, "illegal", ill, "rows", bad[:10], "actions", a[bad[:10]], "legal", [hex(x) for x in l[bad[:10]]], "st", s[bad[:10]])
        ag = ses.agents[t % 2]
        q = torch.empty((n, 20), device="cuda")
        from hanabi_agents_b200 import _lib
        for r in bad[:3]:
            qq = ag.network.forward(obs[r:r + 1])
            print("q row", r, qq)
        prev = ill
    if t % 50 == 0:
        print(t, ses.stats(), float(a0.last_loss.item()) if a0.last_loss is not None else None)
print("final illegal", env.illegal_count(), ses.stats())
