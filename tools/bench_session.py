"""DQN self-play through HanabiParallelSession: ms per tick at the cfg-4 shape
(Hanabi-Full 2p, 658-512-512-(20x51) noisy C51 net), one GPU.
usage: python tools/bench_session.py [n_games] [ticks] [updates_per_tick]"""
import sys
import time

import torch

sys.path.insert(0, ".")
from hanabi_agents_b200.env import HanabiParallelEnv          # noqa: E402
from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams  # noqa: E402
from hanabi_agents_b200.session import HanabiParallelSession   # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 200
upd = int(sys.argv[3]) if len(sys.argv) > 3 else 1
shared = "--shared" in sys.argv

env = HanabiParallelEnv(n, "Hanabi-Full", 2, n_decks=4 * n, seed=1234)
params = RlaxRainbowParams(train_batch_size=256, layers=[512, 512], experience_buffer_size=2 ** 21, seed=1)
a0 = DQNAgent(env.observation_spec(), env.action_spec(), params)
a1 = a0 if shared else DQNAgent(env.observation_spec(), env.action_spec(), params._replace(seed=2))
ses = HanabiParallelSession(env, [a0, a1])
ses.reset()


def timed(fn, reps):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, (time.perf_counter() - t0) * 1e3 / reps


ses.run(20, train=True)
a0.update(); a1.update()
dev, wall = timed(lambda: ses.step(train=False), ticks)
print(f"eval tick  (exploit + env step + accounting): {dev:.3f} ms device, {wall:.3f} ms wall -> {n / dev * 1e3:.3e} env-steps/s")
dev, wall = timed(lambda: ses.step(train=True), ticks)
print(f"train tick (+ replay insert, both seats)    : {dev:.3f} ms device, {wall:.3f} ms wall -> {n / dev * 1e3:.3e} env-steps/s")


def tick_and_learn():
    ses.step(train=True)
    for _ in range(upd):
        a0.update()
        if a1 is not a0:
            a1.update()


dev, wall = timed(tick_and_learn, ticks)
print(f"train tick + {upd} update(s) per agent        : {dev:.3f} ms device, {wall:.3f} ms wall -> {n / dev * 1e3:.3e} env-steps/s")
print("stats", ses.stats(), "illegal", env.illegal_count(), "replay", a0.experience.size, a1.experience.size)
from hanabi_agents_b200 import _lib  # noqa: E402
for a in (a0, a1):
    print("graph captures", int(_lib.lib().hb_learner_info(a._learner, 1)), "eager steps",
          int(_lib.lib().hb_learner_info(a._learner, 2)), "train_step", a.train_step)
