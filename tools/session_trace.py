#!/usr/bin/env python
"""Device timeline of one self-play session tick (CUPTI through torch.profiler).
python tools/session_trace.py [train|eval|learn]"""
import json, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from hanabi_agents_b200.env import HanabiParallelEnv
from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams
from hanabi_agents_b200.session import HanabiParallelSession

mode = sys.argv[1] if len(sys.argv) > 1 else "train"
n = 65536
env = HanabiParallelEnv(n, "Hanabi-Full", 2, n_decks=4 * n, seed=1234)
params = RlaxRainbowParams(train_batch_size=256, layers=[512, 512], experience_buffer_size=2 ** 20,
                           n_atoms=51, atom_vmax=25, seed=1234)
agents = [DQNAgent(env.observation_spec(), env.action_spec(), params),
          DQNAgent(env.observation_spec(), env.action_spec(), params._replace(seed=1235))]
ses = HanabiParallelSession(env, agents)
ses.reset()
ses.run(24, train=True)
for _ in range(8):
    for a in agents:
        a.update()


def tick():
    ses.step(train=(mode != "eval"))
    if mode == "learn":
        ses.update_learners()


for _ in range(20):
    tick()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(50):
    tick()
th = time.perf_counter() - t0
torch.cuda.synchronize()
print("%s tick: host enqueue %.1f us, wall %.1f us" % (mode, th / 50 * 1e6, (time.perf_counter() - t0) / 50 * 1e6))
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(4):
        tick()
    torch.cuda.synchronize()
path = os.path.join(tempfile.mkdtemp(), "trace.json")
prof.export_chrome_trace(path)
tr = json.load(open(path))["traceEvents"]
ev = [e for e in tr if e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy")]
ev.sort(key=lambda e: e["ts"])
sync = [e for e in tr if e.get("cat") == "cuda_runtime" and ("Synchronize" in e["name"] or "cudaMemcpy" == e["name"])]
print("kernels traced: %d; host-side synchronising runtime calls in 4 ticks: %d (%s)"
      % (len(ev), len(sync), sorted(set(e["name"] for e in sync))))
# one tick = from one env step kernel to the next
starts = [i for i, e in enumerate(ev) if "env_step_kernel" in e["name"]]
if len(starts) >= 3:
    a, b = starts[-2], starts[-1]
    t0 = ev[a]["ts"]
    print("tick period %.1f us" % (ev[b]["ts"] - t0))
    for e in ev[a:b]:
        nm = e["name"].replace("(anonymous namespace)::", "").replace("void ", "")
        print("  +%6.1f  %6.1f us  s%-3s %s" % (e["ts"] - t0, e["dur"], e.get("args", {}).get("stream", "?"), nm[:80]))
