#!/usr/bin/env python
"""Headline benchmark: Hanabi env-steps/s incl. observation encoding (BASELINE.json metric).

Workload (BASELINE.json configs[1]): rule-based (piers) self-play, Hanabi-Full 2 players,
65,536 parallel games PER GPU, every env step = rule agent move + HanabiState transition +
chance deal + canonical observation (bit-packed) + legal-move mask + reward + step type.
One bench "step" = one fused launch = --fuse (64) env steps of every game, for both arms.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun ... bench.py --gpus N ...     (one rank per GPU, games sharded, no data collective)

Prints ONE JSON line (contract in the task statement).  `--impl reference` times the CPU
restatement of the reference path (oracle/) on the host cores: the reference's own engine is
the third-party HLE, which is absent, so kind = "port".
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GAMES_PER_GPU = 65536
WORKLOAD = "rule_based(piers) self-play, Hanabi-Full 2p, 65536 games/GPU, env step + obs encoding"
METRIC = "hanabi_env_steps_per_sec_incl_obs_encoding"
UNIT = "env-steps/s"
SEED = 1234


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (profiling recipe): one
    `nvidia-smi -lms` process started before the region and stopped after it."""
    Q = ("index,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index, period_ms=20):
        self.index, self.period, self.proc, self.rows = index, period_ms, None, []
        self.t0 = self.t1 = None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=timestamp," + self.Q,
                 "--format=csv,noheader,nounits", "-lms", str(self.period)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            time.sleep(0.3)          # let the first samples arrive before the region starts
        except Exception:
            self.proc = None
        return self

    def mark_start(self):
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    def __exit__(self, *a):
        if self.proc is None:
            return
        time.sleep(2.5 * self.period / 1e3)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        import datetime
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
            except Exception:
                ts = None
            self.rows.append((ts, f[1:]))

    def summary(self):
        rows = [r for ts, r in self.rows
                if ts is None or self.t0 is None or (self.t0 - 0.05 <= ts <= self.t1 + 0.05)]
        if not rows:
            rows = [r for _, r in self.rows[-3:]]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        num = lambda x: x.replace(".", "").isdigit()  # noqa: E731
        sm = sorted(float(r[1]) for r in rows if num(r[1]))
        mx = max([float(r[2]) for r in rows if num(r[2])] or [0.0])
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": reasons,
                "samples": len(rows)}


def cpu_port_throughput(n_games, seconds, threads, steps_cap=None, warm=2):
    """env-steps/s of the CPU restatement (rule agent + env step + encoding), `threads` threads."""
    from oracle.hanabi import HanabiOracleEnv, PRESETS, game_params, make_decks
    e = HanabiOracleEnv(n_games, threads=threads, **game_params("Hanabi-Full", 2))
    e.set_decks(make_decks(5, 5, 4 * n_games, SEED))
    e.reset()
    rules = PRESETS["piers"]
    t = 0
    for _ in range(warm):
        a, _ = e.rules_act(rules, SEED, t)
        e.step(a)
        t += 1
    t0 = time.perf_counter()
    done = 0
    while True:
        a, _ = e.rules_act(rules, SEED, t)
        e.step(a)
        t += 1
        done += 1
        el = time.perf_counter() - t0
        if (steps_cap is not None and done >= steps_cap) or (steps_cap is None and el >= seconds):
            break
    return n_games * done / el, done, el



def bench_learner(env, obs, legal, reward, stype, action, T, world, args):
    """Learner updates/s (second half of the BASELINE metric) and acting rows/s for the Rainbow
    configuration of BASELINE configs[3]: C51, Hanabi-Full 2p, 512x512 noisy MLP, batch 256,
    prioritized replay shard filled from the self-play outputs of the timed region."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams
    n = env.n
    params = RlaxRainbowParams(train_batch_size=256, layers=[512, 512], experience_buffer_size=2 ** 19,
                               n_atoms=51, atom_vmax=25, seed=SEED)
    agent = DQNAgent(env.observation_spec(), env.action_spec(), params)
    exp = agent.experience
    # transitions between consecutive self-play steps of the same game
    fill = min(T - 1, (2 ** 19) // n)
    for t in range(fill):
        exp.add_transitions(obs[t], action[t + 1], reward[t + 1], obs[t + 1], legal[t + 1],
                            (stype[t + 1] == 2).to(torch.uint8))
    np.random.seed(SEED)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(5):
        agent.update()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k = args.learner_steps
    ev0.record()
    for _ in range(k):
        agent.update()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    tt = torch.tensor([ms], device=env.device)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms = float(tt.item())
    # forward 2.243 MFLOP/sample (folded noisy layers), update = online fwd on 2B rows +
    # target fwd on B rows + backward (2 fwd-equivalents on B rows) = 5 fwd-equivalents per sample
    flop_fwd = 2.0 * (658 * 512 + 512 * 512 + 512 * 1020)
    learner = {"updates_per_s": k / (ms * 1e-3), "ms_per_update": ms / k, "batch_per_gpu": 256,
               "global_batch": 256 * world, "steps": k, "replay_rows": exp.size,
               "net": "658-512-512-(20x51), noisy, C51 double-Q, Adam",
               "algorithmic_tflops": 5 * flop_fwd * 256 * world * k / (ms * 1e-3) / 1e12,
               "allreduce_bytes": agent.network.n_grad * 4 if world > 1 else 0,
               "loss": float(agent.last_loss)}
    if world > 1:
        learner.update(allreduce_selfcheck(agent, world, env.device))
    # acting: explore() over the n games of this GPU, packed rows resident
    o, l = obs[T - 1], legal[T - 1]
    for _ in range(3):
        agent.explore((None, (o, l)))
    barrier()
    ka = args.act_steps
    ev0.record()
    for _ in range(ka):
        agent.explore((None, (o, l)))
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    tt = torch.tensor([ms], device=env.device)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms = float(tt.item())
    tf = flop_fwd * n * ka / (ms * 1e-3) / 1e12
    peak_tf = 1409.6
    try:
        peak_tf = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"])
    except Exception:
        pass
    acting = {"rows_per_s": world * n * ka / (ms * 1e-3), "ms_per_call": ms / ka, "rows_per_gpu": n,
              "roofline": {"bound": "tensor", "kernel": "gemm_f16_kernel (3 layers) + unpack + policy",
                           "achieved": tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": tf / peak_tf,
                           "peak_source": "measured sustained cuBLAS bf16 (MEASURED_PEAKS.json)",
                           "traffic": None}}
    return learner, acting


def allreduce_selfcheck(agent, world, device):
    """Which collective the learner's update() ran, and proof that it sums: the NVLink kernel
    (csrc/comm.cu) and torch.distributed.all_reduce (NCCL) on copies of ONE rank-dependent bucket
    of the gradient's size.  bit-equal at world 2 (one addition per element, same on both); for more
    ranks the two may add in different orders: every element must agree within 1 ulp of the sum of the
    absolute values, and the NVLink result must be identical on every rank (fixed reduction order)."""
    import torch
    import torch.distributed as dist
    rank = dist.get_rank()
    impl = getattr(agent, "allreduce_impl", None)
    out = {"allreduce_impl": impl, "allreduce_fallback_reason": getattr(agent, "allreduce_fallback_reason", None)}
    n = agent.network.grad_bucket.numel()
    g = torch.Generator(device=device)
    g.manual_seed(1234 + rank)
    src = torch.randn(n, generator=g, device=device, dtype=torch.float32)
    ref = src.clone()
    dist.all_reduce(ref)
    mag = src.abs()
    dist.all_reduce(mag)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(fn, reps=20):
        for _ in range(3):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps * 1e3], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    scratch = src.clone()
    out["nccl_allreduce_us"] = timed(lambda: dist.all_reduce(scratch))
    if getattr(agent, "_comm", None):
        comm = agent._comm
        bucket = agent.network.grad_bucket
        keep = bucket.clone()
        bucket.copy_(src)
        comm()
        got = bucket.clone()
        # identical on every rank?
        lo, hi = got.clone(), got.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same_everywhere = bool(torch.equal(lo, hi))
        # the exact sum (float64 all-reduce of the same data) and the forward error bound of ANY order of
        # W float32 additions: |fl(sum) - sum| <= (W - 1) * eps/2 * sum|x_i| (+ the final rounding)
        exact = src.double()
        dist.all_reduce(exact)
        eps = torch.finfo(torch.float32).eps
        bound = (0.5 * (world - 1) * eps * mag.double() + 0.5 * eps * exact.abs()) * 1.0000001
        err_ours = (got.double() - exact).abs()
        err_nccl = (ref.double() - exact).abs()
        within = bool((err_ours <= bound).all().item())
        unit = (eps * mag.double()).clamp_min(1e-300)
        if world == 2:
            ok = within and bool(torch.equal(got, ref))
            how = "bit-equal to NCCL and within the rounding bound of the exact sum"
        else:
            ok = within
            how = "within the rounding bound (W-1)/2 eps sum|x| of the exact (float64) sum"
        out["allreduce_check"] = "ok" if (ok and same_everywhere) else "FAILED"
        out["allreduce_check_detail"] = (
            "%s: %s; identical on every rank (fixed reduction order): %s; max error in units of eps*sum|x|: "
            "ours %.3f, NCCL %.3f (bound %.1f); max |ours - NCCL| %.3g" % (
                how, ok, same_everywhere, float((err_ours / unit).max().item()),
                float((err_nccl / unit).max().item()), 0.5 * (world - 1) + 0.5,
                float((got - ref).abs().max().item())))
        out["allreduce_us"] = timed(lambda: comm())
        bucket.copy_(keep)
    else:
        out["allreduce_check"] = "nccl (library collective; nothing of ours to check)"
        out["allreduce_us"] = out["nccl_allreduce_us"]
    out["allreduce_payload_bytes"] = n * 4
    return out


def cpu_learner_baseline(seconds):
    """CPU baseline of the learner step ('PyTorch-CPU restatement, all cores; JAX CPU
    learner unavailable on this box'): oracle/learner.py OracleQNet, same net / batch as `learner`."""
    import numpy as np
    import torch
    from hanabi_agents_b200.rlax_dqn.network import init_params
    from oracle.learner import OracleQNet
    rng = np.random.default_rng(SEED)
    B, L, A = 256, 658, 20
    net = OracleQNet(init_params(L, [512, 512], A, 51, noisy=True, seed=SEED), A, 51, 25.0, noisy=True,
                     mirror_fp16=False)
    o1 = torch.tensor((rng.random((B, L)) < 0.15).astype(np.float32))
    o2 = torch.tensor((rng.random((B, L)) < 0.15).astype(np.float32))
    a = torch.tensor(rng.integers(0, A, B))
    r = torch.tensor(rng.integers(0, 2, B).astype(np.float32))
    term = torch.zeros(B, dtype=torch.bool)
    pr = torch.tensor(rng.random(B).astype(np.float32) + 0.1)
    done, t0 = 0, time.perf_counter()
    while True:
        _, _, g = net.grads(o1, a, r, o2, term, prios=pr, beta=0.4, discount=0.99)
        net.adam(g, lr=6.25e-5)
        done += 1
        el = time.perf_counter() - t0
        if el >= seconds or done >= 200:
            break
    return {"updates_per_s": done / el, "ms_per_update": 1e3 * el / done, "kind": "port",
            "threads": torch.get_num_threads(),
            "sample": "%d updates of batch 256 in %.1f s (PyTorch-CPU fp32 restatement of update_q + adam; "
                      "the reference's JAX learner is not installable here)" % (done, el)}


def bench_other_shards(world, rank, args):
    """The fused self-play kernel on the per-GPU shards of the two multi-GPU configurations of
    BASELINE.json (informational; the headline stays configs[1]): configs[3] = 1,048,576 Hanabi-Full
    2p games over 8 GPUs (131,072 per GPU), configs[4] = 262,144 Hanabi-Full 5p games (32,768 per GPU)."""
    import torch
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.rule_based import RulebasedAgent, predefined_rules
    out = []
    for players, n, obs_bytes, state_bytes in ((2, 131072, 84, 64), (5, 32768, 160, 128)):
        env = HanabiParallelEnv(n, "Hanabi-Full", players, game_id_base=rank * n, n_total=world * n,
                                n_decks=2 * n, seed=SEED)
        agent = RulebasedAgent(predefined_rules.piers_rules, seed=SEED)
        env.reset()
        T = args.fuse
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
        per_step = obs_bytes + 4 * env.legal_words + 4 + 1 + 4
        gbs = n * (T * per_step + 2 * state_bytes) / (ms * 1e-3) / 1e9
        out.append({"workload": "rule_based(piers) self-play, Hanabi-Full %dp, %d games/GPU" % (players, n),
                    "ms_per_launch": ms, "env_steps_per_s_per_gpu": n * T / (ms * 1e-3),
                    "algorithmic_GBps": gbs, "illegal_moves": env.illegal_count()})
        del env, bufs
        torch.cuda.empty_cache()
    return out


def bench_session(n, world, rank, args):
    """DQN self-play through HanabiParallelSession at the Rainbow configuration
    of BASELINE configs[3]: two learning agents take turns over n games of this GPU; a training
    tick = per-seat accounting + replay insert (both seats) + explore + env step; then the same
    with one learner update per agent and tick (gradient all-reduce across ranks when world > 1)."""
    import torch
    import torch.distributed as dist
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams
    from hanabi_agents_b200.session import HanabiParallelSession
    env = HanabiParallelEnv(n, "Hanabi-Full", 2, n_decks=4 * n * world, seed=SEED,
                            game_id_base=rank * n, n_total=n * world)
    params = RlaxRainbowParams(train_batch_size=256, layers=[512, 512], experience_buffer_size=2 ** 20,
                               n_atoms=51, atom_vmax=25, seed=SEED)
    agents = [DQNAgent(env.observation_spec(), env.action_spec(), params),
              DQNAgent(env.observation_spec(), env.action_spec(), params._replace(seed=SEED + 1))]
    ses = HanabiParallelSession(env, agents)
    ses.reset()
    ses.run(24, train=True)                       # fills both shards past the batch size
    for _ in range(8):                            # past the eager steps and the graph captures of every staging slot
        for a in agents:
            a.update()

    def timed(fn, reps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device=env.device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def learn_tick():
        ses.step(train=True)
        ses.update_learners()        # one update per agent (side by side in a single-process job)

    k = args.session_ticks
    ev = timed(lambda: ses.step(train=False), k)
    tr = timed(lambda: ses.step(train=True), k)
    lt = timed(learn_tick, k)
    st = ses.stats()
    return {"games_per_gpu": n, "ticks": k, "agents": "2 x DQNAgent 658-512-512-(20x51) noisy C51, replay 2^20 rows each",
            "eval_tick_ms": ev, "train_tick_ms": tr, "train_tick_plus_update_ms": lt,
            "env_steps_per_s_eval": world * n / (ev * 1e-3), "env_steps_per_s_train": world * n / (tr * 1e-3),
            "env_steps_per_s_train_plus_update": world * n / (lt * 1e-3),
            "episodes_closed": st["episodes"], "mean_episode_length": st["mean_length"],
            "illegal_moves": env.illegal_count()}


def bench_sum_tree(args):
    """BASELINE configs[2]: capacity 2^22, batch 256 stratified sample + priority update
    (SURVEY 8d cfg 3), device path; CPU side = the reference's own C++ tree (oracle/_ref)."""
    import numpy as np
    import torch
    from hanabi_agents_b200.sum_tree import SumTreef
    cap, B = 2 ** 22, 256
    rng = np.random.default_rng(0)
    tree = SumTreef(cap)
    vals = rng.random(cap).astype(np.float32)
    dv = torch.as_tensor(vals, device="cuda")
    from hanabi_agents_b200 import _lib
    for c0 in range(0, cap, 65536):
        _lib.check(_lib.lib().hb_st_update_range(tree._h, c0, 65536, dv[c0:c0 + 65536].data_ptr(), 0.0,
                                                 _lib.current_stream_ptr()))
    iters = args.tree_iters
    keys = torch.as_tensor((np.linspace(1 / B, 1, B)[None, :] - rng.uniform(0, 1 / B, (iters, B))).astype(np.float32),
                           device="cuda")
    newv = torch.as_tensor(rng.random((iters, B)).astype(np.float32), device="cuda")

    bufs = (torch.empty(B, dtype=torch.int64, device="cuda"), torch.empty(B, dtype=torch.float32, device="cuda"),
            torch.empty(1, dtype=torch.float32, device="cuda"))

    def it(i):
        idx, _, _ = tree.sample(keys[i], out=bufs)
        tree.update_values(idx, newv[i])

    for i in range(10):
        it(i)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(iters):
        it(i)
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    out = {"capacity": cap, "batch": B, "iters": iters, "iters_per_s": iters / (ms * 1e-3),
           "us_per_iter": 1e3 * ms / iters,
           "algorithmic_GBps": 70656 * iters / (ms * 1e-3) / 1e9,
           "note": "latency-bound: 22 dependent levels; 2 launches per iteration"}
    # The loop above is paced by the HOST (two Python / ctypes calls per iteration cost more than the two
    # kernels): the same iterations captured once into a CUDA graph and replayed give the device-side
    # figure -- what a caller inside a captured update (the learner) sees.
    try:
        per = min(iters, 256)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for i in range(3):
                it(i)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                for i in range(per):
                    it(i)
            for _ in range(2):
                g.replay()
            side.synchronize()
            reps = max(1, iters // per)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(side)
            for _ in range(reps):
                g.replay()
            e1.record(side)
            side.synchronize()
        gms = e0.elapsed_time(e1)
        out["graph_replay"] = {"iters": reps * per, "us_per_iter": 1e3 * gms / (reps * per),
                               "iters_per_s": reps * per / (gms * 1e-3),
                               "algorithmic_GBps": 70656 * reps * per / (gms * 1e-3) / 1e9,
                               "note": "%d iterations (sample + update each) captured into one CUDA graph, replayed %d "
                                       "times: no host pacing" % (per, reps)}
        torch.cuda.current_stream().wait_stream(side)
    except Exception as e:  # noqa: BLE001
        out["graph_replay"] = {"unavailable": repr(e)[:200]}
    # throughput variant (SURVEY 8d cfg 3): 1,024 independent batches of 256 keys in flight in ONE
    # sample launch (descent + leaf gather + total: 4*22 + 4 = 92 algorithmic bytes per key)
    nb = 1024
    big = torch.as_tensor((np.linspace(1 / B, 1, B)[None, :] - rng.uniform(0, 1 / B, (nb, B))).astype(np.float32)
                          .reshape(-1), device="cuda")
    bb = (torch.empty(nb * B, dtype=torch.int64, device="cuda"), torch.empty(nb * B, dtype=torch.float32, device="cuda"),
          torch.empty(1, dtype=torch.float32, device="cuda"))
    for _ in range(3):
        tree.sample(big, out=bb)
    torch.cuda.synchronize()
    reps = 20
    ev0.record()
    for _ in range(reps):
        tree.sample(big, out=bb)
    ev1.record()
    torch.cuda.synchronize()
    tms = ev0.elapsed_time(ev1) / reps
    out["throughput_variant"] = {"batches_in_flight": nb, "keys_per_launch": nb * B, "ms_per_launch": tms,
                                 "keys_per_s": nb * B / (tms * 1e-3),
                                 "algorithmic_GBps": 92.0 * nb * B / (tms * 1e-3) / 1e9,
                                 "note": "sampling only: ordered updates of one tree are one dependent chain"}
    # CPU: the reference's own pybind module, one thread (parity mode) -- bounded sample
    try:
        from oracle import ref_pybind_module
        ref_mod = ref_pybind_module()
        if ref_mod is None:
            raise RuntimeError("oracle/_ref/sum_tree*.so was not built")
        t0 = time.perf_counter()
        rt = ref_mod.SumTreef(cap)
        for c0 in range(0, cap, 65536):
            rt.update_values(list(range(c0, c0 + 65536)), vals[c0:c0 + 65536].tolist())
        build_s = time.perf_counter() - t0
        kk, vv = keys.cpu().numpy(), newv.cpu().numpy()
        n_cpu = min(iters, 2000)
        t0 = time.perf_counter()
        for i in range(n_cpu):
            ix = rt.get_indices(kk[i])
            rt.get_values(ix)
            rt.update_values(ix, vv[i])
        el = time.perf_counter() - t0
        out["cpu_reference"] = {"iters_per_s": n_cpu / el, "us_per_iter": 1e6 * el / n_cpu, "kind": "reference",
                                "threads": int(os.environ.get("OMP_NUM_THREADS", os.cpu_count() or 1)),
                                "build_and_fill_s": build_s,
                                "sample": "%d iterations through the reference's pybind module" % n_cpu}
    except Exception as e:  # the prebuilt module may be missing on a box without oracle/_ref
        out["cpu_reference"] = {"unavailable": repr(e)[:200]}
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    # same step as our arm: one bench step = args.fuse env steps of each of the 65,536 games; the
    # sample is bounded to 4096 env steps per game (a few tens of seconds on 8-16 host threads)
    env_steps = min(args.steps * args.fuse, 4096)
    warm = min(args.warmup * args.fuse, 256)
    val, done, el = cpu_port_throughput(GAMES_PER_GPU, None, threads, steps_cap=env_steps, warm=warm)
    bench_steps = done / float(args.fuse)
    sample = ("%d env steps of all %d games = %.1f bench steps of %d env steps (rule agent + step + int8 "
              "[N,658] encoding), %d threads" % (done, GAMES_PER_GPU, bench_steps, args.fuse, threads))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * el / bench_steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "host": "cpu", "games_per_gpu": GAMES_PER_GPU,
                   "fused_steps_per_launch": args.fuse,
                   "step": "one bench step = %d env steps of each of the %d games" % (args.fuse, GAMES_PER_GPU),
                   "env_steps_timed_per_game": done},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                         "note": "reference engine is third-party HLE (absent); oracle/ restatement timed"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from hanabi_agents_b200.env import HanabiParallelEnv
    from hanabi_agents_b200.rule_based import RulebasedAgent, predefined_rules

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = args.games
    env = HanabiParallelEnv(n, "Hanabi-Full", 2, game_id_base=rank * n, n_total=world * n,
                            n_decks=4 * n, seed=SEED)
    agent = RulebasedAgent(predefined_rules.piers_rules, seed=SEED)
    env.reset()
    T = args.fuse
    dev = env.device
    obs = torch.empty((T, n, env.obs_stride), dtype=torch.int32, device=dev)
    legal = torch.empty((T, n, env.legal_words), dtype=torch.int32, device=dev)
    reward = torch.empty((T, n), dtype=torch.float32, device=dev)
    stype = torch.empty((T, n), dtype=torch.uint8, device=dev)
    action = torch.empty((T, n), dtype=torch.int32, device=dev)

    def launch():
        env.selfplay(agent, T, obs=obs, legal=legal, reward=reward, step_type=stype, action=action)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # One bench step = ONE fused launch = T env steps of every game (config.fused_steps_per_launch).
    # The launches of the timed region are queued behind a stream gate (hb_gate_*: a one-thread kernel
    # polling a host flag) before the first event is recorded, so the region holds device time only:
    # at --steps 20 the host's launch latency used to be more than half of it.
    from hanabi_agents_b200._lib import StreamGate
    gate = StreamGate(local)
    launches = args.steps
    queued = min(launches, 96)           # the launch queue holds ~1000 entries; 96 x 0.17 ms is ample lead
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        clocks.mark_start()
        # warm-up: W launches, and at least ~0.3 s of the same kernel so that the clocks the sampler
        # sees (20 ms period) are the clocks of the timed region
        t_w = time.perf_counter()
        done_w = 0
        while done_w < max(3, args.warmup) or time.perf_counter() - t_w < 0.3:
            launch()
            done_w += 1
            if done_w % 64 == 0:
                torch.cuda.synchronize()
        barrier()
        if not args.no_gate:
            gate.close()
        ev0.record()
        for _ in range(queued):
            launch()
        gate.open()
        for _ in range(launches - queued):
            launch()
        ev1.record()
        barrier()
        clocks.mark_end()
    ms = ev0.elapsed_time(ev1)
    # an isolated launch (nothing before or after it on the device): what ncu's serialised
    # gpu__time_duration measures; consecutive launches of the timed region overlap (PDL hand-over)
    iso = []
    eva, evb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(5):
        torch.cuda.synchronize()
        if not args.no_gate:
            gate.close()
        eva.record()
        launch()
        evb.record()
        gate.open()
        torch.cuda.synchronize()
        iso.append(eva.elapsed_time(evb))
    iso_ms = sorted(iso)[len(iso) // 2]
    t = torch.tensor([ms, iso_ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, iso_ms = float(t[0].item()), float(t[1].item())
    env_steps = launches * T
    value = world * n * env_steps / (ms * 1e-3)

    # ---- e2e: same workload through the host-buffer API: actions come from pinned host memory
    # each step, observation / legal / reward / step type are read back to the host each step.
    e_steps = min(env_steps, args.e2e_steps)
    sets = [env.host_buffers(), env.host_buffers()]              # pinned host arrays, two sets
    acts = [sets[0]["actions"], sets[0]["next_actions"], sets[1]["actions"]]
    acts[0].copy_(agent.explore(env).cpu())
    torch.cuda.synchronize()
    e2e_t = [0]

    def e2e_step():
        # observations = env.step(actions); actions = agent.explore(observations), host arrays both ways,
        # double-buffered (include/hanabi_b200.h: hb_env_step_host_begin): the next step's actions go up
        # while this step's 6 MB of outputs are still coming down; the host waits for the actions only
        t = e2e_t[0]
        cur, nxt, hs = acts[t % 3], acts[(t + 1) % 3], sets[t & 1]
        env.step_host_begin(cur)
        agent.explore_host_async(env, nxt)
        env.step_host_outputs(hs["obs"], hs["legal"], hs["reward"], hs["step_type"])
        env.host_wait_actions()
        e2e_t[0] = t + 1

    for _ in range(3):
        e2e_step()
    env.host_wait()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e_steps):
        e2e_step()
    env.host_wait()                                                # every output of every step is on the host
    barrier()
    e_s = time.perf_counter() - t0
    te = torch.tensor([e_s], device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_val = world * n * e_steps / float(te.item())
    h2d = 4 * n
    d2h = n * (4 * env.obs_words + 4 * env.legal_words + 4 + 1) + 4 * n
    # The ceiling of this leg: the same bytes per env step, nothing else -- every rank copies d2h bytes from
    # the device into its pinned host buffer and h2d bytes back, back to back on one stream, all ranks at the
    # same time (they share the host's memory system and PCIe root complexes).  e2e / ceiling says how much of
    # what the host side allows the pipelined loop reaches; a ceiling that does not scale with the ranks is
    # the host's, not the kernels'.
    blob_d = torch.empty(d2h, dtype=torch.uint8, device=dev)
    blob_h = torch.empty(d2h, dtype=torch.uint8).pin_memory()
    act_d = torch.empty(h2d, dtype=torch.uint8, device=dev)
    act_h = torch.empty(h2d, dtype=torch.uint8).pin_memory()
    s_down, s_up = torch.cuda.Stream(), torch.cuda.Stream()      # the two directions use separate copy engines

    def copies():
        with torch.cuda.stream(s_down):
            blob_h.copy_(blob_d, non_blocking=True)
        with torch.cuda.stream(s_up):
            act_d.copy_(act_h, non_blocking=True)

    for _ in range(5):
        copies()
    barrier()
    reps_c = 200
    t0 = time.perf_counter()
    for _ in range(reps_c):
        copies()
    torch.cuda.synchronize()
    c_s = time.perf_counter() - t0
    barrier()
    tc = torch.tensor([c_s], device=dev)
    if world > 1:
        dist.all_reduce(tc, op=dist.ReduceOp.MAX)
    ceiling_val = world * n * reps_c / float(tc.item())
    ceiling_gbs = world * (d2h + h2d) * reps_c / float(tc.item()) / 1e9
    cpu_aff = None
    try:
        cpu_aff = len(os.sched_getaffinity(0))
    except Exception:
        pass

    learner = acting = None
    if not args.no_learner:
        learner, acting = bench_learner(env, obs, legal, reward, stype, action, T, world, args)
    session = None if args.no_session else bench_session(n, world, rank, args)
    other = None if args.no_session else bench_other_shards(world, rank, args)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    tree = None if args.no_tree else bench_sum_tree(args)
    if learner is not None and args.cpu_seconds > 0:
        try:
            learner["cpu_baseline"] = cpu_learner_baseline(min(args.cpu_seconds, 10.0))
        except Exception as e:
            learner["cpu_baseline"] = {"unavailable": repr(e)[:200]}

    # ---- roofline of the dominant kernel (selfplay_kernel), live from the timed region
    peak, which = peaks()
    # algorithmic bytes per env step: packed obs row 84 B (658 bits) + legal 4 +
    # reward 4 + step type 1 + action 4 = 97 B, plus the game state (64 B in + 64 B out) once
    # per launch, i.e. 128/T per step.
    per_step = 84 + 4 + 4 + 1 + 4
    bytes_per_launch = n * (T * per_step + 2 * env.state_words * 4)
    launch_ms = ms / launches                      # the launches actually timed
    achieved = bytes_per_launch / (launch_ms * 1e-3) / 1e9
    prof = {}
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            prof = json.load(open(tp)).get("selfplay_kernel") or {}
            if not isinstance(prof, dict):
                prof = {"dram_bytes_per_launch": prof}
        except Exception:
            prof = {}
    same_shape = prof.get("games") == n and prof.get("fused_steps") == T
    roof = {"bound": "hbm", "kernel": "selfplay_kernel", "achieved": achieved, "peak": peak,
            "peak_source": which + " copy bandwidth (MEASURED_PEAKS.json)", "unit": "GB/s",
            "frac": achieved / peak, "peak_nominal": 8000.0, "frac_nominal": achieved / 8000.0,
            # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch from an ncu --set full capture
            # of this kernel at this shape (profiles/traffic.json names the capture); null otherwise
            "traffic": prof.get("dram_bytes_per_launch") if same_shape else None,
            "traffic_source": prof.get("source") if same_shape else None,
            "algorithmic_bytes_per_launch": bytes_per_launch,
            "launches_timed": launches, "launch_ms": launch_ms,
            "isolated_launch_ms": iso_ms,
            "isolated_achieved": bytes_per_launch / (iso_ms * 1e-3) / 1e9,
            "isolated_frac": bytes_per_launch / (iso_ms * 1e-3) / 1e9 / peak,
            "note": "launch_ms = timed region / launches (consecutive launches overlap on the device); "
                    "isolated_* = one launch alone between two events, which is what ncu's serialised "
                    "gpu__time_duration sees"}
    # The kernel's binding resource is not HBM but the ALU pipe (LOP3 / SHF / ISETP / SEL: one warp
    # instruction per TWO cycles per SM sub-partition).  alu_issue = ALU-pipe warp instructions of one
    # launch (ncu sm__inst_executed_pipe_alu.sum of the capture named in profiles/traffic.json) / live
    # launch time, against 148 SMs x 4 sub-partitions x 0.5 / cycle at the SM clock sampled in the run.
    roof_alu = None
    if same_shape and prof.get("alu_warp_inst_per_launch"):
        clk = (clocks.summary().get("sm_mhz") or 1965.0) * 1e6
        n_sm = torch.cuda.get_device_properties(local).multi_processor_count
        alu_peak = n_sm * 4 * 0.5 * clk / 1e9
        alu = float(prof["alu_warp_inst_per_launch"]) / (launch_ms * 1e-3) / 1e9
        roof_alu = {"bound": "alu_issue", "kernel": "selfplay_kernel", "achieved": alu, "peak": alu_peak,
                    "unit": "G warp-inst/s", "frac": alu / alu_peak,
                    "alu_warp_inst_per_launch": prof["alu_warp_inst_per_launch"],
                    "alu_warp_inst_per_warp_step": float(prof["alu_warp_inst_per_launch"]) / (n / 32 * T),
                    "isolated_frac": float(prof["alu_warp_inst_per_launch"]) / (iso_ms * 1e-3) / 1e9 / alu_peak,
                    "source": prof.get("source"),
                    "peak_note": "%d SMs x 4 sub-partitions x 1 ALU warp-instruction / 2 cycles x %.0f MHz"
                                 % (n_sm, clk / 1e6)}

    cores = os.cpu_count() or 1
    cval, cdone, cel = cpu_port_throughput(n, args.cpu_seconds, cores)
    cpu = {"value": cval, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": "%d env steps of all %d games in %.1f s (rule agent + step + int8 encoding, "
                     "OpenMP over games)" % (cdone, n, cel),
           "note": "reference engine is third-party HLE (absent here); oracle/ restatement timed"}
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / launches, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32-bitfield", "data": "synthetic",
        "config": {"workload": WORKLOAD, "games_per_gpu": n, "fused_steps_per_launch": T,
                   "step": "one bench step = one fused launch = %d env steps of each of the %d games per GPU "
                           "(%d env-steps per GPU); the reference arm's step is the same %d env steps per game"
                           % (T, n, T * n, T),
                   "env_steps_timed_per_game": env_steps,
                   "timing": "CUDA events on the launching stream; the first %d launches are queued behind a "
                             "stream gate before the first event, so no host launch latency is inside" % queued,
                   "warmup_launches": done_w,
                   "l2": "outputs of one launch (%.0f MB) exceed the 126 MB L2; no flush needed"
                         % (T * n * (env.obs_stride * 4 + env.legal_words * 4 + 9) / 1e6),
                   "decks": "numpy default_rng(%d) permutations, pool of %d" % (SEED, 4 * n),
                   "launch_overlap": "consecutive self-play launches of one stream overlap (programmatic dependent "
                                     "launch + per-group FIFO hand-over); roofline.launch_ms = timed region / launches"},
        "clocks": clocks.summary(),
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d * T, "d2h_bytes_per_step": d2h * T,
                "h2d_bytes_per_env_step": h2d, "d2h_bytes_per_env_step": d2h,
                "env_steps_timed": e_steps, "ms_per_env_step": 1e3 * float(te.item()) / e_steps,
                "ceiling": {"value": ceiling_val, "unit": UNIT, "frac": e2e_val / ceiling_val,
                            "copy_GBps_all_ranks": ceiling_gbs, "host_cpus_per_rank": cpu_aff,
                            "how": "the same D2H and H2D bytes per env step copied back to back (one stream per "
                                   "direction) between the device and pinned host memory by every rank at once, "
                                   "nothing else running (max over ranks)"},
                "path": "per step, host in the loop, pinned host arrays both ways, double-buffered: env.step_host_begin "
                "(actions H2D, env_step_kernel) | agent.explore_host_async (rules_act_kernel + next actions D2H) | "
                "env.step_host_outputs (D2H of observation rows (%d B) + legal mask + reward + step type in one copy, "
                "alternating two host sets) | env.host_wait_actions; env.host_wait at the end of the timed region"
                % (4 * env.obs_words)},
        "gpu_launches": launches,
        "roofline": roof,
        "roofline_alu": roof_alu,
        "cpu_baseline": cpu,
        "learner": learner,
        "acting": acting,
        "sum_tree": tree,
        "session": session,
        "other_shards": other,
    }
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1024,
                    help="bench steps; one step = one fused launch of --fuse env steps of every game")
    ap.add_argument("--warmup", type=int, default=16)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--games", type=int, default=GAMES_PER_GPU)
    ap.add_argument("--fuse", type=int, default=64, help="env steps fused per launch")
    ap.add_argument("--e2e-steps", type=int, default=512, help="env steps of the host-in-the-loop leg")
    ap.add_argument("--e2e-chunks", type=int, default=1)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--learner-steps", type=int, default=300)
    ap.add_argument("--act-steps", type=int, default=20)
    ap.add_argument("--tree-iters", type=int, default=3000)
    ap.add_argument("--no-learner", action="store_true")
    ap.add_argument("--no-tree", action="store_true")
    ap.add_argument("--no-session", action="store_true")
    ap.add_argument("--no-gate", action="store_true",
                    help="do not hold the stream behind a gate (for runs under ncu, which executes every launch "
                         "synchronously: a gate kernel would wait for a host that is blocked in its launch)")
    ap.add_argument("--session-ticks", type=int, default=100)
    args = ap.parse_args()
    # Only the JSON line may reach stdout: libraries print there too (NCCL announces its version on
    # stdout when NCCL_DEBUG is set), so file descriptor 1 points at stderr while the benchmark runs.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import io
    captured = io.StringIO()
    py_stdout, sys.stdout = sys.stdout, captured
    try:
        if args.impl == "reference":
            run_reference(args)
        else:
            run_ours(args)
    finally:
        sys.stdout = py_stdout
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    lines = [ln for ln in captured.getvalue().splitlines() if ln.strip()]
    for ln in lines[:-1]:
        sys.stderr.write(ln + "\n")
    if lines:
        print(lines[-1], flush=True)


if __name__ == "__main__":
    main()
