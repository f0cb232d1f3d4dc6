"""2+ GPUs (torchrun): the NVLink all-reduce kernel (csrc/comm.cu) against NCCL on the same data,
then DQNAgent.update with either collective -- parameters must stay bit-identical across ranks, and for
world size 2 (a + b is commutative) identical between the two collectives.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/comm_check.py"""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hanabi_agents_b200.dist import NvlinkAllReduce  # noqa: E402
from hanabi_agents_b200.rlax_dqn import DQNAgent, RlaxRainbowParams  # noqa: E402
from hanabi_agents_b200.rlax_dqn.network import QNetwork  # noqa: E402
from hanabi_agents_b200.env import ObservationSpec, ActionSpec  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)

net = QNetwork(658, 24, [512, 512], 20, 51, 25.0, True)
bucket = net.grad_bucket
comm = NvlinkAllReduce(bucket)
g = torch.Generator(device=dev).manual_seed(100 + rank)
for it in range(40):
    x = torch.randn(bucket.numel(), device=dev, generator=g) * (1.0 + it)
    bucket.copy_(x)
    ref = x.clone()
    dist.all_reduce(ref)
    comm()
    torch.cuda.synchronize()
    if world == 2:
        assert torch.equal(bucket, ref), (rank, it, float((bucket - ref).abs().max()))
    else:
        assert torch.allclose(bucket, ref, rtol=1e-5, atol=1e-5 * (1 + it)), (rank, it)
    # every rank holds the same bits
    chk = bucket.view(torch.int32).to(torch.int64).sum().reshape(1)
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert int(lo) == int(hi), (rank, it)


# ranges on independent channels, in flight together on two streams, captured into ONE graph and replayed:
# every replay must give the sum again (nothing of a call is frozen into the graph but pointers)
from hanabi_agents_b200 import _lib  # noqa: E402
L = _lib.lib()
n_all = bucket.numel()
cut = (n_all // 2) // 4 * 4
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
graph = torch.cuda.CUDAGraph()
x = torch.randn(n_all, device=dev, generator=g)
bucket.copy_(x)
torch.cuda.synchronize()
with torch.cuda.graph(graph):
    cur = torch.cuda.current_stream()
    s1.wait_stream(cur)
    s2.wait_stream(cur)
    _lib.check(L.hb_comm_allreduce_range(comm._h, 0, 0, cut, s1.cuda_stream))
    _lib.check(L.hb_comm_allreduce_range(comm._h, 1, cut, n_all - cut, s2.cuda_stream))
    cur.wait_stream(s1)
    cur.wait_stream(s2)
for it in range(5):
    x = torch.randn(n_all, device=dev, generator=g) * (1.0 + it)
    bucket.copy_(x)
    ref = x.clone()
    dist.all_reduce(ref)
    graph.replay()
    torch.cuda.synchronize()
    if world == 2:
        assert torch.equal(bucket, ref), ("graph replay", rank, it, float((bucket - ref).abs().max()))
    else:
        assert torch.allclose(bucket, ref, rtol=1e-5, atol=1e-5 * (1 + it)), ("graph replay", rank, it)
if rank == 0:
    print("two channels on two streams, captured and replayed 5 times: ok")


def timed(fn, reps=200):
    for _ in range(10):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


small = QNetwork(171, 8, [64], 11, 51, 25.0, False)
comm_small = NvlinkAllReduce(small.grad_bucket)
t_small = timed(lambda: comm_small())
t_small_nc = timed(lambda: dist.all_reduce(small.grad_bucket))
if rank == 0:
    print("all-reduce of %.3f MB: nvlink kernel %.1f us, nccl %.1f us (fixed cost)" % (
        small.grad_bucket.numel() * 4 / 1e6, t_small * 1e3, t_small_nc * 1e3))
t_nv = timed(lambda: comm())
t_nc = timed(lambda: dist.all_reduce(bucket))
if rank == 0:
    print("all-reduce of %.2f MB over %d GPUs: nvlink kernel %.1f us, nccl %.1f us" % (bucket.numel() * 4 / 1e6, world, t_nv * 1e3, t_nc * 1e3))

# agents: same seeds on every rank, different replay contents per rank
def make(force_nccl):
    params = RlaxRainbowParams(train_batch_size=64, layers=[256], experience_buffer_size=4096, seed=7)
    a = DQNAgent(ObservationSpec((128, 171)), ActionSpec(11), params)
    if force_nccl:
        a._comm = False
    rng = np.random.default_rng(50 + rank)
    n = 1024
    o1 = (rng.random((n, 171)) < 0.2).astype(np.int8)
    o2 = (rng.random((n, 171)) < 0.2).astype(np.int8)
    lm = (rng.random((n, 11)) < 0.6).astype(np.int8)
    a.experience.add_transitions(o1, rng.integers(0, 11, n), rng.integers(0, 2, n).astype(np.float32), o2, lm,
                                 np.zeros(n, np.uint8))
    return a


a_nv, a_nc = make(False), make(True)
for step in range(6):
    u = np.random.default_rng(step).uniform(0, 1 / 64, 64)
    a_nv.update(uniforms=u)
    a_nc.update(uniforms=u)
torch.cuda.synchronize()
assert a_nv._comm and a_nv._fused_dp and not a_nc._comm
for pa, pb in zip(a_nv.network.get_params(), a_nc.network.get_params()):
    for k in pa:
        if world == 2:
            assert np.array_equal(pa[k], pb[k]), k
        else:
            np.testing.assert_allclose(pa[k], pb[k], rtol=1e-4, atol=1e-6)
# identical parameters on every rank
flat = torch.as_tensor(np.concatenate([v.ravel() for l in a_nv.network.get_params() for v in l.values()]), device=dev)
lo, hi = flat.clone(), flat.clone()
dist.all_reduce(lo, op=dist.ReduceOp.MIN)
dist.all_reduce(hi, op=dist.ReduceOp.MAX)
assert torch.equal(lo, hi)

B = 256
params = RlaxRainbowParams(train_batch_size=B, layers=[512, 512], experience_buffer_size=2 ** 15, seed=1)
res = {}
for name, force in (("nvlink", False), ("nvlink-unfused", False), ("nccl", True)):
    os.environ["HB_ALLREDUCE"] = name if name != "nccl" else "nvlink"
    ag = DQNAgent(ObservationSpec((1024, 658)), ActionSpec(20), params)
    if force:
        ag._comm = False
    rng = np.random.default_rng(rank)
    n = 8192
    ag.experience.add_transitions((rng.random((n, 658)) < 0.15).astype(np.int8), rng.integers(0, 20, n),
                                  rng.integers(0, 2, n).astype(np.float32), (rng.random((n, 658)) < 0.15).astype(np.int8),
                                  (rng.random((n, 20)) < 0.6).astype(np.int8), np.zeros(n, np.uint8))
    res[name] = timed(lambda: ag.update(), reps=200)
if rank == 0:
    print("learner update, batch %d per GPU, %d GPUs: nvlink per layer inside the update %.1f us (%.0f updates/s), "
          "nvlink whole bucket between two graphs %.1f us (%.0f updates/s), nccl %.1f us (%.0f updates/s)" % (
              B, world, res["nvlink"] * 1e3, 1e3 / res["nvlink"], res["nvlink-unfused"] * 1e3,
              1e3 / res["nvlink-unfused"], res["nccl"] * 1e3, 1e3 / res["nccl"]))
    print("comm ok")
dist.destroy_process_group()
