"""GPU: the NVLink all-reduce of the gradient bucket (csrc/comm.cu).  On one GPU: the C-ABI life cycle
(create / export / world-1 no-op / argument checks).  With two or more GPUs: tools/comm_check.py under
torchrun -- bit-equality with NCCL at world size 2, identical bits on every rank, DQNAgent.update with
either collective."""
import ctypes as C
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_comm_single_gpu_lifecycle():
    import torch
    from hanabi_agents_b200 import _lib
    from hanabi_agents_b200.rlax_dqn.network import QNetwork
    L = _lib.lib()
    net = QNetwork(171, 8, [64], 11, 51, 25.0, False)
    b = net.grad_bucket
    b.fill_(1.5)
    h = C.c_void_p()
    _lib.check(L.hb_comm_create(0, 1, 0, b.data_ptr(), b.numel(), C.byref(h)))
    blob = (C.c_char * 128)()
    _lib.check(L.hb_comm_export(h, blob))
    assert any(x != b"\x00" for x in blob)
    _lib.check(L.hb_comm_allreduce(h, None))          # world 1: nothing to add
    torch.cuda.synchronize()
    assert float(b.min()) == float(b.max()) == 1.5
    assert (L.hb_comm_info(h, 0), L.hb_comm_info(h, 1)) == (0, 1)
    _lib.check(L.hb_comm_destroy(h))
    h2 = C.c_void_p()
    with pytest.raises(ValueError):
        _lib.check(L.hb_comm_create(2, 2, 0, b.data_ptr(), b.numel(), C.byref(h2)))
    with pytest.raises(ValueError):
        _lib.check(L.hb_comm_create(0, 2, 0, b.data_ptr() + 4, b.numel() - 4, C.byref(h2)))
    _lib.check(L.hb_comm_create(0, 2, 0, b.data_ptr(), b.numel(), C.byref(h2)))
    with pytest.raises(_lib.HanabiB200Error):           # peers not mapped yet
        _lib.check(L.hb_comm_allreduce(h2, None))
    _lib.check(L.hb_comm_destroy(h2))


def test_comm_two_gpus_against_nccl():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs on one node")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "tools", "comm_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0 and "comm ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
