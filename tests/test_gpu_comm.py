"""The fused gradient exchange + Adam kernel (dppo_allreduce_adam).  On a single GPU the W
ranks are emulated by ONE cooperative launch over all ranks' data (dppo_*_loopback); the real
one-process-per-GPU path is exercised by test_two_gpus_* when two devices are visible."""
import os
import subprocess
import sys

import pytest
import torch

from oracle import ppo_oracle as orc
from util import assert_close_scaled

pytestmark = pytest.mark.gpu
DEV = "cuda"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("W,nparts", [(1, 5), (2, 147), (4, 16), (8, 3)])
def test_loopback_sum_order_and_identical_replicas(W, nparts):
    """G = gshared + g_0 + g_1 + ... in rank order (model.py:53-55 with arrival order := rank order),
    buffers start at ones (model.py:51), every replica ends bit-identical; five epochs so that both
    receive-slot parities are reused."""
    import b200_native as nv
    P = 71720
    g = torch.Generator().manual_seed(W * 100 + nparts)
    theta0 = torch.randn(P, generator=g)
    comms = nv.comm_create_loopback(W, P)
    th = [theta0.to(DEV).clone() for _ in range(W)]
    m = [torch.zeros(P, device=DEV) for _ in range(W)]
    v = [torch.zeros(P, device=DEV) for _ in range(W)]
    prev = [torch.empty(P, device=DEV) for _ in range(W)]
    gout = [torch.empty(P, device=DEV) for _ in range(W)]
    gsh = [torch.ones(P, device=DEV) for _ in range(W)]
    rt, rm, rv = theta0.clone(), torch.zeros(P), torch.zeros(P)
    rg = torch.ones(P)
    try:
        for step in range(1, 6):
            parts = [torch.randn(nparts, P, generator=g) * 0.01 for _ in range(W)]
            nv.allreduce_adam_loopback(comms, [p.to(DEV) for p in parts], gsh, th, prev, m, v, step, 3e-4, grad_out=gout)
            torch.cuda.synchronize()
            G = rg.clone()
            for r in range(W):                         # rank order
                G += parts[r].sum(0)
            before = rt.clone()
            orc.adam_step(rt, G, rm, rv, step, 3e-4)
            rg.zero_()
            for r in range(W):
                assert torch.equal(th[r], th[0]) and torch.equal(m[r], m[0]) and torch.equal(v[r], v[0])
                assert torch.equal(prev[r].cpu(), before)
                assert float(gsh[r].abs().max()) == 0.0
            assert_close_scaled(gout[0].cpu(), G, 1e-5, "G step %d" % step)
            assert_close_scaled(th[0].cpu(), rt, 1e-5, "theta step %d" % step)
            rt.copy_(th[0].cpu()); rm.copy_(m[0].cpu()); rv.copy_(v[0].cpu())
    finally:
        for c in comms:
            nv.comm_destroy(c)


def test_world_one_is_bitwise_the_single_gpu_path():
    import b200_native as nv
    P, nparts = 71720, 147
    g = torch.Generator().manual_seed(3)
    theta0 = torch.randn(P, generator=g)
    parts = (torch.randn(nparts, P, generator=g) * 0.01).to(DEV)
    outs = []
    for use_comm in (False, True):
        th, m, v = theta0.to(DEV).clone(), torch.zeros(P, device=DEV), torch.zeros(P, device=DEV)
        gsh = torch.ones(P, device=DEV)
        if use_comm:
            c = nv.comm_create(0, 1, P)
            for step in (1, 2, 3):
                nv.allreduce_adam(c, parts, nparts, gsh, th, None, m, v, step, 3e-4)
            torch.cuda.synchronize()
            nv.comm_destroy(c)
        else:
            for step in (1, 2, 3):
                nv.reduce_adam(parts, nparts, gsh, th, None, m, v, step, 3e-4)
            torch.cuda.synchronize()
        outs.append((th.cpu(), m.cpu(), v.cpu()))
    for a, b in zip(*outs):
        assert torch.equal(a, b)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("mode", ["p2p", "nccl"])
def test_two_gpus_match_single_process(mode):
    env = dict(os.environ, DPPO_COMM=mode)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29613",
                          os.path.join(ROOT, "tests", "multi_gpu_worker.py")], env=env, capture_output=True, text=True,
                         timeout=300)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "MULTI_GPU_OK" in out.stdout


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_gpus_main_launcher():
    """pytorch-dppo_b200/main.py `run` under torchrun: train.train on every rank, exchange inside K3."""
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29617",
                          os.path.join(ROOT, "tests", "multi_gpu_main_worker.py")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "MAIN_LAUNCHER_OK" in out.stdout
