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


@pytest.mark.parametrize("algo", ["push_all", "owner"])
@pytest.mark.parametrize("W,nparts", [(1, 5), (2, 147), (4, 16), (8, 3), (8, 147), (3, 20)])
def test_loopback_sum_order_and_identical_replicas(W, nparts, algo):
    """G = gshared + g_0 + g_1 + ... in rank order (model.py:53-55 with arrival order := rank order),
    buffers start at ones (model.py:51), every replica ends bit-identical; five epochs so that both
    receive-slot parities are reused."""
    import b200_native as nv
    P = 71720
    g = torch.Generator().manual_seed(W * 100 + nparts)
    theta0 = torch.randn(P, generator=g)
    comms = nv.comm_create_loopback(W, P)
    for c in comms:                                    # both exchange algorithms must give the same sum in the same order
        nv.comm_set_algorithm(c, nv.COMM_ALGO_PUSH_ALL if algo == "push_all" else nv.COMM_ALGO_OWNER)
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


def test_owner_and_push_all_algorithms_are_bitwise_equal():
    """The OWNER algorithm (slice to one owner, sum back to all) and PUSH_ALL (everything to everyone) add the same
    values in the same order: identical parameters and Adam state after five epochs at W = 8."""
    import b200_native as nv
    P, W, nparts = 71720, 8, 37
    res = []
    for algo in (nv.COMM_ALGO_PUSH_ALL, nv.COMM_ALGO_OWNER):
        g = torch.Generator().manual_seed(5)
        theta0 = torch.randn(P, generator=g)
        comms = nv.comm_create_loopback(W, P)
        for c in comms:
            nv.comm_set_algorithm(c, algo)
        th = [theta0.to(DEV).clone() for _ in range(W)]
        m = [torch.zeros(P, device=DEV) for _ in range(W)]
        v = [torch.zeros(P, device=DEV) for _ in range(W)]
        gsh = [torch.ones(P, device=DEV) for _ in range(W)]
        try:
            for step in range(1, 6):
                parts = [(torch.randn(nparts, P, generator=g) * 0.01).to(DEV) for _ in range(W)]
                nv.allreduce_adam_loopback(comms, parts, gsh, th, None, m, v, step, 3e-4)
            torch.cuda.synchronize()
        finally:
            for c in comms:
                nv.comm_destroy(c)
        res.append((th[0].clone(), m[3].clone(), v[7].clone()))
    for a, b in zip(*res):
        assert torch.equal(a, b)


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
@pytest.mark.parametrize("mode", ["p2p", "p2p-owner", "nccl"])
def test_two_gpus_match_single_process(mode):
    env = dict(os.environ, DPPO_COMM=mode.split("-")[0])
    if mode.endswith("owner"):
        env["DPPO_COMM_ALGO"] = "owner"
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29613",
                          os.path.join(ROOT, "tests", "multi_gpu_worker.py")], env=env, capture_output=True, text=True,
                         timeout=300)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "MULTI_GPU_OK" in out.stdout and "SHARED_STATS_OK" in out.stdout


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_gpus_lagging_rank_is_detected_not_waited_for_forever():
    """A rank that makes one call fewer: the other rank's poll runs out of its clock budget, the kernel returns and
    dppo_comm_check reports it (unbounded spin)."""
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29619",
                          os.path.join(ROOT, "tests", "multi_gpu_lag_worker.py")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "LAG_DETECTED" in out.stdout and "LAG_NOT_DETECTED" not in out.stdout


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_gpus_main_launcher():
    """pytorch-dppo_b200/main.py `run` under torchrun: train.train on every rank, exchange inside K3."""
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29617",
                          os.path.join(ROOT, "tests", "multi_gpu_main_worker.py")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "MAIN_LAUNCHER_OK" in out.stdout


# ---------------------------------------------------------------- statistics shared by the workers (model.py:61-80, main.py:54)
def _shared_stats_loopback(nv, raws, state0=None):
    """W workers' rollouts through dppo_welford_totals -> all-gather (loopback: W virtual ranks on this GPU) ->
    dppo_welford_shared, every rank on its own replica of the statistics.  Returns (normalised rollouts, states)."""
    W, O = len(raws), raws[0].shape[1]
    comms = nv.comm_create_loopback(W, 8)
    try:
        X = [torch.as_tensor(r, dtype=torch.float32).to(DEV).contiguous() for r in raws]
        st = [[(torch.zeros(O, device=DEV) if state0 is None else state0[i].to(DEV).clone()) for i in range(4)] for _ in range(W)]
        mine = [torch.zeros(1 + 2 * O, dtype=torch.float64, device=DEV) for _ in range(W)]
        every = [torch.zeros(W, 1 + 2 * O, dtype=torch.float64, device=DEV) for _ in range(W)]
        for r in range(W):
            T = X[r].shape[0]
            nv.welford_totals(X[r], st[r][1], mine[r], torch.empty(max(nv.welford_totals_scratch_bytes(T, O), 16),
                                                                    dtype=torch.uint8, device=DEV))
        nv.comm_allgather_loopback(comms, mine, every)
        torch.cuda.synchronize()
        for r in range(1, W):
            assert torch.equal(every[r], every[0])
        outs = []
        for r in range(W):
            T = X[r].shape[0]
            out = torch.empty_like(X[r])
            nv.welford_shared(X[r], out, st[r][0], st[r][1], st[r][2], st[r][3], every[r], r,
                              torch.empty(max(nv.welford_scratch_bytes(T, O), 16), dtype=torch.uint8, device=DEV))
            outs.append(out)
        torch.cuda.synchronize()
        return [o.cpu() for o in outs], st
    finally:
        for c in comms:
            nv.comm_destroy(c)


def test_shared_stats_two_workers_match_the_reference_run(golden_dir):
    """dppo_w2_o17a6 was recorded from the reference with ONE Shared_obs_stats updated by worker 0's rollout and
    then worker 1's: worker 1's states continue worker 0's statistics.  Rank order here gives the same."""
    import numpy as np
    import b200_native as nv
    g = np.load(os.path.join(golden_dir, "dppo_w2_o17a6.npz"))
    outs, st = _shared_stats_loopback(nv, [g["w0_raw_obs"], g["w1_raw_obs"]])
    for w in range(2):
        Tn = g["w%d_states" % w].shape[0]
        assert_close_scaled(outs[w][-Tn:], g["w%d_states" % w], 1e-5, "states of worker %d" % w)
    for i, k in enumerate(("n", "mean", "mean_diff", "var")):
        assert torch.equal(st[0][i], st[1][i]), k                                  # replicas end identical
        assert_close_scaled(st[1][i].cpu(), g["w1_stats_" + k], 1e-5, "final " + k)
    assert torch.equal(st[0][0].cpu(), torch.from_numpy(g["w1_stats_n"]))


@pytest.mark.parametrize("W,T,O", [(4, 1500, 17), (3, 700, 40), (8, 300, 376)])
def test_shared_stats_rank_order_vs_oracle(W, T, O):
    """The sequential oracle (model.py:68-80 row by row) over the W rollouts concatenated in rank order, after an
    earlier rollout has left a non-trivial state."""
    import b200_native as nv
    g = torch.Generator().manual_seed(W * 1000 + O)
    loc = torch.rand(O, generator=g) * 4 - 2
    scale = torch.rand(O, generator=g) * 3 + 0.05
    first = loc + scale * torch.randn(500, O, generator=g)
    raws = [loc + scale * torch.randn(T + 17 * r, O, generator=g) for r in range(W)]     # ragged rollout lengths
    ref = orc.welford_new(O)
    orc.welford_observe_normalize(ref, first)
    state0 = [ref[k].clone() for k in ("n", "mean", "mean_diff", "var")]
    want = [orc.welford_observe_normalize(ref, r) for r in raws]
    outs, st = _shared_stats_loopback(nv, raws, state0)
    for r in range(W):
        assert_close_scaled(outs[r], want[r], 1e-5, "rank %d" % r)
    for i, k in enumerate(("n", "mean", "mean_diff", "var")):
        for r in range(1, W):
            assert torch.equal(st[r][i], st[0][i])
        assert_close_scaled(st[0][i].cpu(), ref[k], 1e-5, k)


def test_shared_stats_world_one_equals_plain_welford():
    import b200_native as nv
    g = torch.Generator().manual_seed(2)
    raw = 0.3 + 1.7 * torch.randn(5000, 17, generator=g)
    outs, st = _shared_stats_loopback(nv, [raw])
    X = raw.to(DEV)
    out = torch.empty_like(X)
    s2 = [torch.zeros(17, device=DEV) for _ in range(4)]
    nv.welford(X, out, s2[0], s2[1], s2[2], s2[3], torch.empty(nv.welford_scratch_bytes(5000, 17), dtype=torch.uint8, device=DEV))
    torch.cuda.synchronize()
    assert torch.equal(outs[0], out.cpu())
    assert torch.equal(st[0][0], s2[0])
    for i in range(1, 4):
        assert_close_scaled(st[0][i].cpu(), s2[i].cpu(), 1e-6, "state %d" % i)
