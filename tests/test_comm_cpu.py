"""Host-side logic of the multi-GPU path on CPU: two gloo ranks exchange their handles in rank
order (the device buffers themselves need GPUs: tests/test_gpu_comm.py)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    for p in (ROOT, os.path.join(ROOT, "pytorch-dppo_b200")):
        sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import b200_comm
    comm = b200_comm.PeerComm.create()
    got = b200_comm.exchange(comm.group, b"handle-of-rank-%d" % rank)
    # the gradient sum itself: rank order, identical on every rank (what K3 does over NVLink)
    g = torch.full((8,), float(rank + 1))
    parts = [torch.zeros(8) for _ in range(world)]
    dist.all_gather(parts, g)
    total = torch.ones(8)                              # model.py:51
    for r in range(world):
        total += parts[r]
    q.put((rank, comm.rank, comm.world, comm.mode, got, total.tolist(), comm.describe()["mode"]))
    dist.destroy_process_group()


def test_two_gloo_ranks_exchange_in_rank_order():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, 29651, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, crank, cworld, mode, got, total, dmode in res:
        assert (crank, cworld) == (rank, world) and mode == "p2p" and dmode == "p2p"
        assert got == [b"handle-of-rank-0", b"handle-of-rank-1"]
        assert total == [4.0] * 8                      # 1 (ones-init) + 1 + 2


def test_rejects_unknown_mode(monkeypatch):
    sys.path.insert(0, os.path.join(ROOT, "pytorch-dppo_b200"))
    import b200_comm
    monkeypatch.setenv("DPPO_COMM", "mpi")
    with pytest.raises(ValueError):
        b200_comm.PeerComm.create(group=object())
