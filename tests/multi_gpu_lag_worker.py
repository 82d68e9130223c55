"""torchrun target (2 GPUs): rank 1 makes one exchange call FEWER than rank 0.  Rank 0's kernel must give up polling after
its clock budget (a few seconds), record the error and return - dppo_comm_check then raises - instead of spinning forever
inside the kernel (the reference deadlocks when a worker dies: chief.py:13, train.py:174-175)."""
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "pytorch-dppo_b200")):
    sys.path.insert(0, p)


def main():
    import b200_native as nv
    from b200_comm import PeerComm
    rank = int(os.environ["RANK"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    P = 4096
    comm = PeerComm.create()
    comm._ensure(P)
    th, m, v = torch.zeros(P, device=dev), torch.zeros(P, device=dev), torch.zeros(P, device=dev)
    parts = torch.ones(3, P, device=dev)
    nv.allreduce_adam(comm.handle, parts, 3, None, th, None, m, v, 1, 1e-3)          # both ranks: completes
    torch.cuda.synchronize()
    nv.comm_check(comm.handle)
    dist.barrier()
    if rank == 0:
        t0 = time.time()
        nv.allreduce_adam(comm.handle, parts, 3, None, th, None, m, v, 2, 1e-3)      # rank 1 never makes this call
        torch.cuda.synchronize()
        waited = time.time() - t0
        try:
            nv.comm_check(comm.handle)
            print("LAG_NOT_DETECTED", flush=True)
        except nv.DppoError as e:
            print("LAG_DETECTED after %.1f s: %s" % (waited, e), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
