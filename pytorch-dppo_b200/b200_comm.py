"""Multi-GPU plumbing for K3: one process per GPU, `torch.distributed` only for the bootstrap
(exchange of 64-byte IPC handles, barriers); the gradient exchange itself is the fused
peer-memory kernel `dppo_allreduce_adam` (csrc/dppo_comm.cu).

Replaces the reference's shared-memory bucket + Counter/TrafficLight/chief hand-shake
(model.py:47-59, utils.py:4-39, chief.py:7-21, train.py:169-175) for W workers on W GPUs.

`DPPO_COMM=nccl` selects the baseline instead: sum the local partials, NCCL all-reduce the
gradient bucket, then the single-GPU K3.  It exists to measure the fused kernel against."""
import os

import torch
import torch.distributed as dist

import b200_native as nv


def exchange(group, payload):
    """all-gather one picklable object per rank, returned in rank order (works on gloo and nccl)."""
    out = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, payload, group=group)
    return out


class PeerComm:
    def __init__(self, group, rank, world, mode):
        self.group, self.rank, self.world, self.mode = group, rank, world, mode
        self.handle = None
        self.P_pad = None

    @classmethod
    def create(cls, group=None, mode=None):
        group = group if group is not None else dist.group.WORLD
        mode = mode or os.environ.get("DPPO_COMM", "p2p")
        if mode not in ("p2p", "nccl"):
            raise ValueError("DPPO_COMM must be 'p2p' or 'nccl'")
        return cls(group, dist.get_rank(group), dist.get_world_size(group), mode)

    def _ensure(self, P_pad):
        if self.mode != "p2p" or self.handle is not None:
            return
        self.P_pad = P_pad
        self.handle = nv.comm_create(self.rank, self.world, P_pad)
        handles = exchange(self.group, nv.comm_local_handle(self.handle))
        if self.world > 1:
            nv.comm_open_peers(self.handle, handles)
        algo = os.environ.get("DPPO_COMM_ALGO")            # "push_all" | "owner"; default by world size (owner for W > 4)
        if algo:
            nv.comm_set_algorithm(self.handle, {"push_all": nv.COMM_ALGO_PUSH_ALL, "owner": nv.COMM_ALGO_OWNER}[algo])
        dist.barrier(group=self.group)

    def allreduce_adam(self, eng, k, gshared):
        """Epoch k of the engine: exchange + Adam.  Returns the number of kernels launched."""
        old = eng.theta_old if eng.flavour == nv.FLAVOUR_DPPO else None
        if self.mode == "p2p":
            self._ensure(eng.P_pad)
            nv.allreduce_adam(self.handle, eng.partials, eng.nparts, gshared, eng.theta, old, eng.m, eng.v, 0, eng.lr,
                              eng.betas[0], eng.betas[1], eng.eps, grad_out=eng.grad, loss_partials=eng.loss_partials,
                              loss_out=eng.loss_out[k], scalars_dev=eng.scalars_dev[k])
            return 1
        nv.reduce_partials(eng.partials, eng.nparts, eng.grad)
        dist.all_reduce(eng.grad, op=dist.ReduceOp.SUM, group=self.group)
        nv.reduce_adam(eng.grad.view(1, -1), 1, gshared, eng.theta, old, eng.m, eng.v, 0, eng.lr, eng.betas[0],
                       eng.betas[1], eng.eps, grad_out=None, loss_partials=eng.loss_partials, loss_out=eng.loss_out[k],
                       scalars_dev=eng.scalars_dev[k])
        return 2

    def allgather(self, src, dst):
        """Every rank's ``src`` (float64 vector) into ``dst`` (W, len) in rank order - the observation statistics that
        the reference's workers share through one shm object (main.py:54).  Returns the number of kernels launched."""
        if self.mode == "p2p":
            if self.handle is None:
                raise nv.DppoError("PeerComm.allgather before the communicator was set up (build the UpdateEngine first)")
            nv.comm_allgather(self.handle, src, dst)
        else:
            dist.all_gather_into_tensor(dst, src, group=self.group)
        return 1

    def describe(self):
        algo = None
        if self.mode == "p2p" and self.handle is not None:
            algo = {nv.COMM_ALGO_PUSH_ALL: "push_all", nv.COMM_ALGO_OWNER: "owner"}.get(nv.comm_get_algorithm(self.handle))
        return {"mode": self.mode, "world": self.world, "algorithm": algo,
                "kernel": "dppo_allreduce_adam (push over NVLink peer memory + flags, fused with Adam)"
                if self.mode == "p2p" else "dppo_reduce_partials + NCCL all_reduce + dppo_reduce_adam"}

    def close(self):
        if self.handle is not None:
            nv.comm_destroy(self.handle)
            self.handle = None
