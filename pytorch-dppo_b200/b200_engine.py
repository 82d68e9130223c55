"""Host-side orchestration of the update path: SoA rollout buffer in HBM and the
K1 -> E x (K2 -> K3) sequence, captured once in a CUDA graph and replayed per
rollout.  torch supplies device memory and streams; all arithmetic is in
libdppo_b200.so (b200_native).  There is no CPU path."""
import random

import torch

import b200_native as nv


class RolloutBuffer:
    """Struct-of-arrays replacement of ReplayMemory's list of per-step tuples
    (train.py:17-33): one contiguous device array per column, so the minibatch
    gather is an int32 index lookup inside K2.

    raw_obs (T,O) raw observations (input of K4)      obs (T,O) normalised states
    act (T,A)   rew (T) clipped rewards   val (T) V(s_t)   seg_end (T) uint8   boot (T)
    adv (T), ret (T) written by K1        logp (T) stored log-probs (ppo.py flavour)"""

    FIELDS = ("raw_obs", "obs", "act", "rew", "val", "boot", "seg_end", "adv", "ret", "logp")

    def __init__(self, T, O, A, device, pinned_host=False):
        self.T, self.O, self.A = T, O, A
        kw = dict(device="cpu", pin_memory=True) if pinned_host else dict(device=device)
        f32 = torch.float32
        self.raw_obs = torch.zeros(T, O, dtype=f32, **kw)
        self.obs = torch.zeros(T, O, dtype=f32, **kw)
        self.act = torch.zeros(T, A, dtype=f32, **kw)
        self.rew = torch.zeros(T, dtype=f32, **kw)
        self.val = torch.zeros(T, dtype=f32, **kw)
        self.boot = torch.zeros(T, dtype=f32, **kw)
        self.seg_end = torch.zeros(T, dtype=torch.uint8, **kw)
        self.adv = torch.zeros(T, dtype=f32, **kw)
        self.ret = torch.zeros(T, dtype=f32, **kw)
        self.logp = torch.zeros(T, dtype=f32, **kw)

    def copy_from(self, other, fields, non_blocking=True):
        n = 0
        for f in fields:
            src = getattr(other, f)
            getattr(self, f).copy_(src, non_blocking=non_blocking)
            n += src.numel() * src.element_size()
        return n


class UpdateEngine:
    """One worker's update machinery (train.py:108-175 + chief.py:13-20, or
    ppo.py:118-174) on one GPU.

    theta        = model.flat (updated in place by K3)
    theta_old    = the lagging copy of train.py:128-129,164 (K3 writes it)
    m, v, step   = Adam moments and update count (chief's optimizer state)
    partials     = [nparts][P_pad] per-CTA gradient partials (K2 -> K3)
    grad         = the summed (and clipped) gradient of the last epoch; also
                   published as model.flat_grad for Shared_grad_buffers.add_gradient"""

    def __init__(self, model, flavour, batch_size, num_epoch, lr, clip=0.2, ent_coeff=0.0, max_grad_norm=0.0,
                 gamma=0.99, gae_param=0.95, betas=(0.9, 0.999), eps=1e-8, comm=None):
        if not model.flat.is_cuda:
            raise nv.DppoError("UpdateEngine needs the model on a CUDA device - there is no CPU path")
        self.model, self.flavour = model, flavour
        self.dims, self.P_pad = model.dims, model.P_pad
        self.B, self.E = int(batch_size), int(num_epoch)
        self.lr, self.clip, self.ent_coeff, self.max_grad_norm = lr, clip, ent_coeff, max_grad_norm
        self.gamma, self.gae_param, self.betas, self.eps = gamma, gae_param, betas, eps
        self.comm = comm
        dev = model.flat.device
        self.device = dev
        self.theta = model.flat
        self.theta_old = torch.empty_like(self.theta)
        self.m = torch.zeros_like(self.theta)
        self.v = torch.zeros_like(self.theta)
        self.step = 0
        self.nparts = nv.update_nparts(self.dims, self.B)
        self.partials = torch.empty(self.nparts, self.P_pad, device=dev)
        self.loss_partials = torch.empty(self.nparts, 4, device=dev)
        self.loss_out = torch.zeros(self.E, 4, device=dev)
        self.grad = torch.zeros(self.P_pad, device=dev)
        model.flat_grad = self.grad
        self.idx_dev = torch.zeros(self.E, self.B, dtype=torch.int32, device=dev)
        self.idx_host = torch.zeros(self.E, self.B, dtype=torch.int32, pin_memory=True)
        self._idx_next = torch.zeros(self.E, self.B, dtype=torch.int32, pin_memory=True)
        self._prefetched = None                  # (T, id(rng)) of the positions already drawn into _idx_next
        self.scalars_dev = torch.zeros(self.E, 8, device=dev)
        self.scalars_host = torch.zeros(self.E, 8, pin_memory=True)
        self.loss_host = torch.zeros(self.E, 4, pin_memory=True)
        self._adam_scratch = torch.empty(max(nv.reduce_adam_scratch_bytes(self.P_pad), 16), dtype=torch.uint8, device=dev)
        self._gae_scratch = None
        self._graphs = {}
        self.launches_per_cycle = 0

    # ---- K1 ---------------------------------------------------------------------------------
    def gae(self, buf):
        need = nv.gae_scratch_bytes(buf.T)
        if self._gae_scratch is None or self._gae_scratch.numel() < need:
            self._gae_scratch = torch.empty(need, dtype=torch.uint8, device=self.device)
        nv.gae_scan(buf.rew, buf.val, buf.boot, buf.seg_end, buf.adv, buf.ret, self.gamma, self.gae_param,
                    self._gae_scratch)

    # ---- minibatch positions (a4) --------------------------------------------------------------
    def draw_indices(self, T, rng=random):
        """E minibatches of B positions, exactly ReplayMemory.sample's index stream
        (``random.sample`` without replacement, independent per epoch; train.py:31-33),
        generated by the C restatement of CPython's algorithm on ``rng``'s own state."""
        nv.sample_indices(T, self.B, self.E, rng=rng, out=self.idx_host)
        return self.idx_host

    # ---- E x (K2 -> K3) -----------------------------------------------------------------------
    def _launch_epochs(self, buf, gshared):
        n = 0
        if self.flavour == nv.FLAVOUR_DPPO:
            self.theta_old.copy_(self.theta)                       # model_old = copy of model, train.py:128-129
            n += 1
        for k in range(self.E):
            nv.update_fused(self.flavour, self.dims, buf.obs, buf.act, buf.adv, buf.ret,
                            buf.logp if self.flavour == nv.FLAVOUR_PPO else None, self.idx_dev[k], self.theta,
                            self.theta_old if self.flavour == nv.FLAVOUR_DPPO else None, self.clip, self.ent_coeff,
                            self.partials, self.loss_partials)
            n += 1
            if self.comm is not None:
                n += self.comm.allreduce_adam(self, k, gshared)
            else:
                nv.reduce_adam(self.partials, self.nparts, gshared, self.theta,
                               self.theta_old if self.flavour == nv.FLAVOUR_DPPO else None, self.m, self.v,
                               0, self.lr, self.betas[0], self.betas[1], self.eps,
                               max_grad_norm=self.max_grad_norm, grad_out=self.grad, loss_partials=self.loss_partials,
                               loss_out=self.loss_out[k], scratch=self._adam_scratch, scalars_dev=self.scalars_dev[k])
                n += 2 if self.max_grad_norm > 0 else 1
        return n

    def _stage_scalars(self):
        for k in range(self.E):
            s = nv.adam_scalars(self.step + k + 1, self.lr, self.betas[0], self.betas[1], self.eps)
            self.scalars_host[k, :6] = torch.tensor(s)
        self.scalars_dev.copy_(self.scalars_host, non_blocking=True)

    def run_epochs(self, buf, gshared=None, use_graph=True, indices_staged=False):
        """The E epochs of one outer iteration on minibatch positions already in
        ``idx_host`` (see draw_indices).  Returns the (E,4) device tensor of per-epoch
        loss terms (clip, value, entropy, -)."""
        if not indices_staged:
            self.idx_dev.copy_(self.idx_host, non_blocking=True)
        self._stage_scalars()
        if not use_graph or (self.comm is not None and self.comm.mode == "nccl"):
            self.launches_per_cycle = self._launch_epochs(buf, gshared)
        else:
            key = (id(buf), None if gshared is None else gshared.data_ptr())
            if key not in self._graphs:
                # warm-up launch outside capture (sets kernel attributes), on scratch copies of the state
                saved = [t.clone() for t in (self.theta, self.theta_old, self.m, self.v)]
                gsaved = None if gshared is None else gshared.clone()
                self._launch_epochs(buf, gshared)
                torch.cuda.synchronize()
                for t, s in zip((self.theta, self.theta_old, self.m, self.v), saved):
                    t.copy_(s)
                if gshared is not None:
                    gshared.copy_(gsaved)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    n = self._launch_epochs(buf, gshared)
                for t, s in zip((self.theta, self.theta_old, self.m, self.v), saved):
                    t.copy_(s)                                     # capture does not execute, but be explicit
                self._graphs[key] = (g, n)
            g, n = self._graphs[key]
            g.replay()
            self.launches_per_cycle = n
        self.step += self.E
        return self.loss_out

    # ---- whole outer iteration from a host rollout (the end-to-end call) ----------------------
    def update_from_host(self, host, dev, obs_stats=None, rng=random, use_graph=True, gshared=None,
                         prefetch_next=False):
        """H2D copy of one rollout (pinned ``host`` buffer -> ``dev``), K4 (if obs_stats),
        K1, E x (K2,K3), D2H copy of the per-epoch losses.  Returns (loss_host, h2d_bytes,
        d2h_bytes).  With ``prefetch_next`` the minibatch positions of the NEXT call are drawn
        from ``rng`` while the GPU works on this one (same ``random.sample`` stream, drawn
        earlier); the next call must then use the same ``rng`` and rollout length."""
        fields = ["act", "rew", "val", "boot", "seg_end"] + (["raw_obs"] if obs_stats is not None else ["obs"])
        if self.flavour == nv.FLAVOUR_PPO:
            fields.append("logp")
        h2d = dev.copy_from(host, fields)
        if obs_stats is not None:
            obs_stats.observes_batch(dev.raw_obs, out=dev.obs)
        self.gae(dev)
        # the host draws the minibatch positions while the GPU is busy with the copies, K4 and K1
        if self._prefetched == (dev.T, id(rng)):
            self.idx_host, self._idx_next = self._idx_next, self.idx_host
        else:
            self.draw_indices(dev.T, rng)
        self._prefetched = None
        self.idx_dev.copy_(self.idx_host, non_blocking=True)
        h2d += self.idx_host.numel() * 4 + self.scalars_host.numel() * 4
        self.run_epochs(dev, gshared=gshared, use_graph=use_graph, indices_staged=True)
        self.loss_host.copy_(self.loss_out, non_blocking=True)
        if prefetch_next:
            nv.sample_indices(dev.T, self.B, self.E, rng=rng, out=self._idx_next)
            self._prefetched = (dev.T, id(rng))
        torch.cuda.current_stream().synchronize()
        return self.loss_host, h2d, self.loss_host.numel() * 4
