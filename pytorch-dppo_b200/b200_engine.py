"""Host-side orchestration of the update path: SoA rollout buffer in HBM and the
K1 -> E x (K2 -> K3) sequence, captured once in a CUDA graph and replayed per
rollout.  torch supplies device memory and streams; all arithmetic is in
libdppo_b200.so (b200_native).  There is no CPU path."""
import random
from collections import OrderedDict

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
        if comm is not None:
            comm._ensure(self.P_pad)                 # collective: every rank builds its engine at the same point
        dev = model.flat.device
        self.device = dev
        self.theta = model.flat
        self.theta_old = torch.empty_like(self.theta)
        self.m = torch.zeros_like(self.theta)
        self.v = torch.zeros_like(self.theta)
        self.step = 0
        nv.update_prepare(self.dims, self.B)
        self.k2_launches = nv.update_launches(self.dims, self.B)
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
        self._graphs = OrderedDict()             # captured cycles, see _graph_key; at most MAX_GRAPHS, oldest evicted
        self._rollout = None                     # the engine's own rollout buffer (rollout_buffer())
        self.launches_per_cycle = 0

    MAX_GRAPHS = 4

    def rollout_buffer(self, T):
        """The engine's persistent rollout buffer of T steps.  Training loops fill THIS buffer every outer
        iteration (train.collect_rollout(out=...)) instead of allocating a new one: its addresses stay fixed, so
        the cycle captured for it is replayed for as long as the loop runs."""
        if self._rollout is None or self._rollout.T != T:
            self._rollout = RolloutBuffer(T, self.dims.O, self.dims.A, self.device)
        return self._rollout

    def _graph_key(self, buf, gshared, obs_stats, with_gae):
        """A captured cycle bakes in device ADDRESSES, so that is what identifies it: every rollout column, the
        shared gradient bucket, the statistics vectors and the scratch buffers.  (Python object ids are reused
        once an object is freed; a new buffer with a recycled id must never hit an old graph.)"""
        ptrs = [getattr(buf, f).data_ptr() for f in RolloutBuffer.FIELDS]
        ptrs.append(0 if gshared is None else gshared.data_ptr())
        if obs_stats is not None:
            ptrs += [t.data_ptr() for t in (obs_stats.n, obs_stats.mean, obs_stats.mean_diff, obs_stats.var,
                                            obs_stats._scratch_for(buf.T))]
            if self.comm is not None and self.comm.world > 1:
                ptrs += [t.data_ptr() for t in obs_stats._shared_for(buf.T, self.comm.world)]
        if with_gae:
            ptrs.append(self._gae_scratch.data_ptr())
        return (buf.T, bool(with_gae), obs_stats is not None) + tuple(ptrs)

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
        self._wait_idx_upload()
        nv.sample_indices(T, self.B, self.E, rng=rng, out=self.idx_host)
        return self.idx_host

    def draw_indices_device(self, T, sampler, out=None):
        """The same E x B positions drawn ON THE DEVICE (b200_native.DeviceSampler: CPython's random.sample with the
        Mersenne Twister state in device memory), written straight into the index block K2 reads: nothing crosses
        PCIe."""
        sampler.sample(T, self.B, self.E, out=self.idx_dev if out is None else out)

    def _wait_idx_upload(self):
        """The pinned index block is uploaded asynchronously; it may only be rewritten once that upload has run
        (a loop of draw_indices / run_epochs without a synchronise in between would otherwise train on a mixture
        of two index streams, depending on timing)."""
        ev = getattr(self, "_idx_read", None)
        if ev is not None:
            ev.synchronize()
            self._idx_read = None

    def _upload_indices(self):
        self.idx_dev.copy_(self.idx_host, non_blocking=True)
        self._idx_read = torch.cuda.Event()
        self._idx_read.record(torch.cuda.current_stream())

    # ---- E x (K2 -> K3) -----------------------------------------------------------------------
    def _launch_epochs(self, buf, gshared, obs_stats=None, with_gae=False):
        n = 0
        if obs_stats is not None:                              # K4: raw observations -> normalised states
            obs_stats.observes_batch(buf.raw_obs, out=buf.obs, comm=self.comm)
            n += 1 if self.comm is None or self.comm.world == 1 else 4
        if with_gae:                                           # K1
            self.gae(buf)
            n += 1
        if self.flavour == nv.FLAVOUR_DPPO:
            self.theta_old.copy_(self.theta)                       # model_old = copy of model, train.py:128-129
            n += 1
        if nv.update_nparts(self.dims, self.B) != self.nparts:
            # the partial buffers were sized for another form of K2 (update_set_path / update_set_batch_dw after __init__)
            raise RuntimeError("UpdateEngine: the K2 path changed after the engine was built (partials sized for %d rows, "
                               "K2 now writes %d); build a new engine" % (self.nparts, nv.update_nparts(self.dims, self.B)))
        for k in range(self.E):
            # epoch 0: model_old is a copy of model (train.py:128-129) - passing theta itself lets K2 skip the
            # second forward; from epoch 1 on theta_old is the lagging copy K3 maintains (train.py:164)
            old = None if self.flavour != nv.FLAVOUR_DPPO else (self.theta if k == 0 else self.theta_old)
            nv.update_fused(self.flavour, self.dims, buf.obs, buf.act, buf.adv, buf.ret,
                            buf.logp if self.flavour == nv.FLAVOUR_PPO else None, self.idx_dev[k], self.theta,
                            old, self.clip, self.ent_coeff, self.partials, self.loss_partials)
            n += self.k2_launches
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
        ev = getattr(self, "_scalars_read", None)
        if ev is not None:
            ev.synchronize()                          # the previous upload of the pinned block has been consumed
        for k in range(self.E):
            s = nv.adam_scalars(self.step + k + 1, self.lr, self.betas[0], self.betas[1], self.eps)
            self.scalars_host[k, :6] = torch.tensor(s)
        self.scalars_dev.copy_(self.scalars_host, non_blocking=True)
        self._scalars_read = torch.cuda.Event()
        self._scalars_read.record(torch.cuda.current_stream())

    def run_epochs(self, buf, gshared=None, use_graph=True, indices_staged=False, obs_stats=None, with_gae=False,
                   scalars_staged=False):
        """The E epochs of one outer iteration on minibatch positions already in
        ``idx_host`` (see draw_indices).  Returns the (E,4) device tensor of per-epoch
        loss terms (clip, value, entropy, -)."""
        if not indices_staged:
            self._upload_indices()
        if not scalars_staged:
            self._stage_scalars()
        if not use_graph or (self.comm is not None and self.comm.mode == "nccl"):
            self.launches_per_cycle = self._launch_epochs(buf, gshared, obs_stats, with_gae)
        else:
            # scratch buffers are allocated (or grown) outside the capture; their addresses are part of the key
            if with_gae and (self._gae_scratch is None or self._gae_scratch.numel() < nv.gae_scratch_bytes(buf.T)):
                self._gae_scratch = torch.empty(nv.gae_scratch_bytes(buf.T), dtype=torch.uint8, device=self.device)
            key = self._graph_key(buf, gshared, obs_stats, with_gae)
            if key in self._graphs:
                self._graphs.move_to_end(key)
            else:
                # warm-up launch outside capture (sets kernel attributes), on scratch copies of the state
                state = [self.theta, self.theta_old, self.m, self.v]
                if obs_stats is not None:
                    state += [obs_stats.n, obs_stats.mean, obs_stats.mean_diff, obs_stats.var]
                saved = [t.clone() for t in state]
                gsaved = None if gshared is None else gshared.clone()
                self._launch_epochs(buf, gshared, obs_stats, with_gae)
                torch.cuda.synchronize()
                for t, s in zip(state, saved):
                    t.copy_(s)
                if gshared is not None:
                    gshared.copy_(gsaved)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    n = self._launch_epochs(buf, gshared, obs_stats, with_gae)
                for t, s in zip(state, saved):
                    t.copy_(s)                                     # capture does not execute, but be explicit
                # the entry keeps the tensors whose addresses the graph holds alive for as long as it exists
                self._graphs[key] = (g, n, buf, gshared, obs_stats, self._gae_scratch)
                while len(self._graphs) > self.MAX_GRAPHS:
                    self._graphs.popitem(last=False)
            g, n = self._graphs[key][:2]
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
            obs_stats.observes_batch(dev.raw_obs, out=dev.obs, comm=self.comm)
        self.gae(dev)
        # the host draws the minibatch positions while the GPU is busy with the copies, K4 and K1
        if self._prefetched == (dev.T, id(rng)):
            self.idx_host, self._idx_next = self._idx_next, self.idx_host
        else:
            self.draw_indices(dev.T, rng)
        self._prefetched = None
        self._upload_indices()
        h2d += self.idx_host.numel() * 4 + self.scalars_host.numel() * 4
        self.run_epochs(dev, gshared=gshared, use_graph=use_graph, indices_staged=True)
        self.loss_host.copy_(self.loss_out, non_blocking=True)
        if prefetch_next:
            nv.sample_indices(dev.T, self.B, self.E, rng=rng, out=self._idx_next)
            self._prefetched = (dev.T, id(rng))
        torch.cuda.current_stream().synchronize()
        return self.loss_host, h2d, self.loss_host.numel() * 4


    def update_stream(self, hosts, devs, n_steps, obs_stats=None, rng=random, gshared=None, on_step=None, sampler=None):
        """``n_steps`` consecutive outer iterations from HOST rollouts, software-pipelined the way a training
        loop would run them: a copy stream uploads rollout i+1 (pinned ``hosts[(i+1) % len(hosts)]`` ->
        ``devs[(i+1) % 2]``) while the compute stream runs K4, K1 and the E epochs of rollout i, and the host
        draws the next minibatch positions meanwhile.  Every step still copies its own inputs H2D and its
        per-epoch losses D2H.  With ``sampler`` (a b200_native.DeviceSampler) the minibatch positions are drawn on the
        GPU instead - on a side stream, one step ahead - and no index block is uploaded.
        Returns (list of per-step loss tensors on the host, h2d bytes/step, d2h bytes/step)."""
        assert len(devs) == 2
        if sampler is not None:
            return self._update_stream_device_sampler(hosts, devs, n_steps, obs_stats, gshared, on_step, sampler)
        fields = ["act", "rew", "val", "boot", "seg_end"] + (["raw_obs"] if obs_stats is not None else ["obs"])
        if self.flavour == nv.FLAVOUR_PPO:
            fields.append("logp")
        main = torch.cuda.current_stream()
        if not hasattr(self, "_copy_stream"):
            self._copy_stream = torch.cuda.Stream()
            self._loss_ring = [torch.zeros(self.E, 4, pin_memory=True) for _ in range(4)]
        cs = self._copy_stream
        copied = [torch.cuda.Event(), torch.cuda.Event()]
        consumed = [torch.cuda.Event(), torch.cuda.Event()]
        h2d = 0
        cs.wait_stream(main)
        with torch.cuda.stream(cs):
            h2d = devs[0].copy_from(hosts[0], fields)
            copied[0].record(cs)
        # minibatch positions are drawn by a worker thread (the C sampler releases the GIL) one step ahead,
        # into the pinned block that is not being uploaded
        if not hasattr(self, "_sampler_pool"):
            from concurrent.futures import ThreadPoolExecutor
            self._sampler_pool = ThreadPoolExecutor(max_workers=1)
        T = devs[0].T
        self.draw_indices(T, rng)
        losses = []
        idx_read = [None, None]                       # per pinned block: event after the upload that read it
        cur = 0
        blocks = [self.idx_host, self._idx_next]
        fut = None
        for i in range(n_steps):
            b = i & 1
            if fut is not None:
                fut.result()                          # positions of this step are in blocks[cur]
            main.wait_event(copied[b])
            self.idx_host = blocks[cur]
            self.idx_dev.copy_(self.idx_host, non_blocking=True)
            idx_read[cur] = torch.cuda.Event()
            idx_read[cur].record(main)
            if i + 1 < n_steps:
                nxt = 1 - cur
                if idx_read[nxt] is not None:
                    idx_read[nxt].synchronize()       # the upload that read that block two steps ago is done
                fut = self._sampler_pool.submit(nv.sample_indices, T, self.B, self.E, rng, blocks[nxt])
                cur_next = nxt
            self._stage_scalars()
            small_done = torch.cuda.Event()
            small_done.record(main)                   # this step's small uploads have left the H2D engine
            self.run_epochs(devs[b], gshared=gshared, use_graph=True, indices_staged=True, obs_stats=obs_stats,
                            with_gae=True, scalars_staged=True)   # K4 + K1 + E x (K2, K3): one graph replay
            consumed[b].record(main)
            if i + 1 < n_steps:
                # upload the next rollout behind this step's compute.  It may only start once this step's own
                # small uploads are through: there is ONE H2D copy engine, and the 7 MB upload becomes runnable
                # at the same moment as the index block of this step - queued ahead of it, it held every step
                # back by 0.2 ms (tools/e2e_gaps.py)
                with torch.cuda.stream(cs):
                    cs.wait_event(small_done)
                    if i >= 1:
                        cs.wait_event(consumed[1 - b])
                    devs[1 - b].copy_from(hosts[(i + 1) % len(hosts)], fields)
                    copied[1 - b].record(cs)
            out = self._loss_ring[i % len(self._loss_ring)]
            out.copy_(self.loss_out, non_blocking=True)
            losses.append(out)
            if on_step is not None:
                on_step(i)
            if i + 1 < n_steps:
                cur = cur_next
        self.idx_host, self._idx_next = blocks[cur], blocks[1 - cur]
        main.synchronize()
        return losses, h2d + self.idx_host.numel() * 4 + self.scalars_host.numel() * 4, self.E * 16


    def _stream_fields(self, obs_stats):
        fields = ["act", "rew", "val", "boot", "seg_end"] + (["raw_obs"] if obs_stats is not None else ["obs"])
        if self.flavour == nv.FLAVOUR_PPO:
            fields.append("logp")
        return fields


def _update_stream_device_sampler(self, hosts, devs, n_steps, obs_stats, gshared, on_step, sampler):
    fields = self._stream_fields(obs_stats)
    main = torch.cuda.current_stream()
    if not hasattr(self, "_copy_stream"):
        self._copy_stream = torch.cuda.Stream()
        self._loss_ring = [torch.zeros(self.E, 4, pin_memory=True) for _ in range(4)]
    if not hasattr(self, "_sample_stream"):
        self._sample_stream = torch.cuda.Stream()
        self._idx_next_dev = torch.zeros_like(self.idx_dev)
    cs, ss = self._copy_stream, self._sample_stream
    copied = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    cs.wait_stream(main)
    ss.wait_stream(main)
    with torch.cuda.stream(cs):
        h2d = devs[0].copy_from(hosts[0], fields)
        copied[0].record(cs)
    T = devs[0].T
    drawn, taken = torch.cuda.Event(), torch.cuda.Event()
    with torch.cuda.stream(ss):
        self.draw_indices_device(T, sampler, out=self._idx_next_dev)
        drawn.record(ss)
    losses = []
    for i in range(n_steps):
        b = i & 1
        main.wait_event(copied[b])
        main.wait_event(drawn)
        self.idx_dev.copy_(self._idx_next_dev)                    # device to device, 4*E*B bytes
        taken.record(main)
        if i + 1 < n_steps:
            with torch.cuda.stream(ss):                           # next step's positions while this step computes
                ss.wait_event(taken)
                self.draw_indices_device(T, sampler, out=self._idx_next_dev)
                drawn = torch.cuda.Event()
                drawn.record(ss)
        self._stage_scalars()
        small_done = torch.cuda.Event()
        small_done.record(main)
        self.run_epochs(devs[b], gshared=gshared, use_graph=True, indices_staged=True, obs_stats=obs_stats,
                        with_gae=True, scalars_staged=True)
        consumed[b].record(main)
        if i + 1 < n_steps:
            with torch.cuda.stream(cs):
                cs.wait_event(small_done)
                if i >= 1:
                    cs.wait_event(consumed[1 - b])
                devs[1 - b].copy_from(hosts[(i + 1) % len(hosts)], fields)
                copied[1 - b].record(cs)
        out = self._loss_ring[i % len(self._loss_ring)]
        out.copy_(self.loss_out, non_blocking=True)
        losses.append(out)
        if on_step is not None:
            on_step(i)
    main.synchronize()
    return losses, h2d + self.scalars_host.numel() * 4, self.E * 16


UpdateEngine._update_stream_device_sampler = _update_stream_device_sampler


# ---- checkpoint / resume (the reference has none: SURVEY section 5; "next" row N4) --------------------
def save_checkpoint(path, engine, obs_stats=None, rng=None):
    """Parameters, Adam moments and update count, running observation statistics and the Python
    generator that draws the minibatch positions - everything a bit-identical resume needs."""
    state = {"theta": engine.theta.detach().cpu(), "m": engine.m.cpu(), "v": engine.v.cpu(), "step": engine.step,
             "dims": tuple(getattr(engine.dims, f) for f in ("O", "A", "H1", "H2", "G1", "G2"))}
    if obs_stats is not None:
        state["obs_stats"] = {k: getattr(obs_stats, k).cpu() for k in ("n", "mean", "mean_diff", "var")}
    if rng is not None:
        state["rng"] = rng.getstate()
    torch.save(state, path)


def load_checkpoint(path, engine, obs_stats=None, rng=None):
    state = torch.load(path, weights_only=False)
    if tuple(state["dims"]) != tuple(getattr(engine.dims, f) for f in ("O", "A", "H1", "H2", "G1", "G2")):
        raise nv.DppoError("checkpoint was written for a different network shape")
    engine.theta.copy_(state["theta"])
    engine.m.copy_(state["m"])
    engine.v.copy_(state["v"])
    engine.step = int(state["step"])
    engine._prefetched = None
    if obs_stats is not None and "obs_stats" in state:
        for k, t in state["obs_stats"].items():
            getattr(obs_stats, k).copy_(t)
    if rng is not None and "rng" in state:
        rng.setstate(state["rng"])
