#!/usr/bin/env python
"""Times the UNMODIFIED reference (oracle/_ref, copied by oracle/build_ref.py) on the host cores:
W forked workers running train.train + one chief running chief.chief, exactly as main.py:62-71 starts
them, one thread per process (main.py:42).

THIS IS TEST INFRASTRUCTURE (the CPU baseline of bench.py); the product never imports it.
It is run as a SUBPROCESS with a clean module path, because the reference's module names
(model, train, chief, utils, main) are also the drop-in package's.

    python oracle/ref_runner.py --O 17 --A 6 --W 1 --B 4096 --E 10 --T 4096 --epochs 25 [--as-is]

What the harness adds around the reference (nothing in oracle/_ref is edited; same shims as
tests/golden/make_golden.py):
  * stub gym / mujoco_py / pybullet_envs modules; the "environment" replays the synthetic rollout of
    pytorch-dppo_b200/synthetic_data.py (same seed -> the same raw observations and rewards the GPU arm
    consumes; actions are the reference's own samples);
  * torch 2.x shims: backward(retain_variables=) -> retain_graph, load_state_dict by .data.copy_,
    torch.set_num_threads(1) before fork; ReplayMemory.push detaches what it stores;
  * timing hooks: random.sample (start of an epoch, train.py:140), Shared_grad_buffers.add_gradient
    (end of the worker's part, train.py:169), optimizer.step (the chief's part, chief.py:17),
    ReplayMemory.clear (end of a cycle, train.py:178);
  * without --as-is the chief's time.sleep(1) (chief.py:10) is shortened to 2 ms: "compute-only".
Prints ONE JSON object: per-epoch compute spans (max over workers + the chief's step), per-cycle
wall times, process and thread counts."""
import argparse
import json
import os
import random
import sys
import tempfile
import time
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")
PKG = os.path.join(os.path.dirname(HERE), "pytorch-dppo_b200")


class _Space:
    def __init__(self, n):
        self.shape = (n,)


class ReplayEnv:
    """Serves the synthetic rollout's raw observations and rewards row by row; an episode ends where the
    synthetic rollout ends a segment with done (every `seg_len` steps with probability 1/2)."""

    def __init__(self, raw_obs, rew, coin, A, seg_len=1000):
        self.observation_space, self.action_space = _Space(raw_obs.shape[1]), _Space(A)
        self.obs, self.rew, self.coin, self.seg_len, self.t = raw_obs.numpy(), rew.numpy(), coin.numpy(), seg_len, 0

    def reset(self):
        return self.obs[self.t % len(self.obs)]

    def step(self, action):
        i = self.t % len(self.obs)
        self.t += 1
        done = (self.t % self.seg_len == 0) and bool(self.coin[i] < 0.5)
        return self.obs[self.t % len(self.obs)], float(self.rew[i]), done, {}


_ENV = {"make": None}


def install_stubs():
    gym = types.ModuleType("gym")
    gym.make = lambda name: _ENV["make"]()
    gym.wrappers = types.ModuleType("gym.wrappers")
    sys.modules.update({"gym": gym, "gym.wrappers": gym.wrappers, "mujoco_py": types.ModuleType("mujoco_py"),
                        "pybullet_envs": types.ModuleType("pybullet_envs")})
    sys.path.insert(0, REF)


def install_shims(model_mod):
    orig_backward = torch.Tensor.backward

    def backward(self, gradient=None, retain_graph=None, create_graph=False, inputs=None, retain_variables=None):
        if retain_variables is not None:
            retain_graph = retain_variables
        orig_backward(self, gradient, retain_graph, create_graph, inputs)

    torch.Tensor.backward = backward

    def load_state_dict(self, sd, strict=True):
        own = dict(self.named_parameters())
        for k, v in sd.items():
            own[k].data.copy_(v.data)

    model_mod.Model.load_state_dict = load_state_dict


class _Stop(Exception):
    pass


def worker(rank, a, params, traffic_light, counter, shared_model, sgb, sos, test_n, outdir):
    import model as model_mod
    import train as train_mod
    sys.path.insert(0, PKG)
    from synthetic_data import host_arrays                       # torch only; the CUDA library is not loaded
    install_shims(model_mod)
    raw, rew, _, _, coin = host_arrays(a.O, a.A, a.T + 2 * params.num_steps, 1 + rank)
    _ENV["make"] = lambda: ReplayEnv(raw, rew, coin, a.A)
    rec = dict(epochs=[], cycles=[])
    cur = {}
    orig_sample, orig_add = random.sample, model_mod.Shared_grad_buffers.add_gradient
    orig_push = train_mod.ReplayMemory.push

    def sample(pop, k):
        cur["t0"] = time.perf_counter()
        cur.setdefault("cycle0", cur["t0"])
        return orig_sample(pop, k)

    def add_gradient(self, model):
        orig_add(self, model)
        rec["epochs"].append((cur["t0"], time.perf_counter()))

    def push(self, events):                                      # synthetic rollouts are plain tensors
        orig_push(self, [[t.detach() for t in col] for col in events])

    def clear(self):
        rec["cycles"].append((cur.pop("cycle0"), time.perf_counter()))
        if len(rec["cycles"]) >= a.cycles:
            raise _Stop()
        self.memory = []

    random.sample = sample
    model_mod.Shared_grad_buffers.add_gradient = add_gradient
    train_mod.ReplayMemory.push = push
    train_mod.ReplayMemory.clear = clear
    random.seed(1 + rank)
    try:
        train_mod.train(rank, params, traffic_light, counter, shared_model, sgb, sos, test_n)
    except _Stop:
        pass
    torch.save(rec, os.path.join(outdir, "worker%d.pt" % rank))


def chief_proc(a, params, traffic_light, counter, shared_model, sgb, optimizer, outdir):
    import chief as chief_mod
    if not a.as_is:
        real_sleep = time.sleep
        chief_mod.time.sleep = lambda s: real_sleep(0.002)
    steps = []
    orig_step = optimizer.step

    def step(*args, **kw):
        t0 = time.perf_counter()
        orig_step(*args, **kw)
        steps.append((t0, time.perf_counter()))
        if len(steps) % params.num_epoch == 0:
            torch.save(steps, os.path.join(outdir, "chief.pt"))

    optimizer.step = step
    chief_mod.chief(0, params, traffic_light, counter, shared_model, sgb, optimizer)


def main():
    ap = argparse.ArgumentParser()
    for k, d in (("O", 17), ("A", 6), ("W", 1), ("B", 4096), ("E", 10), ("T", 4096), ("epochs", 20)):
        ap.add_argument("--" + k, type=int, default=d)
    ap.add_argument("--as-is", action="store_true")
    a = ap.parse_args()
    a.cycles = max(1, -(-a.epochs // a.E))
    import warnings
    warnings.filterwarnings("ignore")
    install_stubs()
    import torch.multiprocessing as mp
    import torch.optim as optim
    import main as main_mod
    import model as model_mod
    import utils as utils_mod
    torch.set_num_threads(1)                                     # before fork (main.py:42 comes too late in torch 2)
    params = main_mod.Params()
    params.batch_size, params.num_epoch, params.exploration_size = a.B, a.E, a.T
    params.num_processes, params.update_treshold = a.W, a.W - 1
    torch.manual_seed(params.seed)
    traffic_light, counter = utils_mod.TrafficLight(), utils_mod.Counter()
    shared_model = model_mod.Model(a.O, a.A)
    shared_model.share_memory()
    sgb = model_mod.Shared_grad_buffers(shared_model)
    sos = model_mod.Shared_obs_stats(a.O)
    optimizer = optim.Adam(shared_model.parameters(), lr=params.lr)
    test_n = torch.Tensor([0])
    test_n.share_memory_()
    outdir = tempfile.mkdtemp(prefix="dppo_ref_")
    ctx = mp.get_context("fork")
    pc = ctx.Process(target=chief_proc, args=(a, params, traffic_light, counter, shared_model, sgb, optimizer, outdir))
    pc.start()
    ws = [ctx.Process(target=worker, args=(r, a, params, traffic_light, counter, shared_model, sgb, sos, test_n, outdir))
          for r in range(a.W)]
    t_start = time.perf_counter()
    for p in ws:
        p.start()
    for p in ws:
        p.join()
        if p.exitcode != 0:
            pc.terminate()
            raise SystemExit("reference worker failed with exit code %s" % p.exitcode)
    wall = time.perf_counter() - t_start
    pc.terminate()
    pc.join()
    recs = [torch.load(os.path.join(outdir, "worker%d.pt" % r), weights_only=False) for r in range(a.W)]
    chief = torch.load(os.path.join(outdir, "chief.pt"), weights_only=False)
    n_ep = min(min(len(r["epochs"]) for r in recs), len(chief))
    epoch_s = []
    for k in range(n_ep):
        work = max(r["epochs"][k][1] - r["epochs"][k][0] for r in recs)          # slowest worker: sample .. add_gradient
        epoch_s.append(work + (chief[k][1] - chief[k][0]))                       # + the chief's Adam step
    cycle_wall = [max(r["cycles"][c][1] for r in recs) - min(r["cycles"][c][0] for r in recs) for c in range(a.cycles)]
    print(json.dumps({"epoch_s": epoch_s, "cycle_wall_s": cycle_wall, "workers": a.W, "processes": a.W + 1,
                      "threads_per_process": 1, "as_is": bool(a.as_is), "epochs": n_ep, "cycles": a.cycles,
                      "total_wall_s": wall, "rollout_steps_per_worker": a.T, "torch": torch.__version__}), flush=True)


if __name__ == "__main__":
    main()
