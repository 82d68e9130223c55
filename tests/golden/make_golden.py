#!/usr/bin/env python
"""Generate the golden fixtures in tests/golden/ by running the UNMODIFIED
reference (/root/reference) in the build container.

Run once, here (the reference does not travel to the GPU box):

    python tests/golden/make_golden.py

Nothing in the reference is edited.  What the harness adds around it
(SURVEY.md section 8(c)):
  * stub ``gym`` / ``gym.wrappers`` / ``mujoco_py`` / ``pybullet_envs`` modules
    and a synthetic environment (gym is not installed, there is no network);
  * three compatibility shims for torch 2.x: ``backward(retain_variables=)``
    -> ``retain_graph`` (train.py:166), ``load_state_dict`` by ``.data.copy_``
    (train.py:135,164 overwrite weights between forward and backward, which
    torch 2 version-checks), ``torch.set_num_threads(1)`` before fork;
  * ``ReplayMemory.push`` detaches what it stores (synthetic rollouts are plain
    tensors; the rollout-graph leak of train.py:89,121 is not part of the path);
  * recording hooks only: ``random.sample`` (index stream), ``push``/``sample``
    (SoA buffer), ``observes`` (raw observations), ``backward``/``add_gradient``
    (loss terms from the caller's frame, the 13 gradients), ``optimizer.step``
    (parameters and Adam moments);
  * deterministic scheduling of the reference's unlocked shared-memory writes:
    workers roll out one after the other and add gradients in rank order
    (model.py:53-55 and 68-75 are racy by design), ``time.sleep`` in the chief
    is shortened.
"""
import os
import random
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


# ----------------------------------------------------------------------------
# synthetic environment + stub modules
# ----------------------------------------------------------------------------
class _Space:
    def __init__(self, n):
        self.shape = (n,)


class FakeEnv:
    """Observations ~ N(loc_f, scale_f) per feature, rewards ~ 0.8*N(0,1)
    (so the [-1,1] clip at train.py:96 bites), episodes end after a fixed
    cycle of lengths.  All values are fp32-representable Python floats."""

    def __init__(self, O, A, seed, ep_lens):
        self.observation_space = _Space(O)
        self.action_space = _Space(A)
        self.rng = np.random.RandomState(seed)
        self.loc = self.rng.uniform(-2, 2, O).astype(np.float32)
        self.scale = self.rng.uniform(0.05, 3, O).astype(np.float32)
        self.ep_lens = list(ep_lens)
        self.ep = 0
        self.t = 0

    def _obs(self):
        return (self.loc + self.scale * self.rng.randn(len(self.loc)).astype(np.float32)).astype(np.float32)

    def reset(self):
        self.t = 0
        return self._obs()

    def step(self, action):
        self.t += 1
        r = float(np.float32(0.8 * self.rng.randn()))
        done = self.t >= self.ep_lens[self.ep % len(self.ep_lens)]
        if done:
            self.ep += 1
        return self._obs(), r, done, {}


_ENV_FACTORY = {"make": None}


def install_stubs():
    gym = types.ModuleType("gym")
    gym.make = lambda name: _ENV_FACTORY["make"]()
    wrappers = types.ModuleType("gym.wrappers")
    gym.wrappers = wrappers
    sys.modules["gym"] = gym
    sys.modules["gym.wrappers"] = wrappers
    sys.modules["mujoco_py"] = types.ModuleType("mujoco_py")
    sys.modules["pybullet_envs"] = types.ModuleType("pybullet_envs")
    if REF not in sys.path:
        sys.path.insert(0, REF)


class _Stop(Exception):
    pass


def _np(t):
    return t.detach().cpu().numpy().copy()


def _cat(col):
    return _np(torch.cat([c.detach() for c in col], 0))


# ----------------------------------------------------------------------------
# shims for torch 2.x
# ----------------------------------------------------------------------------
def install_shims(model_mod):
    orig_backward = torch.Tensor.backward

    def backward(self, gradient=None, retain_graph=None, create_graph=False, inputs=None, retain_variables=None):
        if retain_variables is not None:
            retain_graph = retain_variables
        hook = getattr(backward, "hook", None)
        frame = sys._getframe(1)
        orig_backward(self, gradient, retain_graph, create_graph, inputs)
        if hook is not None:
            hook(frame)

    torch.Tensor.backward = backward

    def load_state_dict(self, sd, strict=True):
        own = dict(self.named_parameters())
        for k, v in sd.items():
            own[k].data.copy_(v.data)

    model_mod.Model.load_state_dict = load_state_dict
    return backward


# ----------------------------------------------------------------------------
# DPPO: train.train x W + chief.chief, one outer iteration
# ----------------------------------------------------------------------------
def dppo_worker(rank, W, params, traffic_light, counter, shared_model, sgb, sos, test_n, turn, O, A, outdir):
    import model as model_mod
    import train as train_mod
    bw = install_shims(model_mod)
    rec = dict(idx=[], terms=[], grads=[], segs=[], raw_obs=[])
    _ENV_FACTORY["make"] = lambda: FakeEnv(O, A, 1000 + rank, [90, 37, 160, 400, 55])

    # roll out one worker after the other (Shared_obs_stats is unlocked)
    orig_observes = model_mod.Shared_obs_stats.observes
    state = dict(started=False)

    def observes(self, obs):
        if not state["started"]:
            while turn.value != rank:
                pass
            state["started"] = True
        rec["raw_obs"].append(_np(obs.data.squeeze()))
        orig_observes(self, obs)

    model_mod.Shared_obs_stats.observes = observes

    orig_push = train_mod.ReplayMemory.push

    def push(self, events):
        f = sys._getframe(1).f_locals
        L = len(f["rewards"])
        rec["segs"].append(dict(
            rewards=np.asarray(f["rewards"], np.float64),
            values=np.asarray([float(v.data[0, 0]) for v in f["values"][:L]], np.float32),
            boot=np.float32(float(f["values"][L].data[0, 0])),
            done=bool(f["done"]),
            returns=_cat(f["returns"]).reshape(-1), advantages=_cat(f["advantages"]).reshape(-1)))
        orig_push(self, [[t.detach() for t in col] for col in events])

    train_mod.ReplayMemory.push = push

    orig_sample = random.sample

    def sample(pop, k):
        if not rec["idx"]:
            cols = list(zip(*pop))
            rec["buffer"] = dict(states=_cat(cols[0]), actions=_cat(cols[1]),
                                 returns=_cat(cols[2]), advantages=_cat(cols[3]))
            rec["obs_stats"] = dict(n=_np(sos.n), mean=_np(sos.mean), mean_diff=_np(sos.mean_diff), var=_np(sos.var))
            with turn.get_lock():
                turn.value = rank + 1
        idx = orig_sample(range(len(pop)), k)
        rec["idx"].append(np.asarray(idx, np.int32))
        return [pop[i] for i in idx]

    random.sample = sample

    def after_backward(frame):
        f = frame.f_locals
        rec["terms"].append(np.asarray([float(f["loss_clip"]), float(f["loss_value"]),
                                        float(f["loss_ent"]), float(f["total_loss"])], np.float64))
        rec["grads"].append(np.concatenate([_np(p.grad).reshape(-1) for _, p in f["model"].named_parameters()]))

    bw.hook = after_backward

    orig_add = model_mod.Shared_grad_buffers.add_gradient

    def add_gradient(self, model):
        while counter.get() != rank:        # arrival order := rank order
            pass
        orig_add(self, model)

    model_mod.Shared_grad_buffers.add_gradient = add_gradient

    def clear(self):
        raise _Stop()

    train_mod.ReplayMemory.clear = clear
    random.seed(7 + rank)                    # the reference never seeds `random`
    try:
        train_mod.train(rank, params, traffic_light, counter, shared_model, sgb, sos, test_n)
    except _Stop:
        pass
    torch.save(rec, os.path.join(outdir, "worker%d.pt" % rank))


def dppo_chief(params, traffic_light, counter, shared_model, sgb, optimizer, outdir):
    import time
    import chief as chief_mod
    real_sleep = time.sleep
    chief_mod.time.sleep = lambda s: real_sleep(0.002)
    steps = []
    orig_step = optimizer.step

    def step(*a, **k):
        G = np.concatenate([_np(sgb.grads[n + "_grad"]).reshape(-1) for n, _ in shared_model.named_parameters()])
        orig_step(*a, **k)
        ps = list(shared_model.parameters())
        steps.append(dict(
            G=G,
            theta=np.concatenate([_np(p).reshape(-1) for p in ps]),
            m=np.concatenate([_np(optimizer.state[p]["exp_avg"]).reshape(-1) for p in ps]),
            v=np.concatenate([_np(optimizer.state[p]["exp_avg_sq"]).reshape(-1) for p in ps]),
            step=float(optimizer.state[ps[0]]["step"])))
        torch.save(steps, os.path.join(outdir, "chief.pt"))

    optimizer.step = step
    chief_mod.chief(0, params, traffic_light, counter, shared_model, sgb, optimizer)


def run_dppo(name, O, A, W, T, num_steps, B, E):
    import tempfile
    import torch.multiprocessing as mp
    import torch.optim as optim
    import main as main_mod                       # Params of main.py:18-39 (gym import is the stub)
    import model as model_mod
    import utils as utils_mod

    torch.set_num_threads(1)
    params = main_mod.Params()
    params.batch_size = B
    params.num_epoch = E
    params.num_steps = num_steps
    params.exploration_size = T
    params.num_processes = W
    params.update_treshold = W - 1
    torch.manual_seed(params.seed)
    traffic_light = utils_mod.TrafficLight()
    counter = utils_mod.Counter()
    shared_model = model_mod.Model(O, A)
    shared_model.share_memory()
    theta0 = np.concatenate([_np(p).reshape(-1) for p in shared_model.parameters()])
    names = [n for n, _ in shared_model.named_parameters()]
    sgb = model_mod.Shared_grad_buffers(shared_model)
    sos = model_mod.Shared_obs_stats(O)
    optimizer = optim.Adam(shared_model.parameters(), lr=params.lr)
    test_n = torch.Tensor([0])
    test_n.share_memory_()
    turn = mp.Value("i", 0)
    outdir = tempfile.mkdtemp(prefix="golden_")
    ctx = mp.get_context("fork")
    pc = ctx.Process(target=dppo_chief, args=(params, traffic_light, counter, shared_model, sgb, optimizer, outdir))
    pc.start()
    ws = []
    for r in range(W):
        p = ctx.Process(target=dppo_worker, args=(r, W, params, traffic_light, counter, shared_model, sgb, sos,
                                                  test_n, turn, O, A, outdir))
        p.start()
        ws.append(p)
    for p in ws:
        p.join()
        assert p.exitcode == 0, p.exitcode
    pc.terminate()
    pc.join()

    out = dict(O=O, A=A, W=W, T=T, B=B, E=E, lr=params.lr, gamma=params.gamma, gae_param=params.gae_param,
               clip=params.clip, ent_coeff=params.ent_coeff, theta0=theta0, names=np.asarray(names))
    chief = torch.load(os.path.join(outdir, "chief.pt"), weights_only=False)
    assert len(chief) == E, len(chief)
    for k, s in enumerate(chief):
        out["G_%d" % k] = s["G"]
        out["theta_%d" % k] = s["theta"]
    out["m_final"] = chief[-1]["m"]
    out["v_final"] = chief[-1]["v"]
    out["step_final"] = chief[-1]["step"]
    for r in range(W):
        rec = torch.load(os.path.join(outdir, "worker%d.pt" % r), weights_only=False)
        pre = "w%d_" % r
        for kk, vv in rec["buffer"].items():
            out[pre + kk] = vv
        for kk, vv in rec["obs_stats"].items():
            out[pre + "stats_" + kk] = vv
        out[pre + "raw_obs"] = np.stack(rec["raw_obs"])
        out[pre + "idx"] = np.stack(rec["idx"])
        out[pre + "terms"] = np.stack(rec["terms"])
        for k, g in enumerate(rec["grads"]):
            out[pre + "grad_%d" % k] = g
        # GAE inputs, all segments concatenated
        segs = rec["segs"]
        out[pre + "gae_rewards"] = np.concatenate([s["rewards"] for s in segs]).astype(np.float32)
        out[pre + "gae_values"] = np.concatenate([s["values"] for s in segs])
        se = np.zeros(len(out[pre + "gae_rewards"]), np.uint8)
        bo = np.zeros(len(se), np.float32)
        o = 0
        for s in segs:
            o += len(s["rewards"])
            se[o - 1] = 1
            bo[o - 1] = s["boot"]
        out[pre + "gae_seg_end"] = se
        out[pre + "gae_boot"] = bo
        out[pre + "gae_done"] = np.asarray([s["done"] for s in segs])
        out[pre + "gae_returns"] = np.concatenate([s["returns"] for s in segs])
        out[pre + "gae_advantages"] = np.concatenate([s["advantages"] for s in segs])
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


# ----------------------------------------------------------------------------
# ppo.py: one horizon
# ----------------------------------------------------------------------------
def run_ppo(name, O, A, T, B, E, keep_epochs):
    import torch.nn as nn
    import torch.optim as optim
    import model as model_mod
    import ppo as ppo_mod

    torch.set_num_threads(1)
    params = ppo_mod.Params()
    params.batch_size = B
    params.num_epoch = E
    params.num_steps = T
    params.time_horizon = 1
    ppo_mod.params = params                       # ppo.train reads a module global (ppo.py:63)
    torch.manual_seed(params.seed)
    env = FakeEnv(O, A, 4242, [150, 41, 300, 77])
    model = model_mod.Model(O, A)
    theta0 = np.concatenate([_np(p).reshape(-1) for p in model.parameters()])
    names = [n for n, _ in model.named_parameters()]
    sos = model_mod.Shared_obs_stats(O)
    optimizer = optim.Adam(model.parameters(), lr=params.lr)
    rec = dict(idx=[], terms=[], g_raw=[], g_clip=[], norm=[], theta=[], segs=[], raw_obs=[])

    orig_observes = model_mod.Shared_obs_stats.observes

    def observes(self, obs):
        rec["raw_obs"].append(_np(obs.data.squeeze()))
        orig_observes(self, obs)

    model_mod.Shared_obs_stats.observes = observes

    orig_push = ppo_mod.ReplayMemory.push

    def push(self, events):
        f = sys._getframe(1).f_locals
        L = len(f["rewards"])
        rec["segs"].append(dict(
            rewards=np.asarray(f["rewards"], np.float64),
            values=np.asarray([float(v.data[0, 0]) for v in f["values"][:L]], np.float32),
            boot=np.float32(float(f["values"][L].data[0, 0])), done=bool(f["done"]),
            returns=_cat(f["returns"]).reshape(-1), advantages=_cat(f["advantages"]).reshape(-1)))
        orig_push(self, events)

    ppo_mod.ReplayMemory.push = push

    orig_sample = random.sample

    def sample(pop, k):
        if not rec["idx"]:
            cols = list(zip(*pop))
            rec["buffer"] = dict(states=_cat(cols[0]), actions=_cat(cols[1]), returns=_cat(cols[2]),
                                 advantages=_cat(cols[3]), logprobs=_cat(cols[4]))
            rec["obs_stats"] = dict(n=_np(sos.n), mean=_np(sos.mean), mean_diff=_np(sos.mean_diff), var=_np(sos.var))
        idx = orig_sample(range(len(pop)), k)
        rec["idx"].append(np.asarray(idx, np.int32))
        return [pop[i] for i in idx]

    random.sample = sample

    orig_clip = nn.utils.clip_grad_norm

    def clip_grad_norm(parameters, max_norm, *a, **k):
        ps = list(parameters)
        rec["g_raw"].append(np.concatenate([_np(p.grad).reshape(-1) for p in ps]))
        n = orig_clip(ps, max_norm, *a, **k)
        rec["norm"].append(float(n))
        rec["g_clip"].append(np.concatenate([_np(p.grad).reshape(-1) for p in ps]))
        return n

    nn.utils.clip_grad_norm = clip_grad_norm

    orig_step = optimizer.step

    def step(*a, **k):
        f = sys._getframe(1).f_locals
        rec["terms"].append(np.asarray([float(f["loss_clip"]), float(f["loss_value"]), float(f["loss_ent"]),
                                        float(f["total_loss"])], np.float64))
        orig_step(*a, **k)
        rec["theta"].append(np.concatenate([_np(p).reshape(-1) for p in model.parameters()]))

    optimizer.step = step
    random.seed(11)
    ppo_mod.train(env, model, optimizer, sos)
    random.sample = orig_sample
    nn.utils.clip_grad_norm = orig_clip

    ps = list(model.parameters())
    out = dict(O=O, A=A, T=T, B=B, E=E, lr=params.lr, gamma=params.gamma, gae_param=params.gae_param,
               clip=params.clip, ent_coeff=params.ent_coeff, max_grad_norm=params.max_grad_norm,
               theta0=theta0, names=np.asarray(names))
    for kk, vv in rec["buffer"].items():
        out[kk] = vv
    for kk, vv in rec["obs_stats"].items():
        out["stats_" + kk] = vv
    out["raw_obs"] = np.stack(rec["raw_obs"])
    out["idx"] = np.stack(rec["idx"])
    out["terms"] = np.stack(rec["terms"])
    out["norm"] = np.asarray(rec["norm"])
    for k in keep_epochs:
        out["g_raw_%d" % k] = rec["g_raw"][k]
        out["g_clip_%d" % k] = rec["g_clip"][k]
        out["theta_%d" % k] = rec["theta"][k]
    out["m_final"] = np.concatenate([_np(optimizer.state[p]["exp_avg"]).reshape(-1) for p in ps])
    out["v_final"] = np.concatenate([_np(optimizer.state[p]["exp_avg_sq"]).reshape(-1) for p in ps])
    segs = rec["segs"]
    out["gae_rewards"] = np.concatenate([s["rewards"] for s in segs]).astype(np.float32)
    out["gae_values"] = np.concatenate([s["values"] for s in segs])
    se = np.zeros(len(out["gae_rewards"]), np.uint8)
    bo = np.zeros(len(se), np.float32)
    o = 0
    for s in segs:
        o += len(s["rewards"])
        se[o - 1] = 1
        bo[o - 1] = s["boot"]
    out["gae_seg_end"] = se
    out["gae_boot"] = bo
    out["gae_returns"] = np.concatenate([s["returns"] for s in segs])
    out["gae_advantages"] = np.concatenate([s["advantages"] for s in segs])
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    import warnings
    warnings.filterwarnings("ignore")
    install_stubs()
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "dppo"):
        # W=2 workers, HalfCheetah shape, reference arch (model.py:11-22)
        run_dppo("dppo_w2_o17a6", O=17, A=6, W=2, T=512, num_steps=200, B=128, E=3)
    if which in ("all", "dppo1"):
        # W=1, B == T (main.py:20,28: batch_size == exploration_size), small shape
        run_dppo("dppo_w1_o3a2", O=3, A=2, W=1, T=256, num_steps=100, B=256, E=2)
    if which in ("all", "ppo"):
        import model as model_mod
        install_shims(model_mod)
        run_ppo("ppo_o4a2", O=4, A=2, T=512, B=64, E=10, keep_epochs=(0, 1, 9))
