"""DPPO worker (reference train.py) on the B200 update path.

Kept from the reference: ``ReplayMemory(capacity)`` with push/sample/clear
(train.py:17-33), ``normal(x, mu, std)`` (train.py:41-44) and
``train(rank, params, traffic_light, counter, shared_model, shared_grad_buffers,
shared_obs_stats, test_n)`` (train.py:46).  What changed underneath:
  * the rollout buffer is struct-of-arrays in HBM and ``sample`` draws int32
    positions with the same ``random.sample`` stream (K2 gathers them);
  * GAE is one K1 launch over the whole buffer instead of the Python loop
    (train.py:113-122);
  * the epoch loop (train.py:133-175) is E x (K2 -> K3): the gradient sum into
    the shared buffers and the chief's Adam step are part of K3, so with the
    default ``params.fused_chief = True`` the worker does not wait on the traffic
    light (the chief's work is done in-stream, on every rank identically).
Rollout collection needs an environment (``params.env_fn`` or gym); it steps the
env on the host one action at a time as the reference does and is outside the
accelerated path."""
import random

import numpy as np
import torch

import b200_native as nv
from b200_engine import RolloutBuffer, UpdateEngine
from model import Model


class ReplayMemory(object):
    """FIFO of the last ``capacity`` steps (train.py:17-33), stored as columns."""

    def __init__(self, capacity, device=None):
        self.capacity = capacity
        self.device = device
        self.columns = None                      # list of (n, width) tensors
        self.memory = _LenProxy(self)            # ``len(memory.memory)`` is used by ppo.py:74

    def __len__(self):
        return 0 if self.columns is None else self.columns[0].shape[0]

    def push(self, events):
        """``events``: one list (or (n,w) tensor) per column, as at train.py:124."""
        cols = [c if torch.is_tensor(c) else torch.cat([getattr(e, "data", e).reshape(1, -1) for e in c], 0)
                for c in events]
        cols = [c.detach().reshape(c.shape[0], -1).to(self.device or c.device, torch.float32) for c in cols]
        if self.columns is None:
            self.columns = cols
        else:
            self.columns = [torch.cat([a, b], 0) for a, b in zip(self.columns, cols)]
        if len(self) > self.capacity:            # evict the oldest (train.py:25-26)
            self.columns = [c[-self.capacity:].contiguous() for c in self.columns]

    def clear(self):
        self.columns = None

    def sample_indices(self, batch_size):
        return random.sample(range(len(self)), batch_size)

    def sample(self, batch_size):
        idx = torch.tensor(self.sample_indices(batch_size), dtype=torch.long, device=self.columns[0].device)
        return map(lambda c: c.index_select(0, idx), self.columns)


class _LenProxy:
    def __init__(self, owner):
        self._owner = owner

    def __len__(self):
        return len(self._owner)


def normal(x, mu, std):
    """Per-dimension Gaussian density with ``std`` used as the variance (train.py:41-44)."""
    density = torch.exp(-(x - mu) ** 2 / (2 * std))
    return density / torch.sqrt(2 * std * np.pi)


def _make_env(params):
    if getattr(params, "env_fn", None) is not None:
        return params.env_fn()
    try:
        import gym
    except ImportError as e:
        raise RuntimeError("train(): no environment - set params.env_fn (gym is not installed)") from e
    return gym.make(params.env_name)


def collect_rollout(env, model, shared_obs_stats, params, state, episode_length, std_is_variance=True, out=None):
    """train.py:69-124 for ONE exploration phase: steps the env until at least
    ``exploration_size`` steps are stored, returns a RolloutBuffer holding the last
    ``exploration_size`` of them (raw states are normalised per step through
    Shared_obs_stats as at train.py:84-85).  ``out``: a RolloutBuffer of that length to
    fill (the engine's persistent one, UpdateEngine.rollout_buffer) instead of a new one."""
    dev = model.device
    cols = dict(obs=[], act=[], rew=[], val=[], boot=[], seg_end=[], logp=[])
    w = -1
    done = True
    while w < params.exploration_size:
        n_seg = 0
        for step in range(params.num_steps):
            w += 1
            n_seg += 1
            episode_length += 1
            shared_obs_stats.observes(state)
            s = shared_obs_stats.normalize(state.to(dev))
            with torch.no_grad():
                mu, sigma_sq, v = model(s)
                std = sigma_sq.sqrt() if std_is_variance else sigma_sq
                action = mu + std * torch.randn(mu.size()).to(dev)
                logp = (-0.5 * ((action - mu) / sigma_sq).pow(2) - 0.5 * np.log(2 * np.pi) - model.log_std).sum(-1)
            cols["obs"].append(s.reshape(1, -1)); cols["act"].append(action.reshape(1, -1))
            cols["val"].append(v.reshape(1)); cols["logp"].append(logp.reshape(1))
            nxt, reward, done, _ = env.step(action.reshape(-1).cpu().numpy())
            done = bool(done or episode_length >= params.max_episode_length)
            cols["rew"].append(max(min(float(reward), 1.0), -1.0))          # train.py:96
            if done:
                episode_length = 0
                nxt = env.reset()
            state = torch.as_tensor(np.asarray(nxt), dtype=torch.float32).unsqueeze(0)
            if done:
                break
        boot = 0.0
        if not done:                                                        # train.py:110-112 (raw state)
            with torch.no_grad():
                boot = float(model(state.to(dev))[2])
        cols["seg_end"] += [0] * (n_seg - 1) + [1]
        cols["boot"] += [0.0] * (n_seg - 1) + [boot]
    T = params.exploration_size
    buf = out if out is not None else RolloutBuffer(T, model.dims.O, model.dims.A, dev)
    if buf.T != T:
        raise nv.DppoError("collect_rollout: out holds %d steps, the rollout has %d" % (buf.T, T))
    buf.obs.copy_(torch.cat(cols["obs"], 0)[-T:]); buf.act.copy_(torch.cat(cols["act"], 0)[-T:])
    buf.val.copy_(torch.cat(cols["val"], 0)[-T:]); buf.logp.copy_(torch.cat(cols["logp"], 0)[-T:])
    buf.rew.copy_(torch.tensor(cols["rew"][-T:])); buf.boot.copy_(torch.tensor(cols["boot"][-T:]))
    buf.seg_end.copy_(torch.tensor(cols["seg_end"][-T:], dtype=torch.uint8))
    return buf, state, episode_length


def collect_rollout_vec(envs, model, shared_obs_stats, params, states, episode_lengths, std_is_variance=True, out=None,
                        eps_fn=None):
    """The rollout phase of train.py:69-124 for N environments stepped together.  Per step ONE
    K4 launch feeds the N raw observations to the shared statistics in environment order and normalises them
    (model.py:68-80), and ONE launch of the update path's forward kernel (Model.act) computes mu, v, the action
    ``mu + std*eps`` and its log-probability for all N and writes them straight into the SoA rollout buffer - nothing of
    the per-step arithmetic runs through torch eager.  Only the N actions go back to the host for ``env.step``.

    Environment e owns rows [e*L, (e+1)*L) of the buffer (L = exploration_size / N steps each): its trajectory is
    contiguous, a segment ends where the episode ends (bootstrap 0) or at row (e+1)*L - 1 (bootstrap V of the raw next
    state, train.py:109-112).  ``states``: list of the N current raw observations; returns (buffer, states, lengths)."""
    dev = model.device
    N, O, A = len(envs), model.dims.O, model.dims.A
    T = params.exploration_size
    if T % N:
        raise nv.DppoError("collect_rollout_vec: exploration_size (%d) must be a multiple of the number of environments (%d)" % (T, N))
    L = T // N
    buf = out if out is not None else RolloutBuffer(T, O, A, dev)
    if buf.T != T:
        raise nv.DppoError("collect_rollout_vec: out holds %d steps, the rollout has %d" % (buf.T, T))
    raw_host = torch.empty(N, O, pin_memory=True)
    raw_dev, norm_dev = torch.empty(N, O, device=dev), torch.empty(N, O, device=dev)
    act_dev = torch.empty(N, A, device=dev)
    act_host = torch.empty(N, A, pin_memory=True)
    rew = torch.zeros(T)
    seg_end = torch.zeros(T, dtype=torch.uint8)
    boot = torch.zeros(T)
    for t in range(L):
        for e in range(N):
            raw_host[e] = torch.as_tensor(np.asarray(states[e]), dtype=torch.float32).reshape(-1)
        raw_dev.copy_(raw_host, non_blocking=True)
        shared_obs_stats.observes_batch(raw_dev, out=norm_dev)                      # train.py:84-85 for N rows
        eps = eps_fn(t) if eps_fn is not None else torch.randn(N, A, device=dev)
        model.act(norm_dev, eps, std_is_variance, v=buf.val, act=buf.act, logp=buf.logp, states_out=buf.obs,
                  row0=t, row_stride=L)                                             # train.py:87-90, rows e*L + t
        act_dev.copy_(buf.act.view(N, L, A)[:, t])
        act_host.copy_(act_dev)                                                     # synchronises: the host needs the actions
        for e in range(N):
            nxt, reward, done, _ = envs[e].step(act_host[e].numpy())
            episode_lengths[e] += 1
            done = bool(done or episode_lengths[e] >= params.max_episode_length)
            row = e * L + t
            rew[row] = max(min(float(reward), 1.0), -1.0)                           # train.py:96
            if done:
                seg_end[row] = 1
                episode_lengths[e] = 0
                nxt = envs[e].reset()
            states[e] = nxt
    # bootstrap of the open segments: V of the RAW next state (train.py:110-112), one more batched forward
    for e in range(N):
        raw_host[e] = torch.as_tensor(np.asarray(states[e]), dtype=torch.float32).reshape(-1)
    raw_dev.copy_(raw_host)
    v_next = torch.empty(N, device=dev)
    model.act(raw_dev, None, std_is_variance, v=v_next)
    v_next = v_next.cpu()
    for e in range(N):
        row = (e + 1) * L - 1
        if not seg_end[row]:
            seg_end[row] = 1
            boot[row] = v_next[e]
    buf.rew.copy_(rew); buf.seg_end.copy_(seg_end); buf.boot.copy_(boot)
    return buf, states, episode_lengths


def make_engine(params, model, comm=None):
    return UpdateEngine(model, nv.FLAVOUR_DPPO, params.batch_size, params.num_epoch, params.lr, clip=params.clip,
                        ent_coeff=params.ent_coeff, gamma=params.gamma, gae_param=params.gae_param, comm=comm)


def train(rank, params, traffic_light, counter, shared_model, shared_grad_buffers, shared_obs_stats, test_n):
    torch.manual_seed(params.seed)                                           # train.py:47 (same seed on every rank)
    env = _make_env(params)
    model = shared_model                       # one replica per GPU process; K3 keeps replicas identical
    engine = getattr(params, "engine", None) or make_engine(params, model, getattr(params, "comm", None))
    num_envs = int(getattr(params, "num_envs", 1))
    if num_envs > 1:                           # N environments stepped together, forward + statistics batched on the GPU
        envs = [env] + [_make_env(params) for _ in range(num_envs - 1)]
        states, lengths = [e.reset() for e in envs], [0] * num_envs
    else:
        state = torch.as_tensor(np.asarray(env.reset()), dtype=torch.float32).unsqueeze(0)
    episode_length = 0
    fused = getattr(params, "fused_chief", True)
    outer = 0
    while getattr(params, "max_outer_iters", None) is None or outer < params.max_outer_iters:
        outer += 1
        if num_envs > 1:
            buf, states, lengths = collect_rollout_vec(envs, model, shared_obs_stats, params, states, lengths,
                                                       out=engine.rollout_buffer(params.exploration_size))
        else:
            buf, state, episode_length = collect_rollout(env, model, shared_obs_stats, params, state, episode_length,
                                                         out=engine.rollout_buffer(params.exploration_size))
        engine.gae(buf)                                                      # train.py:108-122
        engine.draw_indices(buf.T)                                           # train.py:140, all epochs
        if fused:
            engine.run_epochs(buf, gshared=shared_grad_buffers.flat,         # train.py:133-175 + chief.py:13-20
                              use_graph=getattr(params, "use_graph", True))
            for _ in range(params.num_epoch):
                counter.increment()
                counter.reset()
                traffic_light.switch()
        else:
            _epochs_with_external_chief(engine, buf, params, traffic_light, counter, shared_grad_buffers)
        test_n += 1                                                          # train.py:177


def _epochs_with_external_chief(engine, buf, params, traffic_light, counter, shared_grad_buffers):
    """Reference hand-shake (train.py:133-175): gradient into the shared buffers, count,
    spin until the chief (chief.chief in another thread) has stepped."""
    engine.idx_dev.copy_(engine.idx_host)
    engine.theta_old.copy_(engine.theta)
    for k in range(params.num_epoch):
        signal_init = traffic_light.get()
        snap = engine.theta.clone()
        nv.update_fused(engine.flavour, engine.dims, buf.obs, buf.act, buf.adv, buf.ret, None, engine.idx_dev[k],
                        engine.theta, engine.theta_old, engine.clip, engine.ent_coeff, engine.partials,
                        engine.loss_partials)
        engine.grad.copy_(engine.partials.sum(0))
        engine.theta_old.copy_(snap)                                         # train.py:164
        shared_grad_buffers.add_gradient(engine.model)
        torch.cuda.synchronize()
        counter.increment()
        while traffic_light.get() == signal_init:
            pass
