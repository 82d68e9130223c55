"""Single-process PPO (reference ppo.py) on the B200 update path: ``Params``
(ppo.py:21-42), ``ReplayMemory`` and ``train(env, model, optimizer,
shared_obs_stats)`` (ppo.py:62-178).  The epoch loop ppo.py:139-174 - stored
log-probs, clipped surrogate, plain value loss, entropy bonus, global-norm clip
0.5, Adam - is E x (K2 flavour PPO -> K3 with max_grad_norm)."""
import numpy as np
import torch

import b200_native as nv
from b200_engine import UpdateEngine
from train import ReplayMemory, collect_rollout  # noqa: F401  (ReplayMemory is part of the reference surface)


class Params():
    def __init__(self):
        self.batch_size = 64
        self.lr = 7e-4
        self.gamma = 0.99
        self.gae_param = 0.95
        self.clip = 0.2
        self.ent_coeff = 0.01
        self.num_epoch = 10
        self.num_steps = 2048
        self.time_horizon = 1000000
        self.max_episode_length = 10000
        self.max_grad_norm = 0.5
        self.seed = 1
        self.env_name = 'HalfCheetah-v1'


params = None          # the reference's train() reads a module-level ``params`` (ppo.py:63)


def make_engine(p, model, optimizer=None):
    lr, betas, eps = p.lr, (0.9, 0.999), 1e-8
    if optimizer is not None:
        g = optimizer.param_groups[0]
        lr, betas, eps = g["lr"], tuple(g["betas"]), g["eps"]
    return UpdateEngine(model, nv.FLAVOUR_PPO, p.batch_size, p.num_epoch, lr, clip=p.clip, ent_coeff=p.ent_coeff,
                        max_grad_norm=p.max_grad_norm, gamma=p.gamma, gae_param=p.gae_param, betas=betas, eps=eps)


def train(env, model, optimizer, shared_obs_stats):
    p = params if params is not None else Params()
    engine = make_engine(p, model, optimizer)
    state = torch.as_tensor(np.asarray(env.reset()), dtype=torch.float32).unsqueeze(0)
    episode_length = 0

    class _Roll:                                   # collect_rollout speaks train.py's names
        exploration_size = p.num_steps
        num_steps = p.num_steps
        max_episode_length = p.max_episode_length

    for t in range(p.time_horizon):
        buf, state, episode_length = collect_rollout(env, model, shared_obs_stats, _Roll, state, episode_length,
                                                     std_is_variance=False,           # ppo.py:93
                                                     out=engine.rollout_buffer(p.num_steps))
        engine.gae(buf)
        engine.draw_indices(buf.T)
        losses = engine.run_epochs(buf, use_graph=getattr(p, "use_graph", True))
        print('horizon', t, 'loss', losses[-1, :3].tolist())
    return engine
