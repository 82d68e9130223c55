"""N2: the rollout's batched forward + action sampling on the update path's kernel
(dppo_policy_forward / Model.act) and the vectorised collector that writes straight into the SoA rollout buffer."""
import math

import numpy as np
import pytest
import torch

from oracle import ppo_oracle as orc
from util import assert_close_scaled

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("O,A,widths,N", [(17, 6, (100, 100, 500, 100), 500), (17, 6, (100, 100, 500, 100), 1),
                                          (4, 1, (64, 64, 64, 64), 37), (376, 17, (256, 256, 256, 256), 200),
                                          (15, 3, (100, 100, 500, 100), 4096)])
@pytest.mark.parametrize("std_is_variance", [True, False])
def test_policy_forward_vs_oracle(O, A, widths, N, std_is_variance):
    """mu, sigma_sq, v = model(state); action = mu + std*eps; log-prob (model.py:35-45, train.py:87-90, ppo.py:91-98)."""
    from model import Model
    torch.manual_seed(3)
    m = Model(O, A, *widths, device=DEV)
    with torch.no_grad():
        m.log_std.copy_(torch.randn(1, A) * 0.3)
        for n_, p_ in m.named_parameters():
            if n_.endswith("bias"):
                p_.copy_(torch.randn_like(p_) * 0.1)
    p = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(N)
    x = torch.randn(N, O, generator=g).clamp_(-5, 5)
    eps = torch.randn(N, A, generator=g)
    mu_ref, sig, v_ref = orc.forward(p, x)
    std = sig.sqrt() if std_is_variance else sig
    act_ref = mu_ref + std * eps
    logp_ref = (-0.5 * ((act_ref - mu_ref) / sig).pow(2) - 0.5 * math.log(2 * math.pi) - p["log_std"]).sum(-1)
    mu, v, act, logp = m.act(x.to(DEV), eps.to(DEV), std_is_variance)
    torch.cuda.synchronize()
    assert_close_scaled(mu.cpu(), mu_ref, 1e-5, "mu")
    assert_close_scaled(v.cpu(), v_ref.reshape(-1), 1e-5, "v")
    assert_close_scaled(act.cpu(), act_ref, 1e-5, "action")
    assert_close_scaled(logp.cpu(), logp_ref, 1e-5, "log-prob")
    # rows scattered into a larger buffer: row r -> 3 + 2 r
    big = torch.full((3 + 2 * N, A), float("nan"), device=DEV)
    vbig = torch.full((3 + 2 * N,), float("nan"), device=DEV)
    m.act(x.to(DEV), eps.to(DEV), std_is_variance, act=big, v=vbig, row0=3, row_stride=2)
    torch.cuda.synchronize()
    assert torch.equal(big[3::2], act) and torch.equal(vbig[3::2], v)
    assert torch.isnan(big[4::2]).all() and torch.isnan(big[:3]).all()


class _Space:
    def __init__(self, n):
        self.shape = (n,)


class ScriptedEnv:
    """Observations and rewards from a fixed table (independent of the actions), episodes of a fixed length."""

    def __init__(self, O, A, seed, ep_len):
        self.observation_space, self.action_space = _Space(O), _Space(A)
        r = np.random.RandomState(seed)
        self.obs = (r.randn(4000, O) * 2 + 1).astype(np.float32)
        self.rew = (r.randn(4000) * 0.8).astype(np.float32)
        self.ep_len, self.t, self.k = ep_len, 0, 0

    def reset(self):
        self.t = 0
        self.k += 1
        return self.obs[self.k % 4000]

    def step(self, action):
        assert np.all(np.isfinite(action))
        self.t += 1
        self.k += 1
        return self.obs[self.k % 4000], float(self.rew[self.k % 4000]), self.t >= self.ep_len, {}


def test_collect_rollout_vec_matches_the_sequential_rollout_loop():
    """Four environments stepped together for 48 steps each against the rollout loop of train.py:82-106 restated with
    the oracle's functions (statistics fed in environment order within a step)."""
    import main as main_mod
    from model import Model, Shared_obs_stats
    from train import collect_rollout_vec
    O, A, N, L = 7, 3, 4, 48
    torch.manual_seed(2)
    m = Model(O, A, device=DEV)
    stats = Shared_obs_stats(O)
    p = main_mod.Params()
    p.exploration_size, p.max_episode_length = N * L, 10000
    g = torch.Generator().manual_seed(9)
    eps_all = torch.randn(L, N, A, generator=g)
    envs = [ScriptedEnv(O, A, 100 + e, 13 + 5 * e) for e in range(N)]
    states = [e.reset() for e in envs]
    buf, states_after, lengths = collect_rollout_vec(envs, m, stats, p, states, [0] * N,
                                                     eps_fn=lambda t: eps_all[t].to(DEV))
    torch.cuda.synchronize()
    # ---- the same loop, sequentially, with the oracle ----
    pr = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    envs = [ScriptedEnv(O, A, 100 + e, 13 + 5 * e) for e in range(N)]
    cur = [e.reset() for e in envs]
    st = orc.welford_new(O)
    T = N * L
    obs = torch.zeros(T, O); act = torch.zeros(T, A); val = torch.zeros(T); rew = torch.zeros(T)
    seg = torch.zeros(T, dtype=torch.uint8); boot = torch.zeros(T)
    ep = [0] * N
    for t in range(L):
        raw = torch.stack([torch.from_numpy(np.asarray(c)) for c in cur])
        norm = orc.welford_observe_normalize(st, raw)                           # model.py:68-80, rows in environment order
        mu, sig, v = orc.forward(pr, norm)
        a = mu + sig.sqrt() * eps_all[t]                                        # train.py:89
        for e in range(N):
            row = e * L + t
            obs[row], act[row], val[row] = norm[e], a[e], v[e, 0]
            nxt, r, done, _ = envs[e].step(a[e].numpy())
            ep[e] += 1
            rew[row] = max(min(float(r), 1.0), -1.0)
            if done:
                seg[row] = 1
                ep[e] = 0
                nxt = envs[e].reset()
            cur[e] = nxt
    raw = torch.stack([torch.from_numpy(np.asarray(c)) for c in cur])
    _, _, vn = orc.forward(pr, raw)                                              # bootstrap on the raw state, train.py:110-112
    for e in range(N):
        row = (e + 1) * L - 1
        if not seg[row]:
            seg[row] = 1
            boot[row] = vn[e, 0]
    assert_close_scaled(buf.obs.cpu(), obs, 1e-5, "states")
    assert_close_scaled(buf.act.cpu(), act, 1e-5, "actions")
    assert_close_scaled(buf.val.cpu(), val, 1e-5, "values")
    assert torch.equal(buf.seg_end.cpu(), seg)
    assert_close_scaled(buf.rew.cpu(), rew, 1e-7, "rewards")
    assert_close_scaled(buf.boot.cpu(), boot, 1e-5, "bootstrap values")
    assert_close_scaled(stats.mean.cpu(), st["mean"], 1e-5, "statistics mean")
    assert float(stats.n[0]) == T


def test_train_with_vectorised_environments():
    """train.train with params.num_envs = 4: rollouts through collect_rollout_vec, two outer iterations."""
    import main as main_mod
    from model import Model, Shared_grad_buffers, Shared_obs_stats
    from train import train
    from utils import Counter, TrafficLight
    O, A = 5, 2
    torch.manual_seed(1)
    model = Model(O, A, device=DEV)
    before = model.flat.clone()
    p = main_mod.Params()
    p.batch_size, p.num_epoch, p.num_steps, p.exploration_size, p.num_processes = 64, 3, 50, 128, 1
    p.num_envs, p.max_outer_iters = 4, 2
    seeds = iter(range(1000))
    p.env_fn = lambda: ScriptedEnv(O, A, 50 + next(seeds), 21)
    stats = Shared_obs_stats(O)
    test_n = torch.zeros(1)
    train(0, p, TrafficLight(), Counter(), model, Shared_grad_buffers(model), stats, test_n)
    torch.cuda.synchronize()
    assert float(test_n) == 2.0 and float(stats.n[0]) == 2 * 128
    assert torch.isfinite(model.flat).all() and not torch.equal(model.flat, before)
