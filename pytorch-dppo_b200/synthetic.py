"""Synthetic rollouts of the named MuJoCo/PyBullet shapes (gym is not available;
only shapes matter to the update path).  SURVEY.md section 8(d): raw observations
~ loc_f + scale_f*N(0,1) normalised by K4 into states clamped to [-5,5],
actions = mu_theta0(state) + std*N(0,1) with the flavour's convention
(train.py:89 / ppo.py:93), rewards ~ clamp(N(0,1),-1,1) (train.py:96), values
= V_theta0(state), segments of ``seg_len`` steps plus a ragged tail, each ending
with done (p = 1/2, bootstrap 0) or a bootstrap value (train.py:109-112)."""
import torch

import b200_native as nv
from b200_engine import RolloutBuffer
from synthetic_data import host_arrays

SHAPES = {                       # env name -> (obs dim, act dim)
    "InvertedPendulum": (4, 1), "HalfCheetah": (17, 6), "HopperBullet": (15, 3), "Humanoid": (376, 17),
}


def make_rollout(model, obs_stats, T, seed, flavour, seg_len=1000, pinned_host=True):
    """Returns (host RolloutBuffer in pinned memory, device RolloutBuffer), both filled
    up to (and excluding) adv/ret, which K1 produces."""
    O, A = model.dims.O, model.dims.A
    dev = model.flat.device
    host = RolloutBuffer(T, O, A, dev, pinned_host=pinned_host)
    raw_obs, rew, eps, boot_noise, coin = host_arrays(O, A, T, seed)
    host.raw_obs.copy_(raw_obs)
    host.rew.copy_(rew)
    d = RolloutBuffer(T, O, A, dev)
    d.raw_obs.copy_(host.raw_obs)
    if obs_stats is not None:
        # normalise with a scratch copy of the statistics so that the caller's stay untouched
        import copy
        st = copy.copy(obs_stats)
        st.n, st.mean, st.mean_diff, st.var = (t.clone() for t in (obs_stats.n, obs_stats.mean, obs_stats.mean_diff, obs_stats.var))
        st.observes_batch(d.raw_obs, out=d.obs)
    else:
        d.obs.copy_(d.raw_obs.clamp(-5, 5))
    # actions, values and stored log-probs from the batched rollout forward of the update path (Model.act, no torch
    # eager / cuBLAS): action = mu + std*eps with the flavour's convention (train.py:89 / ppo.py:93)
    model.act(d.obs, eps.to(dev), flavour == nv.FLAVOUR_DPPO, v=d.val, act=d.act, logp=d.logp)
    seg_end = torch.zeros(T, dtype=torch.uint8)
    seg_end[seg_len - 1::seg_len] = 1
    seg_end[T - 1] = 1
    ends = seg_end.bool()
    boot = torch.where(ends & (coin >= 0.5), boot_noise * 0.5, torch.zeros(T))
    host.seg_end.copy_(seg_end)
    host.boot.copy_(boot)
    for f in ("obs", "act", "val", "logp"):
        getattr(host, f).copy_(getattr(d, f))
    for f in ("rew", "boot", "seg_end"):
        getattr(d, f).copy_(getattr(host, f))
    torch.cuda.synchronize()
    return host, d
