"""Host-side (CPU, torch only) part of the synthetic rollouts: the raw observation, reward, noise and coin
arrays of ``synthetic.make_rollout`` for a given seed.  Kept apart from synthetic.py so that the CPU baseline
(bench.py --impl reference) can feed the reference the same bytes without loading the CUDA library."""
import torch


def host_arrays(O, A, T, seed):
    """raw_obs (T,O) ~ loc_f + scale_f*N(0,1), rew (T) ~ clamp(N(0,1),-1,1) (train.py:96), eps (T,A), boot_noise (T),
    coin (T) - drawn in this order from one generator, as make_rollout consumes them."""
    g = torch.Generator().manual_seed(seed)
    loc = torch.rand(O, generator=g) * 2 - 1
    scale = torch.rand(O, generator=g) * 1.5 + 0.25
    raw_obs = loc + scale * torch.randn(T, O, generator=g)
    rew = torch.randn(T, generator=g).clamp_(-1, 1)
    eps = torch.randn(T, A, generator=g)
    boot_noise = torch.randn(T, generator=g)
    coin = torch.rand(T, generator=g)
    return raw_obs, rew, eps, boot_noise, coin
